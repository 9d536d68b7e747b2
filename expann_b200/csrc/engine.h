// Internal handle behind the C ABI (include/expann_gpu.h). Not installed; nothing here crosses the boundary.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <string>
#include <vector>
#include "../../include/expann_gpu.h"

namespace expann {

struct DevBuf {
	void* p = nullptr;
	size_t cap = 0;
	bool host = false; // pinned host memory instead of device memory
	cudaError_t ensure(size_t bytes) {
		if (bytes <= cap) return cudaSuccess;
		release();
		size_t want = bytes + bytes / 4 + 256;
		cudaError_t e = host ? cudaMallocHost(&p, want) : cudaMalloc(&p, want);
		if (e != cudaSuccess) { p = nullptr; cap = 0; return e; }
		cap = want;
		return cudaSuccess;
	}
	void release() {
		if (p) { if (host) cudaFreeHost(p); else cudaFree(p); }
		p = nullptr; cap = 0;
	}
	template <class T> T* as() const { return reinterpret_cast<T*>(p); }
};

// host copy of the graph in the device layout (fixed-stride rows padded with 0xFFFFFFFF)
struct HostGraph {
	bool present = false;
	size_t n = 0, m0s = 0, ms = 0;
	uint32_t start = 0, max_layer = 0;
	std::vector<uint32_t> bottom, upper, upper_idx, upper_nl;
	std::vector<float> bottom_len, upper_len; // edge lengths when known (GPU build), for the index writer
};

struct Multi; // multi.cu: state of a multi-device handle

struct DeviceGuard {
	int prev = -1;
	explicit DeviceGuard(int dev) { cudaGetDevice(&prev); if (prev != dev) cudaSetDevice(dev); }
	~DeviceGuard() { int cur; cudaGetDevice(&cur); if (prev >= 0 && cur != prev) cudaSetDevice(prev); }
};

} // namespace expann

struct expann_gpu {
	expann_gpu_cfg cfg{};
	int device = 0;
	expann::Multi* mg = nullptr; // non-null: this handle fans out to one child handle per GPU (multi.cu)
	std::string err;
	// base set: row-major float[n][d_pad] in HBM, d_pad = d rounded up to 32 floats (128 B rows-chunks for TMA)
	size_t n = 0, d = 0, d_pad = 0, cap_rows = 0;
	float* base = nullptr;
	std::vector<float> staging; // small host adds are batched here (d floats per row) until flushed
	size_t staged_rows = 0;
	bool built = false;
	CUtensorMap tmap_base{};
	cudaStream_t stream = nullptr;
	expann::DevBuf qpad, partial, keys, ids, dists, counts, scan_cnt; // device work buffers (grow-only)
	expann::DevBuf h_q, h_ids, h_dists, h_counts;            // pinned host staging for the host-pointer API
	// The per-call work buffers of the scans exist twice (brute-force engine): calls issued on two streams use one set each
	// and overlap on the GPU (one batch's conversion / select / re-rank beside the other's filter); a third stream takes over
	// the set used longest ago behind an event. The members above / below are the ACTIVE set, `spare` holds the other one's.
	struct Scratch { expann::DevBuf qpad, partial, scan_cnt, tc_qstats, tc_qb, tc_qn, tc_thr, tc_cnt, tc_cand, tc_gthr, tc_sel, tc_seed; } spare;
	int set_active = 0;
	bool set_used[2] = { false, false };
	cudaStream_t set_stream[2] = { nullptr, nullptr };
	cudaEvent_t set_done[2] = { nullptr, nullptr }; // recorded behind the last call that used the set
	expann_gpu_stats_t stats{};
	// tensor-core candidate filter (K3)
	int opt_tc = 1, opt_tc_min_batch = 5; // from 5 queries on K3 is level with or ahead of K1 (1M x 128: 0.12 vs 0.15 ms at 6, 0.13 vs 0.22 at 12)
	int tc_mode = 0; // 0 off, 1 bf16-exact single GEMM, 2 two-term bf16 split, 3 one bf16 term per value (real-valued data, wider slack), 4 tf32 on the fp32 rows
	int opt_tc_real = 1; // real-valued data: 1 = start in the one-term format (mode 3) and move to the split when a batch is redone, 0 = split from the start
	bool tc_flag_pending = false; // mode 3: a copy of the last batch's flag word is on its way to tc_flag_host
	cudaEvent_t tc_flag_event = nullptr;
	expann::DevBuf tc_flag_host;
	uint32_t tc_kd = 0;
	float tc_xmin = 0.f, tc_xmax = 0.f, tc_xnorm_max = 0.f;
	expann::DevBuf tc_xb, tc_xn, tc_stats, tc_qstats, tc_qb, tc_qn, tc_thr, tc_cnt, tc_cand, tc_gthr, tc_sel, tc_xx, tc_seed, tc_fbcount;
	CUtensorMap tmap_xb{}, tmap_xb128{}, tmap_xx{}, tmap_x32{}, tmap_xx32{}; // (x32 / xx32: format 4, fp32 maps over the base rows / tf32 norm terms)
	expann::DevBuf tc_xx32;
	int opt_tc_tf32 = 1; // real-valued rows: try kind::tf32 (format 4) between the one-term format and the split
	int opt_tc_v1 = 0; // force the chunked first form of K3 (tests)
	int opt_tc_pair = 1; // pair form of K3 (clusters of two CTAs, tcgen05.mma.cta_group::2)
	int opt_tc_astream = 1; // pair form with the query operand streamed through the ring: 0 off, 1 when nothing else fits, 2 always
	int opt_graph_occ = 5; // graph search: resident CTAs of 4 warps per SM (2, 4, 5 or 8); set before build()
	int opt_graph_cta_batch = -1; // graph search: calls with at most this many queries take the one-CTA-per-query latency form (-1 = two waves: twice the SM count; five waves over the neighbourhood blocks)
	int opt_graph_cta_width = 1; // latency form: vertices expanded per hop (1 = the reference's order, the default; 2 = half the dependent hops, recall >= at every beam width measured, +2 % distance evaluations)
	int opt_graph_hash = 0; // warp form of the search: visited set in a per-warp shared-memory hash (exact; falls back to the HBM bytes when full). Measured slower than the HBM bytes (5.28 vs 5.81 M QPS at ef 40): off by default
	int opt_graph_upper_compact = 1; // latency form: upper layers in compact form (set before build())
	int opt_tc_boot = 1; // pair form: seeding launch over the leading tiles of a large base
	int opt_tc_boot_tiles = 128; // tiles (of 256 rows) the seeding launch covers, at most 1/32 of the base (experiments)
	// quantised scan engine (K6 + K4)
	int quantizer = 0;       // 0 = quantizer_simple<uint8_t>, 1 = quantizer_ranged_q8
	int opt_kprime = 0;      // 0 = auto
	size_t dc_pad = 0;       // code row stride (d rounded up to 128)
	float q_scale = 1.f, q_offset = 0.f;
	bool q_range_fixed = false; // ranged q8: scale / offset were set from outside (multi-device handle: one range for all shards)
	expann::DevBuf q_codes, q_cnorm, q_mm, qq_codes, qq_norm, q_cand, q_cnb, q_thrfin, q_zero;
	CUtensorMap tmap_codes{}, tmap_codes128{};
	int opt_quant_tc = 1, opt_quant_tc_min_batch = 32; // code scan through the tensor-core filter (kind::i8) from this many queries on
	bool q_tc_ready = false;
	uint64_t q_tc_launches = 0;
	// graph engine
	expann::HostGraph graph;
	expann::DevBuf g_bottom, g_upper, g_upper_idx, g_upper_nl, g_visited, g_epoch, g_codes, g_up_meta, g_upper_c, g_up_vec;
	size_t g_nup = 0; // vertices with upper layers in the compact form (0 = not built)
	// latency form on neighbourhood blocks (graph_inline.cuh): block(v) = ids of v's neighbours + their vectors, contiguous
	expann::DevBuf g_blk_bottom, g_blk_upper;
	uint32_t g_bb_bytes = 0, g_bb_idb = 0, g_ub_bytes = 0, g_ub_idb = 0;
	bool g_inline = false; // the blocks are built
	int opt_graph_inline_batch = 1; // the one-warp-per-query form reads the blocks too when they exist (+8-12 % QPS: one region of HBM per hop instead of scattered rows)
	int opt_graph_inline = -1; // -1 automatic (built when the blocks take at most a third of the free HBM), 0 off, 1 whenever they fit, 2 = 1 with the visited set kept in HBM bytes (what graphs too large for the shared-memory bitmap get; tests); before build()
	float build_ortho_factor = 0.5f, build_ortho_bias = 0.f;
	uint64_t build_prune_overflow = 0, build_K = 0;
	uint32_t g_ctas = 0;
	// small host calls: where the latency form of the search may decode its results itself (set around one call)
	uint64_t* fuse_ids = nullptr;
	float* fuse_dists = nullptr;
	uint64_t* fuse_counts = nullptr;
	bool fused = false;
	size_t g_vis_stride = 0;
	int opt_nsplit = 0; // 0 = auto
	bool timing = false;
	std::vector<std::pair<cudaEvent_t, cudaEvent_t>> timed; // pending (start, stop) pairs of the dominant kernel
	std::vector<cudaEvent_t> event_pool;
	expann_gpu() { h_q.host = h_ids.host = h_dists.host = h_counts.host = tc_flag_host.host = true; }
};

namespace expann {
int fail(expann_gpu* h, int code, const std::string& msg);
// timing brackets around the dominant kernel (no-ops unless expann_gpu_set_timing(h, 1))
cudaEvent_t timing_begin(expann_gpu* h, cudaStream_t st);
void timing_end(expann_gpu* h, cudaEvent_t start, cudaStream_t st);
int cuda_fail(expann_gpu* h, cudaError_t e, const char* what);
// api.cu
// opt-in dynamic shared memory of a kernel: raised when a launch needs more than any launch before it on that device,
// never lowered (several engines, row lengths and host threads share the kernels of this library)
int ensure_dyn_smem(expann_gpu* h, const void* kernel, size_t bytes);
int query_device_impl(expann_gpu* h, const float* d_q, size_t m, size_t k, size_t ef, uint64_t* d_keys_out, cudaStream_t st);
bool is_pinned_host(const void* p);
// multi.cu — every entry of the C ABI on a multi-device handle
int mg_create(const expann_gpu_cfg* cfg, const int32_t* devices, size_t ndev, expann_gpu** out);
void mg_destroy(expann_gpu* h);
int mg_add(expann_gpu* h, const float* rows, size_t n, size_t d, bool device_rows, cudaStream_t st);
int mg_build(expann_gpu* h);
int mg_build_graph(expann_gpu* h, uint64_t M, uint64_t M0, uint64_t K, float ortho_factor, float ortho_bias, uint64_t prune_overflow, uint64_t seed);
int mg_load_index(expann_gpu* h, const char* path);
int mg_query(expann_gpu* h, const float* q, size_t m, size_t k, size_t ef, uint64_t* ids_out, float* dist_out, uint64_t* counts_out);
int mg_query_device(expann_gpu* h, const float* d_q, size_t m, size_t k, size_t ef, uint64_t* d_keys_out, cudaStream_t st);
int mg_set_option(expann_gpu* h, const char* name, int64_t value);
int mg_set_timing(expann_gpu* h, int on);
int mg_stats(expann_gpu* h, expann_gpu_stats_t* out);
// scan_f32.cu
int launch_scan_f32(expann_gpu* h, const float* d_q_padded, size_t m, size_t k, uint64_t* d_keys_out,
                    cudaStream_t st);
int make_base_tensor_map(expann_gpu* h);
// K1 with its scratch memory in a static device array (no allocation in the call): expann_gpu_scan_device_noalloc
void scan_noalloc_setup(expann_gpu* h);
int scan_noalloc(expann_gpu* h, const float* d_q, size_t m, size_t k, uint64_t* d_keys_out, cudaStream_t st);
// 2D tiled tensor map over fp32 (dtype_u8 = 0) or byte (1) rows; 128-byte swizzle, or 32-byte for 32-byte box rows
int encode_tensor_map_2d(expann_gpu* h, CUtensorMap* out, int dtype_u8, void* base, uint64_t inner, uint64_t rows,
                         uint64_t row_bytes, uint32_t box_inner, uint32_t box_rows, int swizzle32 = 0);
// graph.cu
int graph_set_csr(expann_gpu* h, size_t n, const uint32_t* nlayers, const uint64_t* offsets, const uint32_t* edges,
                  uint64_t start, uint64_t max_layer);
int graph_load_index(expann_gpu* h, const char* path, std::vector<float>& vectors, size_t& n_out, size_t& d_out);
int graph_finalize(expann_gpu* h);
int launch_graph_search(expann_gpu* h, const float* d_q, size_t m, size_t k, size_t ef, uint64_t* d_keys_out, cudaStream_t st);
int graph_read_distcomps(expann_gpu* h);
// tc.cu
int tc_finalize(expann_gpu* h);
int launch_scan_tc(expann_gpu* h, const float* d_q, size_t m, size_t k, uint64_t* d_keys_out, cudaStream_t st);
bool tc_first_form_fits(const expann_gpu* h); // the chunked first form of K3 keeps the query tile resident
uint32_t tc2_pick_rsplit(uint32_t qblocks, uint32_t ntiles, int sms);
// tc_i8.cu: the quantised engine's code scan on the tensor cores (tcgen05.mma.kind::i8)
int tc_code_prepare(expann_gpu* h);
bool tc_code_ready(const expann_gpu* h);
bool tc_code_fits(const expann_gpu* h, size_t kp);
int launch_code_filter(expann_gpu* h, size_t m, size_t kp, uint64_t* d_cand_out, cudaStream_t st, const uint32_t** flags_out);
// quant.cu
int quant_finalize(expann_gpu* h);
int quant_minmax(expann_gpu* h, float* mn_out, float* mx_out);
void quant_set_range(expann_gpu* h, float mn, float mx);
int launch_quant_scan(expann_gpu* h, const float* d_q, size_t m, size_t k, uint64_t* d_keys_out, cudaStream_t st);
// graph_build.cu
int graph_build_gpu(expann_gpu* h, uint64_t M, uint64_t M0, uint64_t K, float ortho_factor, float ortho_bias,
                    uint64_t prune_overflow, uint64_t seed);
int graph_save_index(expann_gpu* h, const char* path);
// merge_topk.cu
int launch_merge(expann_gpu* h, const uint64_t* in, size_t m, size_t nlists, size_t klist, size_t stride_q,
                 size_t stride_s, size_t k, uint32_t id_add, uint64_t* out, cudaStream_t st, const uint32_t* gate = nullptr);
int launch_decode(expann_gpu* h, const uint64_t* keys, size_t m, size_t k, int kind, uint64_t* ids, float* dists,
                  uint64_t* counts, cudaStream_t st);
} // namespace expann

#define EXPANN_CUDA(h, call)                                              \
	do {                                                                  \
		cudaError_t _e = (call);                                          \
		if (_e != cudaSuccess) return expann::cuda_fail((h), _e, #call);  \
	} while (0)
