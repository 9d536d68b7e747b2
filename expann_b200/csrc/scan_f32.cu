// Host side of K1/K2: tensor-map construction and launch-shape selection for the fp32 scan.
#include "engine.h"
#include "scan_f32.cuh"
#include <algorithm>
#include <mutex>

namespace expann {

constexpr int SCAN_TR = 2, SCAN_NWARP = 8, SCAN_STAGES = 3;
constexpr int SCAN_ROWS = 32 * SCAN_TR * SCAN_NWARP; // 512 rows per pipeline stage
constexpr int SCAN_BOXR = SCAN_ROWS > 256 ? 256 : SCAN_ROWS;

typedef CUresult (*encode_tiled_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static encode_tiled_fn get_encode_tiled() {
	static encode_tiled_fn fn = nullptr;
	static std::once_flag once;
	std::call_once(once, [] {
		void* p = nullptr;
		cudaDriverEntryPointQueryResult qres;
		if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
		    qres == cudaDriverEntryPointSuccess)
			fn = reinterpret_cast<encode_tiled_fn>(p);
	});
	return fn;
}

// 2D tiled tensor map with 128B swizzle (16-byte chunk c of box row r lands at chunk c ^ (r & 7) of its line)
int encode_tensor_map_2d(expann_gpu* h, CUtensorMap* out, int dtype_u8, void* base, uint64_t inner, uint64_t rows,
                         uint64_t row_bytes, uint32_t box_inner, uint32_t box_rows, int swizzle32) {
	encode_tiled_fn enc = get_encode_tiled();
	if (!enc) return fail(h, EXPANN_ERR_CUDA, "cuTensorMapEncodeTiled entry point not available");
	cuuint64_t gdim[2] = { (cuuint64_t)inner, (cuuint64_t)rows };
	cuuint64_t gstride[1] = { (cuuint64_t)row_bytes };
	cuuint32_t box[2] = { box_inner, box_rows };
	cuuint32_t estr[2] = { 1, 1 };
	CUresult r = enc(out, dtype_u8 ? CU_TENSOR_MAP_DATA_TYPE_UINT8 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, base, gdim, gstride,
	                 box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle32 ? CU_TENSOR_MAP_SWIZZLE_32B : CU_TENSOR_MAP_SWIZZLE_128B,
	                 CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
	if (r != CUDA_SUCCESS) return fail(h, EXPANN_ERR_CUDA, "cuTensorMapEncodeTiled failed: " + std::to_string((int)r));
	return EXPANN_OK;
}

int encode_tensor_map_bf16(expann_gpu* h, CUtensorMap* out, void* base, uint64_t inner, uint64_t rows, uint32_t box_rows,
                           uint32_t box_inner) {
	encode_tiled_fn enc = get_encode_tiled();
	if (!enc) return fail(h, EXPANN_ERR_CUDA, "cuTensorMapEncodeTiled entry point not available");
	cuuint64_t gdim[2] = { (cuuint64_t)inner, (cuuint64_t)rows };
	cuuint64_t gstride[1] = { (cuuint64_t)inner * 2 };
	cuuint32_t box[2] = { box_inner, box_rows };
	cuuint32_t estr[2] = { 1, 1 };
	CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, base, gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
	                 box_inner == 16 ? CU_TENSOR_MAP_SWIZZLE_32B : CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
	if (r != CUDA_SUCCESS) return fail(h, EXPANN_ERR_CUDA, "cuTensorMapEncodeTiled(bf16) failed: " + std::to_string((int)r));
	return EXPANN_OK;
}

// base set: dim0 = d_pad floats (contiguous), dim1 = n rows; box = [32 floats][BOXR rows]
int make_base_tensor_map(expann_gpu* h) {
	return encode_tensor_map_2d(h, &h->tmap_base, 0, h->base, h->d_pad, h->n, h->d_pad * 4, 32, SCAN_BOXR);
}

template <int QT, int METRIC, int LPL>
static int launch_inst(expann_gpu* h, const ScanParams& p, dim3 grid, cudaStream_t st) {
	using Cfg = ScanCfg<QT, SCAN_TR, SCAN_NWARP, SCAN_STAGES, METRIC, LPL>;
	auto kern = scan_topk_f32_kernel<QT, SCAN_TR, SCAN_NWARP, SCAN_STAGES, METRIC, LPL>;
	const size_t smem = Cfg::smem_bytes(p.d_pad);
	if (smem > 227 * 1024) return fail(h, EXPANN_ERR_UNSUPPORTED, "scan: d too large for the shared-memory query tile");
	if (int rc = ensure_dyn_smem(h, (const void*)kern, smem)) return rc;
	kern<<<grid, Cfg::THREADS, smem, st>>>(h->tmap_base, p);
	EXPANN_CUDA(h, cudaGetLastError());
	return EXPANN_OK;
}

template <int QT, int METRIC> static int launch_k(expann_gpu* h, const ScanParams& p, dim3 grid, cudaStream_t st) {
	if (p.k <= 32) return launch_inst<QT, METRIC, 1>(h, p, grid, st);
	if (p.k <= 128) return launch_inst<QT, METRIC, 4>(h, p, grid, st);
	return fail(h, EXPANN_ERR_UNSUPPORTED, "scan: k > 128 not supported");
}
template <int QT> static int launch_m(expann_gpu* h, const ScanParams& p, dim3 grid, cudaStream_t st) {
	return h->cfg.metric == EXPANN_METRIC_IP ? launch_k<QT, METRIC_IP>(h, p, grid, st)
	                                          : launch_k<QT, METRIC_L2>(h, p, grid, st);
}

int launch_scan_f32_simt(expann_gpu* h, const float* d_q, size_t m, size_t k, uint64_t* d_keys_out, cudaStream_t st,
                         const uint32_t* gate = nullptr, uint32_t* gate_count = nullptr);

// path selection: the tensor-core filter (K3) for large batches when build() prepared its operands, else the SIMT scan
int launch_scan_f32(expann_gpu* h, const float* d_q, size_t m, size_t k, uint64_t* d_keys_out, cudaStream_t st) {
	if (h->tc_mode && h->opt_tc && k <= 128) {
		// second form of K3 (k <= 16, single launch) pays from tc_min_batch queries on; the chunked first form (k <= 128)
		// has ~10 launches and host round trips per batch and only pays for large batches
		const bool v1_fits = tc_first_form_fits(h); // resident query tile: d <= 128 real-valued
		const bool second = k <= 16 && (!h->opt_tc_v1 || !v1_fits);
		if (!second && !v1_fits) return launch_scan_f32_simt(h, d_q, m, k, d_keys_out, st);
		const size_t need = second ? (size_t)h->opt_tc_min_batch : std::max<size_t>((size_t)h->opt_tc_min_batch, 512);
		if (m >= need) return launch_scan_tc(h, d_q, m, k, d_keys_out, st);
	}
	return launch_scan_f32_simt(h, d_q, m, k, d_keys_out, st);
}

int launch_scan_f32_simt(expann_gpu* h, const float* d_q, size_t m, size_t k, uint64_t* d_keys_out, cudaStream_t st,
                         const uint32_t* gate, uint32_t* gate_count) {
	int qt = h->cfg.qtile;
	if (qt == 0) qt = m >= 8 ? 8 : (m >= 3 ? 4 : 1);
	if (qt != 1 && qt != 4 && qt != 8) return fail(h, EXPANN_ERR_ARG, "qtile must be 0, 1, 4 or 8");
	// keep stages + query tile + mailboxes within 227 KB (d = 960: 8 queries do not fit, 4 do)
	auto fits = [&](int t) {
		const uint32_t dp = (uint32_t)h->d_pad;
		const size_t b = t == 8 ? ScanCfg<8, SCAN_TR, SCAN_NWARP, SCAN_STAGES, 0, 1>::smem_bytes(dp)
		               : t == 4 ? ScanCfg<4, SCAN_TR, SCAN_NWARP, SCAN_STAGES, 0, 1>::smem_bytes(dp)
		                        : ScanCfg<1, SCAN_TR, SCAN_NWARP, SCAN_STAGES, 0, 1>::smem_bytes(dp);
		return b <= 227 * 1024;
	};
	if (qt == 8 && !fits(8)) qt = 4;
	if (qt == 4 && !fits(4)) qt = 1;
	const uint32_t qtiles = (uint32_t)((m + qt - 1) / qt);
	const uint32_t groups = (uint32_t)((h->n + SCAN_ROWS - 1) / SCAN_ROWS);
	int dev_sms = 148;
	cudaDeviceGetAttribute(&dev_sms, cudaDevAttrMultiProcessorCount, h->device);
	// whole waves of CTAs (1 CTA per SM): W = ceil(qtiles / SMs) waves, as many row splits as fit in them; every split
	// keeps at least 2 row groups. A small batch (qtiles == 1) is exactly one CTA per SM, each streaming 1/SMs of the rows.
	const uint32_t waves = std::max(1u, (qtiles + dev_sms - 1) / (uint32_t)dev_sms);
	const uint32_t want = std::max(1u, (uint32_t)((uint64_t)dev_sms * waves / qtiles));
	const uint32_t max_split = std::max(1u, groups / 2u);
	uint32_t nsplit = std::max(1u, std::min(want, max_split));
	if (h->opt_nsplit > 0) nsplit = std::min((uint32_t)h->opt_nsplit, groups);
	// rows per split: even share rounded up to 64 rows (TMA box coordinates are arbitrary rows; the last row group of a
	// split is partial and masked)
	uint32_t rps = (uint32_t)(((h->n + nsplit - 1) / nsplit + 63) / 64 * 64);
	nsplit = (uint32_t)((h->n + rps - 1) / rps);

	ScanParams p{};
	p.queries = d_q;
	p.m = (uint32_t)m;
	p.k = (uint32_t)k;
	p.d_pad = (uint32_t)h->d_pad;
	p.n = (uint32_t)h->n;
	p.rows_per_split = rps;
	p.nsplit = nsplit;
	uint64_t* scan_out = d_keys_out;
	const bool need_merge = nsplit > 1 || (uint32_t)h->cfg.id_offset != 0;
	if (need_merge) {
		EXPANN_CUDA(h, h->partial.ensure((size_t)m * nsplit * k * 8));
		scan_out = h->partial.as<uint64_t>();
		if (h->scan_cnt.cap < m * 4) { // tickets of the fused merge: zero once, the kernel leaves them zero
			EXPANN_CUDA(h, h->scan_cnt.ensure(m * 4));
			EXPANN_CUDA(h, cudaMemsetAsync(h->scan_cnt.p, 0, h->scan_cnt.cap, st));
		}
		p.final_out = d_keys_out;
		p.counters = h->scan_cnt.as<uint32_t>();
		p.id_add = (uint32_t)h->cfg.id_offset;
	}
	p.partial = scan_out;
	p.gate = gate;
	p.gate_count = gate_count;
	dim3 grid(qtiles, nsplit);
	if (gate) { // K3's queued redo: does its work only if the device flag is up; not timed, counted as an auxiliary launch
		h->stats.aux_launches++;
		return qt == 8 ? launch_m<8>(h, p, grid, st) : qt == 4 ? launch_m<4>(h, p, grid, st) : launch_m<1>(h, p, grid, st);
	}
	cudaEvent_t t0 = timing_begin(h, st);
	int rc = qt == 8 ? launch_m<8>(h, p, grid, st) : qt == 4 ? launch_m<4>(h, p, grid, st) : launch_m<1>(h, p, grid, st);
	timing_end(h, t0, st);
	if (rc) return rc;
	h->stats.scan_launches++;
	h->stats.last_nsplit = nsplit;
	h->stats.last_qtile = qt;
	return EXPANN_OK; // split merge is fused into the scan (last writer of each query)
}

// ---- K1 without allocations: expann_gpu_scan_device_noalloc (include/expann_gpu.h) -----------------------------------
// The scan's scratch memory -- the per-row-split key lists and the tickets of the fused merge -- lives in ONE file-scope
// array instead of the handle's grow-only buffers, so the call makes no cudaMalloc / cudaFree (callers under CUDA graph
// capture, serving loops and harnesses that forbid allocation). Every GPU that loads the library holds the array until
// exit: 8 MiB of key lists + 64 KiB of tickets. ALL calls on a device share it: calls that overlap on the GPU race.
constexpr size_t NOALLOC_MAX_KEYS = (size_t)1 << 20;  // m * row splits * k keys of 8 bytes
constexpr size_t NOALLOC_MAX_QUERIES = 16384;         // one 4-byte ticket per query
static __device__ __align__(256) unsigned long long g_scan_scratch[NOALLOC_MAX_KEYS + NOALLOC_MAX_QUERIES / 2];

// In host code `g_scan_scratch` is a stand-in, not a device address: the address is asked for once per device, with that
// device current, and kept here.
constexpr int NOALLOC_MAX_DEVICES = 64;
static std::mutex g_noalloc_mu;
static unsigned long long* g_noalloc_ptr[NOALLOC_MAX_DEVICES] = {};
static cudaError_t g_noalloc_err[NOALLOC_MAX_DEVICES] = {};

template <int QT, int METRIC, int LPL> static void noalloc_raise_smem(expann_gpu* h) {
	using Cfg = ScanCfg<QT, SCAN_TR, SCAN_NWARP, SCAN_STAGES, METRIC, LPL>;
	const size_t smem = Cfg::smem_bytes((uint32_t)h->d_pad);
	if (smem <= 227 * 1024) ensure_dyn_smem(h, (const void*)scan_topk_f32_kernel<QT, SCAN_TR, SCAN_NWARP, SCAN_STAGES, METRIC, LPL>, smem);
}
template <int METRIC> static void noalloc_raise_smem_all(expann_gpu* h) {
	noalloc_raise_smem<1, METRIC, 1>(h);
	noalloc_raise_smem<1, METRIC, 4>(h);
	noalloc_raise_smem<4, METRIC, 1>(h);
	noalloc_raise_smem<4, METRIC, 4>(h);
	noalloc_raise_smem<8, METRIC, 1>(h);
	noalloc_raise_smem<8, METRIC, 4>(h);
}

// set-up (expann_gpu_build of a brute-force handle, the handle's device current): resolve the scratch array's device
// address for this device and raise the shared-memory limits the scan's instantiations need for this row length, so that
// the call itself has nothing left to set up. A failure is kept and reported by the call, build() goes on.
void scan_noalloc_setup(expann_gpu* h) {
	if (h->device < 0 || h->device >= NOALLOC_MAX_DEVICES) return;
	{
		std::lock_guard<std::mutex> lk(g_noalloc_mu);
		if (!g_noalloc_ptr[h->device]) {
			unsigned long long* ptr = nullptr;
			const cudaError_t e = cudaGetSymbolAddress((void**)&ptr, g_scan_scratch);
			g_noalloc_err[h->device] = e;
			if (e == cudaSuccess) g_noalloc_ptr[h->device] = ptr;
			else cudaGetLastError();
		}
	}
	if (h->cfg.metric == EXPANN_METRIC_IP) noalloc_raise_smem_all<METRIC_IP>(h);
	else noalloc_raise_smem_all<METRIC_L2>(h);
}

int scan_noalloc(expann_gpu* h, const float* d_q, size_t m, size_t k, uint64_t* d_keys_out, cudaStream_t st) {
	if (h->cfg.engine != EXPANN_ENGINE_BRUTE_FORCE) return fail(h, EXPANN_ERR_UNSUPPORTED, "scan_device_noalloc: brute-force handles only");
	if (!h->built) return fail(h, EXPANN_ERR_STATE, "query before build()");
	if (k == 0) return fail(h, EXPANN_ERR_ARG, "k must be > 0");
	if (k > 128) return fail(h, EXPANN_ERR_UNSUPPORTED, "scan: k > 128 not supported");
	if (m == 0) return EXPANN_OK;
	if (h->d != h->d_pad)
		return fail(h, EXPANN_ERR_UNSUPPORTED, "scan_device_noalloc: d must be a multiple of 32 (padding the queries would need a work buffer)");
	if (m > NOALLOC_MAX_QUERIES)
		return fail(h, EXPANN_ERR_UNSUPPORTED, "scan_device_noalloc: at most " + std::to_string(NOALLOC_MAX_QUERIES) + " queries per call (static ticket array)");
	if (h->device < 0 || h->device >= NOALLOC_MAX_DEVICES) return fail(h, EXPANN_ERR_UNSUPPORTED, "scan_device_noalloc: device ordinal beyond 63");
	unsigned long long* scratch = nullptr;
	{
		std::lock_guard<std::mutex> lk(g_noalloc_mu);
		scratch = g_noalloc_ptr[h->device];
		if (!scratch) return cuda_fail(h, g_noalloc_err[h->device] == cudaSuccess ? cudaErrorNotReady : g_noalloc_err[h->device],
		                               "scan_device_noalloc: cudaGetSymbolAddress of the static scratch array (set-up at build())");
	}
	// grid: the rule of launch_scan_f32_simt
	int qt = h->cfg.qtile;
	if (qt == 0) qt = m >= 8 ? 8 : (m >= 3 ? 4 : 1);
	if (qt != 1 && qt != 4 && qt != 8) return fail(h, EXPANN_ERR_ARG, "qtile must be 0, 1, 4 or 8");
	const uint32_t dp = (uint32_t)h->d_pad;
	if (qt == 8 && ScanCfg<8, SCAN_TR, SCAN_NWARP, SCAN_STAGES, 0, 1>::smem_bytes(dp) > 227 * 1024) qt = 4;
	if (qt == 4 && ScanCfg<4, SCAN_TR, SCAN_NWARP, SCAN_STAGES, 0, 1>::smem_bytes(dp) > 227 * 1024) qt = 1;
	const uint32_t qtiles = (uint32_t)((m + qt - 1) / qt);
	const uint32_t groups = (uint32_t)((h->n + SCAN_ROWS - 1) / SCAN_ROWS);
	int dev_sms = 148;
	cudaDeviceGetAttribute(&dev_sms, cudaDevAttrMultiProcessorCount, h->device);
	const uint32_t waves = std::max(1u, (qtiles + dev_sms - 1) / (uint32_t)dev_sms);
	const uint32_t want = std::max(1u, (uint32_t)((uint64_t)dev_sms * waves / qtiles));
	uint32_t nsplit = std::max(1u, std::min(want, std::max(1u, groups / 2u)));
	uint32_t rps = (uint32_t)(((h->n + nsplit - 1) / nsplit + 63) / 64 * 64);
	nsplit = (uint32_t)((h->n + rps - 1) / rps);
	ScanParams p{};
	p.queries = d_q;
	p.m = (uint32_t)m;
	p.k = (uint32_t)k;
	p.d_pad = dp;
	p.n = (uint32_t)h->n;
	p.rows_per_split = rps;
	p.nsplit = nsplit;
	p.partial = d_keys_out;
	if (nsplit > 1 || (uint32_t)h->cfg.id_offset != 0) {
		if (m * (size_t)nsplit * k > NOALLOC_MAX_KEYS)
			return fail(h, EXPANN_ERR_UNSUPPORTED, "scan_device_noalloc: m * row splits * k = " + std::to_string(m * (size_t)nsplit * k) +
			                                           " keys exceed the static scratch array (" + std::to_string(NOALLOC_MAX_KEYS) + " keys)");
		uint32_t* tickets = reinterpret_cast<uint32_t*>(scratch + NOALLOC_MAX_KEYS);
		// the array keeps the last call's contents: the key lists are overwritten before they are read, the tickets must
		// start at zero (a call that ran to its end leaves them zero; cleared anyway, behind earlier work of this stream)
		EXPANN_CUDA(h, cudaMemsetAsync(tickets, 0, m * 4, st));
		p.partial = reinterpret_cast<uint64_t*>(scratch);
		p.final_out = d_keys_out;
		p.counters = tickets;
		p.id_add = (uint32_t)h->cfg.id_offset;
	}
	const dim3 grid(qtiles, nsplit);
	const int rc = qt == 8 ? launch_m<8>(h, p, grid, st) : qt == 4 ? launch_m<4>(h, p, grid, st) : launch_m<1>(h, p, grid, st);
	if (rc) return rc;
	h->stats.scan_launches++;
	h->stats.last_nsplit = nsplit;
	h->stats.last_qtile = qt;
	return EXPANN_OK;
}

} // namespace expann
