/* libexpann_gpu.so — C ABI of the B200-native expANN query path.
 *
 * This is the drop-in boundary: every entry point replaces one member of the reference's
 * compile-time engine interface  template<class T, class Derived> struct ann_engine  (reference
 * src/ann_engine.h:16-29) or of the engines behind it. No C++ or torch types cross it: opaque handle, plain
 * pointers, sizes, negative status codes. Host-pointer entry points are what a reference-side engine shim
 * binds; the *_device entry points take HBM-resident buffers and a CUDA stream for
 * callers that already own device memory (bench.py / the multi-GPU merge).
 *
 * Limits (the reference's query_k takes any k; beyond these a call returns EXPANN_ERR_UNSUPPORTED, nothing is truncated
 * silently): scans k <= 128; quantised scan k' (survivors re-ranked in fp32) <= 128; graph engine max(k, ef) <= 256 and
 * vertex degree <= 256; GPU graph build 2 <= M, M0 <= 96 and at most 127 candidates per vertex; ids: n + id_offset <
 * 2^32 - 1 per handle or per GPU shard (graph engine: n < 2^31). Memory beyond the vectors: the graph engine keeps one
 * visited byte per vertex per resident search slot (about 2960 slots: 3 GB at n = 1M) — shard larger bases over several
 * GPUs (device list below); the brute-force engine keeps a bf16 copy of the rows (2 or 6 bytes per component).
 *
 * One handle = one engine instance on one GPU (or, with a device list, on several GPUs of this process) = one host thread at a time (the reference engines are not
 * re-entrant either: src/antitopo_engine.h:208-209). Distinct handles are independent.
 */
#ifndef EXPANN_GPU_H
#define EXPANN_GPU_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct expann_gpu expann_gpu_t;

enum { EXPANN_OK = 0, EXPANN_ERR_ARG = -1, EXPANN_ERR_CUDA = -2, EXPANN_ERR_STATE = -3, EXPANN_ERR_UNSUPPORTED = -4,
       EXPANN_ERR_IO = -5 };
enum { EXPANN_ENGINE_BRUTE_FORCE = 0, /* brute_force_engine<float>, src/brute_force_engine.h:10-46 */
       EXPANN_ENGINE_GRAPH = 1,       /* antitopo_engine<float> query half, src/antitopo_engine.h:495-928 */
       EXPANN_ENGINE_QUANT_SCAN = 2   /* quantizer.h compressed scan + fp32 rerank, src/quantizer.h:20-150 */ };
enum { EXPANN_METRIC_L2 = 0, /* dist2, src/vec.h:139-142 */
       EXPANN_METRIC_IP = 1  /* key = -dot, src/vec.h:109-111 (no reference engine ranks by it) */ };

typedef struct expann_gpu_cfg {
	int32_t engine;          /* EXPANN_ENGINE_* */
	int32_t metric;          /* EXPANN_METRIC_* */
	int32_t device;          /* CUDA ordinal; -1 = current device */
	int32_t qtile;           /* queries per CTA in the scan: 0 = auto, else 1, 4 or 8 */
	uint64_t id_offset;      /* global id of local row 0 (base-set sharding) */
	/* graph engine (antitopo_engine_config, src/antitopo_engine.h:66-101) */
	uint64_t M, M0, ef_search_mult;
	int32_t use_compression; /* src/antitopo_engine.h:197-199: quantizer_simple<uint8_t> codes */
	int32_t quantizer;       /* QUANT_SCAN: 0 = quantizer_simple<uint8_t> (src/quantizer.h:125-150),
	                            1 = quantizer_ranged_q8 (src/quantizer.h:186-238) */
	/* device list: n_devices > 1 puts GPUs device .. device + n_devices - 1 behind this ONE handle; the
	   base set is sharded by contiguous row ranges, every call fans out and the per-GPU top-k lists are merged on the
	   first GPU. expann_gpu_create_multi takes an explicit list instead. row_groups: R > 0 = R row shards, each held by
	   n_devices / R GPUs that split the queries of a call between them (R = n_devices: one shard per GPU, nothing
	   replicated); 0 = automatic: shards of up to 2^20 rows, the GPUs that are not needed to hold the base split the
	   queries (option "mgpu_row_groups"). The result does not depend on either. */
	int32_t n_devices, row_groups;
	int32_t reserved[4];
} expann_gpu_cfg;

typedef struct expann_gpu_stats {
	uint64_t n, d, d_pad;
	uint64_t scan_launches, merge_launches, graph_launches, aux_launches; /* kernels launched so far */
	uint64_t last_nsplit, last_qtile;
	uint64_t distcomps; /* RECORD_STATS num_distcomps analogue (src/antitopo_engine.h:125-129), graph engine */
	uint64_t hot_time_ns, hot_timed_launches; /* CUDA-event time of the dominant kernel, see expann_gpu_set_timing */
	uint64_t tc_launches, tc_fallbacks;       /* tensor-core filter launches; batches redone by the SIMT scan (counted on
	                                             the device: expann_gpu_stats() synchronises the device to read it) */
	uint64_t tc_mode;                         /* operand format of the tensor-core filter: 0 off, 1 bf16-exact integers (error 0),
	                                             3 one bf16 term per value (real-valued data; slack 2^-6 |q| max|x|), 2 two-term bf16
	                                             split (3x the flops, slack 2^-14 |q| max|x|), 4 tf32 on the fp32 rows themselves (twice
	                                             the MMA time of 3, slack 2^-8.4 |q| max|x|, no operand copy). build() picks 1, else
	                                             probes 3, then 4, then takes 2; a batch that had to be redone moves the handle on
	                                             (3 -> 4 -> 2) */
	uint64_t peer_bytes;                      /* multi-device handle: bytes moved between GPUs so far (query blocks out over
	                                             NVLink, [m][k] key lists stored into the first GPU's gather buffer) */
	uint64_t query_groups, row_groups;        /* multi-device handle: decomposition n_devices = query_groups x row_groups (else 0) */
} expann_gpu_stats_t;

/* ann_engine() / Engine(config) constructor — src/ann_engine.h:17, src/bench_runner.h:33 */
int expann_gpu_create(const expann_gpu_cfg* cfg, expann_gpu_t** out);
/* the same engine over an explicit list of CUDA ordinals (cfg->device / n_devices are ignored). Every other entry point
 * works on the handle unchanged: add / build shard the rows, query / query_device fan out over the GPUs (device pointers
 * and the stream of query_device belong to devices[0]) and return the merged result, stats() sums the launches and
 * reports the slowest GPU's kernel time. An ordinal may repeat (several shards on one GPU). Multi-device handles are
 * immutable after build(): add() then fails with EXPANN_ERR_STATE; save_index / set_graph are single-device only;
 * load_index replicates the graph on every GPU and splits the queries. stats() of a multi-device handle: hot_time_ns is
 * the slowest GPU's. No host thread per GPU, no NCCL: cross-GPU ordering is by CUDA events, the [m][k] key lists are
 * stored by each GPU's last kernel straight into the first GPU's memory over NVLink. */
int expann_gpu_create_multi(const expann_gpu_cfg* cfg, const int32_t* devices, size_t n_devices, expann_gpu_t** out);
void expann_gpu_destroy(expann_gpu_t* h);
/* last error text of this handle (or of the failed create when h == NULL); never NULL */
const char* expann_gpu_last_error(const expann_gpu_t* h);

/* store_vector(const vec<T>&) × n — src/ann_engine.h:22-24, src/brute_force_engine.h:19-22. Rows are COPIED
 * (the reference's engines copy too); `rows` is host memory, row-major [n][d]. d is fixed by the first call. */
int expann_gpu_add(expann_gpu_t* h, const float* rows, size_t n, size_t d);
/* same, rows already in HBM on the handle's device */
int expann_gpu_add_device(expann_gpu_t* h, const float* d_rows, size_t n, size_t d, void* stream);
/* Graph hand-off (graph engine only). The reference builds its graph on the CPU (src/antitopo_engine.h:310-465)
 * and dumps it with serialize (:931-991); load_index reads that file like deserialize (:994-1074): the file's
 * vectors replace anything stored, exactly as read_index does (:312-313,476-481). set_graph takes the same graph
 * as CSR arrays over (vertex, layer) pairs in vertex-major order: nlayers[n], offsets[sum(nlayers)+1], edges[]. */
int expann_gpu_load_index(expann_gpu_t* h, const char* path);
int expann_gpu_set_graph(expann_gpu_t* h, size_t n, const uint32_t* nlayers, const uint64_t* offsets,
                         const uint32_t* edges, uint64_t starting_vertex, uint64_t max_layer);
/* GPU graph construction from the stored vectors: exact K-NN candidates (K1 scan), the
 * reference's diversity prune (prune_edges, src/antitopo_engine.h:262-308: ortho_factor / ortho_bias /
 * prune_overflow), reverse edges with lazy prune (:442-455), levels floor(-ln U / ln M) from mt19937(seed)
 * (:159,323). K = 0 picks min(127, 2*M0). Batch construction: not bit-identical to the sequential insert.
 * save_index writes the reference's serialize format (:931-991) so the reference can deserialize it (:994-1074). */
int expann_gpu_build_graph(expann_gpu_t* h, uint64_t M, uint64_t M0, uint64_t K, float ortho_factor, float ortho_bias,
                           uint64_t prune_overflow, uint64_t seed);
int expann_gpu_save_index(expann_gpu_t* h, const char* path);
/* build() — src/ann_engine.h:25, src/brute_force_engine.h:24-26 (the reference asserts n > 0; here n == 0
 * returns EXPANN_ERR_STATE) */
int expann_gpu_build(expann_gpu_t* h);

/* query_k(const vec<T>&, size_t k) × m — src/ann_engine.h:26-28, src/brute_force_engine.h:29-46.
 * q: host [m][d]. Outputs (host, caller-owned): ids_out [m][k] nearest first under (key asc, id asc), unused
 * slots UINT64_MAX; dist_out [m][k] (may be NULL) the fp32 ranking key (dist2, or -dot), unused slots +inf;
 * counts_out [m] (may be NULL) = min(k, n). ef: beam width for the graph engine (0 = k*ef_search_mult,
 * src/antitopo_engine.h:858-859), ignored by scans. Includes H2D of q and D2H of the results: page-locked caller
 * buffers (cudaHostAlloc / cudaHostRegister) are used by the copy engines directly, pageable ones are staged through
 * the handle's own page-locked buffers (one extra host memcpy each way). */
int expann_gpu_query(expann_gpu_t* h, const float* q, size_t m, size_t k, size_t ef, uint64_t* ids_out,
                     float* dist_out, uint64_t* counts_out);
/* HBM-resident form: d_q device [m][d] (row stride d), d_keys_out device [m][k] 64-bit keys
 * (orderable(dist) << 32 | local id + id_offset), asynchronous on `stream`. */
int expann_gpu_query_device(expann_gpu_t* h, const float* d_q, size_t m, size_t k, size_t ef,
                            uint64_t* d_keys_out, void* stream);
/* No counterpart in the reference: added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it). The exact fp32 scan (K1,
 * scan_topk_f32_kernel) of expann_gpu_query_device on a built single-device brute-force handle, with the kernel's
 * scratch memory -- the per-row-split key lists and the tickets of its fused merge -- in a static device array of the
 * library (8 MiB + 64 KiB, held by every GPU that loads the library until exit) instead of the handle's grow-only
 * buffers: the call makes no cudaMalloc / cudaFree and uses no allocating library routine. d_q device [m][d], d_keys_out
 * device [m][k], keys bit-identical to expann_gpu_query_device's, asynchronous on `stream`. Limits, beyond which the
 * call returns EXPANN_ERR_UNSUPPORTED with the reason in last_error and launches nothing: m <= 16384 queries;
 * m * row_splits * k <= 2^20 keys whenever the lists pass through the scratch array, i.e. with more than one row split
 * (row_splits = SM count / ceil(m / 8) for batches below ~1184 queries: at most ~1184 * k keys) or on a handle with an
 * id offset (m * k keys); k <= 128; d a multiple of 32 (no padded copy of the queries is made). ALL calls on one
 * GPU share the array (across handles too): calls that overlap on the GPU -- different streams without an ordering
 * between them -- race; calls on one stream, or ordered by events, are safe. */
int expann_gpu_scan_device_noalloc(expann_gpu_t* h, const float* d_q, size_t m, size_t k, uint64_t* d_keys_out, void* stream);
/* K7: merge nlists key lists per query (list s of query q at d_keys_in + q*stride_q + s*stride_s, klist keys
 * each) into the k smallest -> d_keys_out [m][k]. Used for the per-GPU lists gathered over NCCL. */
int expann_gpu_merge_device(expann_gpu_t* h, const uint64_t* d_keys_in, size_t m, size_t nlists, size_t klist,
                            size_t stride_q, size_t stride_s, size_t k, uint64_t* d_keys_out, void* stream);
/* keys -> reference output form, all device pointers; d_dist / d_counts may be NULL */
int expann_gpu_decode_device(expann_gpu_t* h, const uint64_t* d_keys, size_t m, size_t k, uint64_t* d_ids,
                             float* d_dist, uint64_t* d_counts, void* stream);

/* quantizer_ranged_q8 parameters fitted by build() (scale_factor / offset, src/quantizer.h:217-227) */
int expann_gpu_get_quantizer(const expann_gpu_t* h, float* scale, float* offset);
/* Distance primitives on explicit pairs, for parity tests of the reference kernels (host pointers, npairs x d):
 * kind 0 fp32 sum (a-b)^2 (src/vec.h:139-142, src/distance.h:86-148) -> out_f; 1 fp32 dot (src/vec.h:109-111,
 * src/distance.h:181-190) -> out_f; 2 float query vs u8 code, dist2_compressed (src/antitopo_engine.h:38-61) -> out_i;
 * 3 int8 and 4 int16 squared distance in the mathematically correct form (src/distance.h:14-53 intend it) -> out_i. */
int expann_gpu_pair_distances(int device, int kind, const void* a, const void* b, size_t npairs, size_t d,
                              float* out_f, int32_t* out_i);
/* on != 0: bracket every launch of the engine's dominant kernel (scan / graph search) with CUDA events on the
 * launching stream; expann_gpu_stats() then synchronises those events and reports the accumulated device time
 * (total_query_time analogue of RECORD_STATS, src/antitopo_engine.h:125-129,920-926). */
int expann_gpu_set_timing(expann_gpu_t* h, int on);
int expann_gpu_stats(expann_gpu_t* h, expann_gpu_stats_t* out);
/* tuning knobs by name (the reference exposes its knobs as Engine::config members, src/bench_runner.h:133-172):
 * "qtile" (0 auto | 1 | 4 | 8), "nsplit" (0 auto | forced row splits per scan), "kprime" (QUANT_SCAN: survivors
 * re-ranked in fp32, 0 auto = min(128, 10k)), "graph_occ" (GRAPH: resident CTAs of 4 warps per SM, 2 | 4 | 5 | 8,
 * default 5, before build()), "graph_cta_batch" (GRAPH: calls with at most this many queries take the latency form
 * of the search, one CTA per query; -1 automatic = twice the SM count, five times with the neighbourhood blocks of "graph_inline"; 0 never), "graph_cta_width" (vertices the latency form expands per hop: 1 (default) =
 * exactly the reference's visit order, 2 = the two closest unexpanded beam entries together: half the dependent hops,
 * measured recall >= at every beam width for 2 % more distance evaluations), "graph_upper_compact" (0/1, before build():
 * compact copy of the upper layers for the latency form, default 1), "graph_hash" (0/1, default 0: the one-warp-per-query form
 * keeps the visited set in a per-warp shared-memory hash, exact, moving to the per-vertex bytes in HBM when a query fills it;
 * measured 9-20 % slower than the bytes in HBM, which mostly hit L2), "graph_inline" (-1 automatic (default) | 0 | 1 | 2; before
 * build() it decides whether the blocks are built, after it 0 / non-0 switches their use off and on: the latency form reads NEIGHBOURHOOD BLOCKS — per vertex the ids of its neighbours and copies of their vectors,
 * contiguous, n x (M0 x d x 4 + 128) bytes of HBM (16.5 GB per million vertices at M0 = 32, d = 128) — one memory round trip
 * per hop instead of two, same ids and distances bit for bit; automatic = built when the blocks take at most a third of the
 * free device memory, 1 = whenever they fit, 2 = as 1 with the visited set in HBM bytes instead of the shared-memory bitmap (the
 * form graphs above ~1.7 M vertices get; for tests); with use_compression the bottom blocks hold the 8-bit codes (n x (M0 x d + 128)
 * bytes) and only the one-CTA-per-query form reads them; not used with graph_cta_width 2), "graph_inline_batch"
 * (0/1, default 1: the one-warp-per-query form gathers from the blocks too when they exist — same results, +8-12 % QPS).
 * Tensor-core candidate filter of the brute-force engine: "tc" (0/1; operands are prepared by build(), so enabling
 * needs to precede it), "tc_min_batch" (default 5: measured crossover with the SIMT scan on 1M x 128), "tc_pair" (0/1:
 * clusters of two CTAs for batches > 128, default 1), "tc_astream" (0 off | 1 when the query tile cannot stay resident
 * | 2 always), "tc_boot" (seeding launch over the leading rows, one wave of clusters: 0 off | 1 every format but the split, default | 2 always), "tc_boot_tiles" (its 256-row tiles, default 128), "tc_v1" (force the
 * chunked first form), "tc_real" (real-valued rows, before build(): 1 (default) one bf16 term per value, falling back to the
 * next format when a batch had to be redone | 0 the split from the start), "tc_tf32" (0/1, default 1, before build(): tf32 on the
 * fp32 rows (format 4) is the format between the one-term one and the split).
 * Quantised scan engine: "quant_tc" (0/1, default 1: batches of at least "quant_tc_min_batch" queries (default 32) run the
 * code scan as a tensor-core candidate filter, tcgen05.mma.kind::i8 on the 8-bit codes, with the SIMT scan as its queued
 * redo; same result).
 * Multi-device handles: "mgpu_row_groups" (0 (default) automatic: shards of up to 2^20 rows | R row shards, each on
 * n_devices / R GPUs that split the queries of a call; R = n_devices: one shard per GPU; before build()); every other knob is applied to each GPU's engine.
 * Unknown names: EXPANN_ERR_ARG. */
int expann_gpu_set_option(expann_gpu_t* h, const char* name, int64_t value);
/* library build info: "sm_100a;..." */
const char* expann_gpu_version(void);

#ifdef __cplusplus
}
#endif
#endif
