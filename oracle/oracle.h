/* TEST INFRASTRUCTURE — CPU restatement of the expANN query hot path.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load this
 * library, and only as the checker or the timed CPU baseline. The product path (expann_b200/) never links,
 * imports or calls anything under oracle/.
 *
 * Parity status: PINNED by execution — every function here is checked (tests/test_oracle_pin.py, run in the
 * build container) against the reference's own code compiled unmodified into oracle/_ref/ (oracle/Makefile,
 * oracle/ref_driver.cpp) and against the frozen outputs of that code in tests/golden/ (npz files). The reference
 * itself ships no tests or golden vectors, so the pins are minted from the reference run here.
 * Functions with no reference engine behind them (inner-product ranking, sign-corrected int8/int16 scorers)
 * are marked "parity unpinned" below.
 *
 * Every function cites the reference file:line it restates (paths relative to /root/reference).
 */
#ifndef EXPANN_ORACLE_H
#define EXPANN_ORACLE_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

enum { ORC_L2 = 0, ORC_IP = 1 };

/* ---- distance primitives ------------------------------------------------------------------------------ */
/* src/distance.h:136-148 (== :86-111 for d%128==0): 16 fp32 lanes, lane j accumulates fma(diff,diff,acc)
 * over dims j, j+16, ...; lanes reduced as _mm512_reduce_add_ps does (8+8, 4+4, 2+2, 1+1). Tail dims
 * (d%16, never reached by the reference's AVX kernels) are added sequentially. Bit-exact vs the reference for
 * d%16==0. */
float orc_dist2_f32(const float* a, const float* b, size_t d);
/* src/distance.h:181-190 (dot_avx512f_f16), same lane structure. */
float orc_dot_f32(const float* a, const float* b, size_t d);
/* src/antitopo_engine.h:38-61 + query conversion :726-737: sum over dims of (uint32(q_i) - code_i)^2 in
 * wrapping 32-bit arithmetic. Lane order is irrelevant (integer add is associative), so no swizzle. */
int32_t orc_dist2_u8(const float* q, const uint8_t* code, size_t d);
/* src/distance.h:29-53 AS WRITTEN: 8-bit wrapping subtract, zero-extended, squared (wrong for negative
 * differences). Kept only to pin the oracle's understanding of the reference bug. */
int32_t orc_dist2_i8_asref(const int8_t* a, const int8_t* b, size_t d);
/* sign-corrected int8 / int16 squared distance (what src/distance.h:14-53 intend). parity unpinned. */
int32_t orc_dist2_i8(const int8_t* a, const int8_t* b, size_t d);
/* src/distance.h:14-27 AS WRITTEN: 16-bit wrapping sub and mullo, then widened with madd(1, x) (sign-extends
 * the wrapped 16-bit square). */
int32_t orc_dist2_i16_asref(const int16_t* a, const int16_t* b, size_t d);
int32_t orc_dist2_i16(const int16_t* a, const int16_t* b, size_t d);

/* ---- quantizers --------------------------------------------------------------------------------------- */
/* src/quantizer.h:132-141 quantizer_simple<uint8_t>::build: code = (uint8_t)x (C++ float->u8 conversion;
 * defined only for x in [0,256), which is the only domain the compressed path is meaningful on). */
void orc_quant_simple_u8(const float* rows, size_t n, size_t d, uint8_t* codes);
/* src/quantizer.h:206-227 quantizer_ranged_q8::build: global min/max (initial max_val=FLT_MIN, min_val=FLT_MAX
 * as in :210-211), scale = 128/(max-min), offset = -scale*min. */
void orc_quant_ranged_q8_fit(const float* rows, size_t n, size_t d, float* scale, float* offset);
/* src/quantizer.h:196-205 convert_single: size_t(round(scale*x+offset)) clamped to [0,127]; negative
 * pre-clamp values wrap through size_t to 127 exactly as the reference does. */
void orc_quant_ranged_q8_encode(const float* rows, size_t n, size_t d, float scale, float offset,
                                int8_t* codes);

/* ---- top-k container ---------------------------------------------------------------------------------- */
/* src/topk_t.h:9-67: bounded max-heap on (dist,id) + dedupe set. */
typedef struct orc_topk orc_topk;
orc_topk* orc_topk_create(size_t k);
void orc_topk_destroy(orc_topk* t);
int orc_topk_consider(orc_topk* t, float d, uint64_t v);      /* :24-35 */
void orc_topk_discard_until_size(orc_topk* t, size_t goal);  /* :36-41 */
size_t orc_topk_size(const orc_topk* t);
uint64_t orc_topk_worst(const orc_topk* t);                   /* :42 */
float orc_topk_worst_val(const orc_topk* t);                  /* :44 */
int orc_topk_at_capacity(const orc_topk* t);                  /* :43 */
size_t orc_topk_to_combined(const orc_topk* t, float* d_out, uint64_t* id_out); /* :56-66 ascending */

/* ---- exhaustive scans --------------------------------------------------------------------------------- */
/* src/brute_force_engine.h:29-46. metric ORC_L2 ranks by orc_dist2_f32; ORC_IP (parity unpinned — no
 * reference engine ranks by inner product) ranks by -orc_dot_f32. Result: k smallest under
 * (key asc, id asc), nearest first; ids padded with UINT64_MAX, dists with +inf; counts[i]=min(k,n).
 * dists_out holds the ranking key (dist2, or -dot). `threads` host threads split the queries. */
void orc_bf_query(const float* base, size_t n, size_t d, const float* q, size_t m, size_t k, int metric,
                  uint64_t* ids_out, float* dists_out, uint64_t* counts_out, int threads);
/* exhaustive compressed scan (quantizer_simple<u8> codes scored with a8), keeping the
 * kprime smallest (score asc, id asc); then fp32 rerank of those kprime with orc_dist2_f32 / -dot and a
 * final (key asc, id asc) sort, truncated to k. metric ORC_IP scores codes with -(sum uint32(q_i)*code_i). */
void orc_scan_u8_query(const float* base, const uint8_t* codes, size_t n, size_t d, const float* q, size_t m,
                       size_t k, size_t kprime, int metric, uint64_t* ids_out, float* dists_out,
                       int32_t* scores_out /* m*kprime, compressed scores of the survivors, may be NULL */,
                       uint64_t* cand_out /* m*kprime survivor ids, may be NULL */, int threads);

/* ---- graph (antitopo_engine) -------------------------------------------------------------------------- */
typedef struct orc_graph orc_graph;
/* src/antitopo_engine.h:994-1074 deserialize (index file written by :931-991). */
orc_graph* orc_graph_load(const char* path);
/* build from arrays: level[v] = number of layers of v; CSR over (v,layer) pairs in v-major order. */
orc_graph* orc_graph_from_arrays(const float* vectors, size_t n, size_t d, const uint32_t* nlayers,
                                 const uint64_t* offsets /* sum(nlayers)+1 */, const uint32_t* edges,
                                 uint64_t starting_vertex, uint64_t max_layer, uint64_t M, uint64_t M0);
void orc_graph_destroy(orc_graph* g);
uint64_t orc_graph_info(const orc_graph* g, int what); /* 0 start,1 max_layer,2 n,3 M,4 M0,5 d */
/* src/antitopo_engine.h:853-928 _query_k: upper-layer greedy descent (:861-903), then
 * query_k_at_layer<true,false,false> (:495-708) or query_k_bottom_compressed (:710-851) with width ef,
 * truncated to k. std::priority_queue behaviour (libstdc++ push_heap/pop_heap/make_heap) is restated so that
 * tie handling matches the reference build. distcomps (optional) accumulates the RECORD_STATS counter. */
void orc_graph_query(orc_graph* g, const float* q, size_t m, size_t k, size_t ef, int use_compression,
                     uint64_t* ids_out, float* dists_out, uint64_t* counts_out, uint64_t* distcomps);
/* multi-threaded variant for the CPU baseline: one visited array per thread. */
void orc_graph_query_mt(orc_graph* g, const float* q, size_t m, size_t k, size_t ef, int use_compression,
                        uint64_t* ids_out, uint64_t* counts_out, int threads);
/* export for building device layouts in tests: bottom-layer degree and neighbours of v */
size_t orc_graph_neighbours(const orc_graph* g, uint64_t v, uint32_t layer, uint32_t* out, size_t cap);
const float* orc_graph_vectors(const orc_graph* g);

#ifdef __cplusplus
}
#endif
#endif
