// GPU graph construction — kernels.
// The reference inserts vertices one at a time (src/antitopo_engine.h:310-465): candidates from an
// ef_construction-wide search, diversity prune (prune_edges, :262-308), reverse edges with lazy prune (:442-455).
// Here the same three ingredients run as batch passes per layer:
//   candidates = exact K nearest neighbours from the fp32 scan (K1),
//   forward prune = the reference's rule, one CTA per vertex (prune_select_kernel),
//   reverse edges = sort of (target, dist, source) tuples, then a lazy prune of forward ∪ reverse lists.
// Not bit-identical to the sequential build (stated in DESIGN.md); acceptance = recall/QPS curves.
#pragma once
#include "common.cuh"
#include "graph_search.cuh"

namespace expann {

constexpr int GB_THREADS = 128; // one CTA per vertex, one thread per candidate
constexpr int GB_MAXC = 128;    // max candidates per vertex

// scan keys [rows][kk] of the query rows row0.. (self included somewhere) -> cand [n][K] without self
__global__ void knn_strip_self_kernel(const uint64_t* __restrict__ keys, uint32_t kk, uint32_t row0, uint32_t rows,
                                      uint64_t* __restrict__ cand, uint32_t K) {
	const uint32_t r = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
	const int lane = threadIdx.x & 31;
	if (r >= rows) return;
	const uint64_t* src = keys + (size_t)r * kk;
	uint64_t* dst = cand + (size_t)(row0 + r) * K;
	const uint32_t self = row0 + r;
	uint32_t w = 0;
	for (uint32_t j0 = 0; j0 < kk; j0 += 32) {
		const uint32_t j = j0 + lane;
		const uint64_t key = j < kk ? src[j] : KEY_EMPTY;
		const bool ok = key != KEY_EMPTY && key_id(key) != self;
		const unsigned mk = __ballot_sync(FULL, ok);
		const uint32_t pos = w + __popc(mk & ((1u << lane) - 1));
		if (ok && pos < K) dst[pos] = key;
		w += __popc(mk);
	}
	for (uint32_t j = w + lane; j < K; j += 32) dst[j] = KEY_EMPTY;
}

struct PruneParams {
	const float* base;        // vectors, row stride d_pad
	const uint32_t* rowmap;   // local id -> base row (null = identity)
	const uint64_t* candA;    // [n][ka] keys (dist, local id), KEY_EMPTY padded
	const uint64_t* candB;    // [.] keys, segment of vertex v = candB[offB[v] .. offB[v+1]) (may be null)
	const uint32_t* offB;
	uint32_t* out_ids;        // [n][stride] local ids, NO_VERTEX padded
	float* out_len;           // [n][stride]
	uint32_t n, ka, capB, d_pad, stride, mcap, leniency;
	float ortho_factor, ortho_bias;
	int lazy;
};

// prune_edges (src/antitopo_engine.h:262-308): repeatedly take the candidate with the smallest
//   score(c) = d(v,c) + sum over kept p with d(p,c) < d(v,c) of (ortho_factor*(d(v,c)-d(p,c)) + ortho_bias),
// a candidate is dropped for good after `leniency` = prune_overflow+1 such violations; stop at mcap edges.
// lazy (:269-272): lists no longer than mcap are kept as they are.
__global__ void __launch_bounds__(GB_THREADS) prune_select_kernel(const PruneParams p) {
	extern __shared__ uint8_t gb_smem[];
	const uint32_t v = blockIdx.x;
	const int t = threadIdx.x;
	uint64_t* skey = reinterpret_cast<uint64_t*>(gb_smem);               // [GB_MAXC]
	float* svec = reinterpret_cast<float*>(gb_smem + GB_MAXC * 8);       // [GB_MAXC][d_pad+1]
	__shared__ uint32_t s_cnt, s_sel;
	__shared__ uint64_t s_red[GB_THREADS / 32];
	const uint32_t ld = p.d_pad + 1;

	// ---- gather candidate keys, sort (bitonic), drop duplicates / self / empties
	uint64_t key = KEY_EMPTY;
	if (t < (int)p.ka) key = p.candA[(size_t)v * p.ka + t];
	else if (p.candB) {
		const uint32_t b0 = p.offB[v], b1 = p.offB[v + 1];
		const uint32_t j = t - p.ka;
		if (j < p.capB && b0 + j < b1) key = p.candB[b0 + j];
	}
	if (key != KEY_EMPTY && key_id(key) == v) key = KEY_EMPTY;
	skey[t] = key;
	__syncthreads();
	for (int k = 2; k <= GB_MAXC; k <<= 1)
		for (int j = k >> 1; j > 0; j >>= 1) {
			const int ixj = t ^ j;
			if (ixj > t) {
				const uint64_t a = skey[t], b = skey[ixj];
				const bool up = (t & k) == 0;
				if ((a > b) == up) { skey[t] = b; skey[ixj] = a; }
			}
			__syncthreads();
		}
	key = skey[t];
	const bool uniq = key != KEY_EMPTY && (t == 0 || skey[t - 1] != key);
	__syncthreads();
	if (t == 0) s_cnt = 0;
	__syncthreads();
	// compact in order: warp-aggregated prefix
	{
		const unsigned mk = __ballot_sync(FULL, uniq);
		__shared__ uint32_t wbase[GB_THREADS / 32];
		if ((t & 31) == 0) wbase[t >> 5] = __popc(mk);
		__syncthreads();
		uint32_t off = 0;
		for (int w = 0; w < (t >> 5); ++w) off += wbase[w];
		if (uniq) skey[off + __popc(mk & ((1u << (t & 31)) - 1))] = key;
		if (t == GB_THREADS - 1) s_cnt = off + __popc(mk);
		__syncthreads();
	}
	const uint32_t cnt = s_cnt;
	uint32_t* oid = p.out_ids + (size_t)v * p.stride;
	float* olen = p.out_len + (size_t)v * p.stride;
	if (p.lazy && cnt <= p.mcap) {
		for (uint32_t j = t; j < p.stride; j += GB_THREADS) {
			oid[j] = j < cnt ? key_id(skey[j]) : NO_VERTEX;
			olen[j] = j < cnt ? key_dist(skey[j]) : 0.f;
		}
		return;
	}
	// ---- candidate vectors -> smem (row pitch d_pad+1: conflict-free thread-per-row reads)
	for (uint32_t c = 0; c < cnt; ++c) {
		const uint32_t id = key_id(skey[c]);
		const float* src = p.base + (size_t)(p.rowmap ? p.rowmap[id] : id) * p.d_pad;
		for (uint32_t i = t; i < p.d_pad; i += GB_THREADS) svec[c * ld + i] = src[i];
	}
	__syncthreads();
	const bool mine = t < (int)cnt;
	const float basic = mine ? key_dist(skey[t]) : 0.f;
	float score = basic;
	uint32_t viol = 0;
	bool alive = mine;
	uint32_t nsel = 0;
	while (nsel < p.mcap) {
		// argmin over (score, then the candidates' (dist,id) order == position t)
		uint64_t mykey = alive ? (((uint64_t)f32_orderable(score) << 32) | (uint32_t)t) : KEY_EMPTY;
#pragma unroll
		for (int o = 16; o >= 1; o >>= 1) {
			const uint64_t other = shfl_u64(mykey, (t & 31) ^ o);
			if (other < mykey) mykey = other;
		}
		if ((t & 31) == 0) s_red[t >> 5] = mykey;
		__syncthreads();
		if (t == 0) {
			uint64_t best = s_red[0];
			for (int w = 1; w < GB_THREADS / 32; ++w) if (s_red[w] < best) best = s_red[w];
			s_sel = best == KEY_EMPTY ? NO_VERTEX : (uint32_t)best;
		}
		__syncthreads();
		const uint32_t sel = s_sel;
		if (sel == NO_VERTEX) break;
		if (t == 0) { oid[nsel] = key_id(skey[sel]); olen[nsel] = key_dist(skey[sel]); }
		nsel++;
		if (t == (int)sel) alive = false;
		if (alive) {
			const float* a = svec + sel * ld;
			const float* b = svec + t * ld;
			float co = 0.f;
			for (uint32_t i = 0; i < p.d_pad; ++i) { const float df = a[i] - b[i]; co = fmaf(df, df, co); }
			if (co < basic) {
				score += p.ortho_factor * (basic - co) + p.ortho_bias;
				if (++viol >= p.leniency) alive = false;
			}
		}
		__syncthreads();
	}
	for (uint32_t j = nsel + t; j < p.stride; j += GB_THREADS) { oid[j] = NO_VERTEX; olen[j] = 0.f; }
}

// forward lists -> (target<<32 | orderable(dist), source) tuples for the reverse-edge sort; also a histogram
__global__ void reverse_tuples_kernel(const uint32_t* __restrict__ ids, const float* __restrict__ len, uint32_t n,
                                      uint32_t stride, uint64_t* __restrict__ tkey, uint32_t* __restrict__ tval,
                                      uint32_t* __restrict__ hist) {
	const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= (size_t)n * stride) return;
	const uint32_t u = ids[i];
	if (u == NO_VERTEX) { tkey[i] = KEY_EMPTY; tval[i] = NO_VERTEX; return; }
	tkey[i] = ((uint64_t)u << 32) | f32_orderable(len[i]);
	tval[i] = (uint32_t)(i / stride);
	atomicAdd(&hist[u], 1u);
}
// sorted tuples -> candidate keys (dist, source) in place of the sort keys
__global__ void reverse_keys_kernel(const uint64_t* __restrict__ tkey, const uint32_t* __restrict__ tval, size_t total,
                                    uint64_t* __restrict__ out) {
	const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= total) return;
	out[i] = tkey[i] == KEY_EMPTY ? KEY_EMPTY : ((tkey[i] << 32) | tval[i]);
}
// forward (ids,len) lists -> keys [n][stride] (for the candA input of the second prune)
__global__ void lists_to_keys_kernel(const uint32_t* __restrict__ ids, const float* __restrict__ len, size_t total,
                                     uint64_t* __restrict__ out) {
	const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= total) return;
	out[i] = ids[i] == NO_VERTEX ? KEY_EMPTY : make_key(len[i], ids[i]);
}
__global__ void gather_rows_kernel(const float* __restrict__ base, const uint32_t* __restrict__ rowmap, uint32_t ns,
                                   uint32_t d_pad, float* __restrict__ out) {
	const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= (size_t)ns * d_pad) return;
	out[i] = base[(size_t)rowmap[i / d_pad] * d_pad + i % d_pad];
}

} // namespace expann
