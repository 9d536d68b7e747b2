// K5, latency form on NEIGHBOURHOOD BLOCKS (round 2): one CTA per query, one memory round trip per hop, the beam
// bookkeeping of hop h under the memory trip of hop h+1.
//
// Same search as graph_search_cta_kernel (graph_search.cuh) — antitopo_engine<T>::_query_k, reference
// src/antitopo_engine.h:853-928: greedy descent through the upper layers (:878-893), best-first beam search of width ef on
// the bottom layer (query_k_at_layer<true,false,false>, :495-708), same visit order, same admission, same per-vertex
// summation order, hence the same ids and distances bit for bit — over a different layout of the index in HBM and with a
// different division of labour inside the CTA:
//
//   * block(v) = [ids of v's M0 neighbours | THEIR VECTORS], contiguous (M0 = 32, d = 128: 16.5 KB per vertex, 16.5 GB per
//     million vertices: the 180 GB of HBM3e are spent on latency). Expanding v reads ids and vectors in ONE round trip
//     (the addresses of the vectors do not depend on the ids) instead of the chain adjacency row -> gathered vectors; the
//     upper layers have the same form (block(row) = [dense indices | {first upper row, layers, global id} | vectors]).
//   * the visited set is a bitmap of the whole vertex set in shared memory (n <= ~1.7 M: 1 bit per vertex, cleared per
//     query in ~0.5 us) — no memory round trip; larger graphs keep the epoch-stamped bytes in HBM.
//   * warp 0 is the CONTROLLER (owns the beam), warps 1..8 are SCORERS (32 teams of 8 lanes, one neighbour each). They
//     meet on mbarriers, not bar.sync. As soon as the scores of hop h are in, the vertex hop h+1 expands is known WITHOUT
//     touching the beam: it is the smaller of the closest fresh neighbour just scored (if the beam admits it) and the
//     closest unexpanded beam entry as of the previous admission — exactly what "admit, then take the closest unexpanded
//     entry" yields. The controller drives the upper layers; on the bottom layer every scorer warp works that minimum out
//     for itself (the scores and {r1, r2, bound} are in shared memory) and starts its loads, while the controller does the
//     reference's own sequence one hop behind: merge hop h's candidates into the beam, mark the closest unexpanded entry
//     (the vertex the scorers are on), leave the two after it and the new bound for the next decision. That work runs
//     under the scorers' memory trip.
//     The measured cost of a hop was instruction latency in the beam-owning warp, not memory (ncu: 233 k cycles per query,
//     0.4 long-scoreboard stalls per issue); see profiles/r2_notes.md section 12.
#pragma once
#include "graph_search.cuh"

namespace expann {

constexpr int GI_SCORERS = 8;                   // scorer warps (warps 1..8); warp 0 is the controller
constexpr int GI_THREADS = 32 * (GI_SCORERS + 1);

struct InlineParams {
	GraphParams g;
	const uint8_t* blk_bottom; // [n] blocks of bb_bytes: ids (m0s x u32, padded to bb_idb) | vectors (m0s x d_pad x f32)
	const uint8_t* blk_upper;  // [upper rows] blocks of ub_bytes: dense indices (ms x u32, padded to ub_idb) | meta (ms x uint4) | vectors
	uint32_t bb_bytes, bb_idb, ub_bytes, ub_idb;
	uint32_t bitmap_words; // visited bitmap in shared memory (0: epoch-stamped bytes in HBM)
};

__device__ __forceinline__ void bulk_prefetch_l2(const void* gsrc, uint32_t bytes) {
	asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(gsrc), "r"(bytes) : "memory");
}
__device__ __forceinline__ uint64_t warp_min_u64(uint64_t v) {
	const uint32_t hi = (uint32_t)(v >> 32), mh = __reduce_min_sync(FULL, hi);
	const uint32_t ml = __reduce_min_sync(FULL, hi == mh ? (uint32_t)v : 0xFFFFFFFFu);
	return ((uint64_t)mh << 32) | ml;
}

// Rendezvous between the warps of the CTA: mbarrier arrive / try_wait (the waiter sleeps in hardware). -DGI_POLL: counters in
// shared memory that the waiting warp polls instead — measured SLOWER (one query per call 78.8 -> 86.1 us at ef 40: the
// polling warps take issue slots from the warps they wait for). The k-th use (k = 0, 1, ...) of a polled rendezvous with A
// arrivals is complete when its word has reached (k + 1) * A.
#ifndef GI_POLL
#define GI_INIT(bar, arrivals) mbar_init(bar, arrivals)
#define GI_WAIT(bar, k, arrivals) mbar_wait(bar, (k) & 1)
#define GI_ARRIVE(bar) mbar_arrive(bar)
#else
#define GI_INIT(bar, arrivals) (*reinterpret_cast<volatile uint32_t*>(bar) = 0u)
#define GI_WAIT(bar, k, arrivals)                                                            \
	do {                                                                                     \
		while (*reinterpret_cast<volatile uint32_t*>(bar) < ((k) + 1u) * (arrivals)) { }     \
		__threadfence_block();                                                               \
	} while (0)
#define GI_ARRIVE(bar)                                                                       \
	do {                                                                                     \
		__threadfence_block();                                                               \
		atomicAdd(reinterpret_cast<uint32_t*>(bar), 1u);                                     \
	} while (0)
#endif

#ifdef GI_PROFILE
__device__ unsigned gi_prof_calls;
#define GI_T(slot) do { const long long t_ = clock64(); pf[slot] += t_ - pf_t; pf_t = t_; } while (0)
#else
#define GI_T(slot) do { } while (0)
#endif

// The whole kernel is bound by instruction latency in whichever warp is on the critical path (every warp is alone or
// nearly alone on its scheduler), so the per-hop paths are specialised: NSTEP = d_pad / 32 (3 or 4: the lane's share of the
// query lives in registers and the loads are issued unconditionally; 0 = any dimension), rows of at most 32 neighbours
// (one per team) take a path without loops.
template <int LPL, int NSTEP, int NR, bool COMP>
__global__ void __launch_bounds__(GI_THREADS) graph_search_inline_kernel(const InlineParams p) {
	extern __shared__ __align__(128) uint8_t gi_smem[];
	const int tid = threadIdx.x, lane = tid & 31, wib = tid >> 5;
	const uint32_t slot = blockIdx.x;
	const uint32_t d_pad = p.g.d_pad, m0s = p.g.m0s, ms = p.g.ms, ef = p.g.ef;
	const bool one_round = m0s <= 32u * NR && ms <= 32u * NR; // NR = 1: one neighbour per team; NR = 4: up to four (rows of up to 128, the reference's M0 = 120)
	// shared memory: query | keys of a hop's scores x 2 | beam copy | staged candidates | ctl x 2 | state x 2 | barriers | bitmap
	float* qsm = reinterpret_cast<float*>(gi_smem);
	uint8_t* qcs = reinterpret_cast<uint8_t*>(qsm + d_pad);  // [d_pad] the query truncated to u8 (compressed graphs: src/antitopo_engine.h:726-737)
	uint64_t* rk = reinterpret_cast<uint64_t*>(qcs + d_pad); // [2][GS_MAXNB] bottom: beam key of a fresh neighbour; upper: orderable(dist) << 32 | position; KEY_EMPTY
	uint64_t* bm = rk + 2 * GS_MAXNB;                        // [32 * LPL] shared-memory copy of the beam
	uint64_t* ck = bm + 256;                                 // [32] candidate keys of one admission round
	uint4* ctl = reinterpret_cast<uint4*>(ck + 32);          // [2] job for the scorers: block pointer (lo, hi), mode (0 stop, 1 upper, 2 bottom), bound for their prefetch
	uint4* st = ctl + 2;                                     // [2][2] the beam after an admission: r1 (lo, hi), r2 (lo, hi) | bound, -, -, -
	uint64_t* bar_go = reinterpret_cast<uint64_t*>(st + 4);  // [2] controller -> scorers: a job in ctl
	uint64_t* bar_sc = bar_go + 2;                           // [2] scorers -> everybody: the scores of a job are in rk
	uint64_t* bar_adm = bar_sc + 2;                          // [2] controller -> scorers: admission done, st written
	uint32_t* bits = p.bitmap_words ? reinterpret_cast<uint32_t*>(bar_adm + 2) : nullptr;
	uint8_t* vis = p.g.visited + (size_t)slot * p.g.vis_stride;
	uint32_t epoch = p.g.epoch[slot];
	unsigned long long dc_total = 0;
	uint32_t h = 0; // jobs scored so far: job h uses rk / bar_sc h & 1, phase (h >> 1) & 1
	uint32_t Jg = 0, Jb = 0; // jobs handed over through ctl / decisions on the bottom layer so far (same counting in every warp)

	if (tid == 0) {
		for (int b = 0; b < 2; ++b) { GI_INIT(&bar_go[b], 1); GI_INIT(&bar_sc[b], GI_SCORERS); GI_INIT(&bar_adm[b], 1); }
#ifndef GI_POLL
		mbar_fence_init();
#endif
	}
	__syncthreads();

	for (uint32_t qi = slot; qi < p.g.m; qi += gridDim.x) {
		if (wib == 0 && lane == 0) { // the first blocks do not depend on the query
			if (p.g.max_layer > 1 && p.g.up_start_nl > 1) bulk_prefetch_l2(p.blk_upper + (size_t)(p.g.up_start_ui + min(p.g.max_layer, p.g.up_start_nl) - 2) * p.ub_bytes, p.ub_bytes);
			else bulk_prefetch_l2(p.blk_bottom + (size_t)p.g.start * p.bb_bytes, p.bb_bytes);
		}
		if (!bits) {
			epoch++;
			if ((epoch & 0xFF) == 0) {
				epoch++;
				for (size_t i = (size_t)tid * 16; i < p.g.vis_stride; i += GI_THREADS * 16)
					*reinterpret_cast<uint4*>(vis + i) = make_uint4(0, 0, 0, 0);
			}
		} else
			for (uint32_t i = tid * 4; i < p.bitmap_words; i += GI_THREADS * 4) *reinterpret_cast<uint4*>(bits + i) = make_uint4(0, 0, 0, 0);
		const uint8_t ep8 = (uint8_t)(epoch & 0xFF);
		for (uint32_t i = tid; i < d_pad; i += GI_THREADS) {
			const float qv = p.g.queries[(size_t)qi * d_pad + i];
			qsm[i] = qv;
			if (COMP) qcs[i] = (uint8_t)(uint32_t)qv;
		}
		__syncthreads();

		if (wib != 0) {
			// =============================== scorers ===============================
			const uint32_t team = (uint32_t)(tid - 32) >> 3, tl = tid & 7;
			const int nstep = d_pad >> 5;
			float4 qr[NSTEP > 0 ? NSTEP : 1];
#pragma unroll
			for (int s = 0; s < NSTEP; ++s) qr[s] = *reinterpret_cast<const float4*>(qsm + s * 32 + tl * 4);
			// where this lane reads in a block (rows of at most 32 neighbours: team t scores neighbour t)
			const uint32_t off_b = p.bb_idb + team * d_pad * 4 + tl * 16, off_u = p.ub_idb + ms * 16 + team * d_pad * 4 + tl * 16;
			const bool in_b = team < m0s, in_u = team < ms;
			// compressed graphs: a bottom block holds 8-bit codes, d_pad bytes per row, one 16-byte chunk per lane for d_pad <= 128
			const uint32_t nchunk = d_pad >> 4, off_c = p.bb_idb + team * d_pad + tl * 16;
			const uint4 qcr = COMP && tl < nchunk ? *reinterpret_cast<const uint4*>(qcs + tl * 16) : make_uint4(0, 0, 0, 0);
#ifdef GI_PROFILE
			long long sp[6] = { 0, 0, 0, 0, 0, 0 }, sp_t = clock64();
			int sp_jobs = 0;
#define GI_S(slot) do { const long long t_ = clock64(); sp[slot] += t_ - sp_t; sp_t = t_; } while (0)
#else
#define GI_S(slot) do { } while (0)
#endif
			// The controller hands over the steps through the upper layers and the entry into the bottom layer (ctl / bar_go).
			// From there on the scorers drive themselves: once a hop's scores are complete, every scorer warp works out the
			// vertex of the next hop for itself (the same two-instruction minimum the controller computes for its admission)
			// and starts its loads — no hand-over, no wake-up of another warp on the hop's path.
			bool self = false;
			uint4 job = make_uint4(0, 0, 0, 0);
			for (;;) {
				const uint32_t par = h & 1;
				if (!self) {
					GI_WAIT(&bar_go[Jg & 1], Jg >> 1, 1u);
					job = ctl[Jg & 1];
					++Jg;
				}
				GI_S(0);
				const uint8_t* blk = reinterpret_cast<const uint8_t*>(((uint64_t)job.y << 32) | job.x);
				uint64_t* out = rk + par * GS_MAXNB;
				const bool bottom = job.z == 2;
				uint32_t pfv = NO_VERTEX; // a neighbour likely to be expanded next: its block is asked into L2
				if (COMP && bottom && NSTEP > 0 && one_round) {
					// dist2_compressed (src/antitopo_engine.h:38-61) on the block's 8-bit codes: one 16-byte load per lane
					uint32_t v = NO_VERTEX;
					uint4 xc = make_uint4(0, 0, 0, 0);
					if (in_b) {
						v = __ldg(reinterpret_cast<const uint32_t*>(blk) + team);
						if (tl < nchunk) xc = __ldg(reinterpret_cast<const uint4*>(blk + off_c));
					}
					bool seen = false;
					uint8_t seenb = 0;
					if (tl == 0 && v != NO_VERTEX) { // visited test-and-set (:596-605) under the code loads
						if (bits) seen = (atomicOr(bits + (v >> 5), 1u << (v & 31)) >> (v & 31)) & 1u;
						else seenb = vis[v];
					}
					uint32_t sc = 0, t;
					t = __vabsdiffu4(qcr.x, xc.x); sc = __dp4a(t, t, sc); t = __vabsdiffu4(qcr.y, xc.y); sc = __dp4a(t, t, sc);
					t = __vabsdiffu4(qcr.z, xc.z); sc = __dp4a(t, t, sc); t = __vabsdiffu4(qcr.w, xc.w); sc = __dp4a(t, t, sc);
#pragma unroll
					for (int o = 4; o >= 1; o >>= 1) sc += __shfl_xor_sync(FULL, sc, o);
					if (!bits && tl == 0 && v != NO_VERTEX) {
						seen = seenb == ep8;
						if (!seen) vis[v] = ep8;
					}
					const bool fresh = v != NO_VERTEX && !seen;
					const uint32_t od = f32_orderable((float)(int32_t)sc); // the reference converts the int32 score to T = float (:738-743)
					if (tl == 0 && in_b) {
						out[team] = fresh ? ((uint64_t)od << 32) | (v << 1) : KEY_EMPTY;
						if (fresh && od < job.w) pfv = v;
					}
				} else if (NSTEP > 0 && NR > 1 && one_round) {
					// rows of up to 32 * NR neighbours: team t scores neighbours t, t + 32, ...; the loads of all its rounds are
					// issued before the first is used
					const uint32_t rowlen = bottom ? m0s : ms;
					const uint8_t* pv = blk + (bottom ? off_b : off_u);
					uint32_t v[NR];
					float4 xa[NR][NSTEP > 0 ? NSTEP : 1];
#pragma unroll
					for (int r = 0; r < NR; ++r) {
						v[r] = NO_VERTEX;
						if (team + 32u * r < rowlen) {
							v[r] = __ldg(reinterpret_cast<const uint32_t*>(blk) + team + 32 * r);
#pragma unroll
							for (int s = 0; s < NSTEP; ++s) xa[r][s] = __ldg(reinterpret_cast<const float4*>(pv + (size_t)r * 32 * d_pad * 4 + s * 128));
						}
					}
					bool seen[NR];
					uint8_t seenb[NR];
#pragma unroll
					for (int r = 0; r < NR; ++r) { // visited test-and-set (:596-605) as soon as the id is here, under the vector loads
						seen[r] = false;
						seenb[r] = 0;
						if (tl == 0 && bottom && v[r] != NO_VERTEX) {
							if (bits) seen[r] = (atomicOr(bits + (v[r] >> 5), 1u << (v[r] & 31)) >> (v[r] & 31)) & 1u;
							else seenb[r] = vis[v[r]];
						}
					}
					float sa[NR];
#pragma unroll
					for (int r = 0; r < NR; ++r) {
						sa[r] = 0.f;
						if (team + 32u * r < rowlen) {
#pragma unroll
							for (int s = 0; s < NSTEP; ++s) { // same order of operations as gather_dist_f32
								float t;
								t = qr[s].x - xa[r][s].x; sa[r] = fmaf(t, t, sa[r]); t = qr[s].y - xa[r][s].y; sa[r] = fmaf(t, t, sa[r]);
								t = qr[s].z - xa[r][s].z; sa[r] = fmaf(t, t, sa[r]); t = qr[s].w - xa[r][s].w; sa[r] = fmaf(t, t, sa[r]);
							}
						}
					}
#pragma unroll
					for (int o = 4; o >= 1; o >>= 1) {
#pragma unroll
						for (int r = 0; r < NR; ++r) sa[r] += __shfl_xor_sync(FULL, sa[r], o);
					}
#pragma unroll
					for (int r = 0; r < NR; ++r) {
						const uint32_t pos = team + 32u * r;
						if (!bits && tl == 0 && bottom && v[r] != NO_VERTEX) {
							seen[r] = seenb[r] == ep8;
							if (!seen[r]) vis[v[r]] = ep8;
						}
						const bool fresh = v[r] != NO_VERTEX && !seen[r];
						const uint32_t od = f32_orderable(sa[r]);
						const uint64_t key = fresh ? ((uint64_t)od << 32) | (bottom ? v[r] << 1 : pos) : KEY_EMPTY; // bottom: beam_key(sa, v)
						if (tl == 0 && pos < rowlen) {
							out[pos] = key;
							if (bottom && fresh && od < job.w && pfv == NO_VERTEX) pfv = v[r];
						}
					}
				} else if (NSTEP > 0 && one_round) {
					const bool in = bottom ? in_b : in_u;
					const uint8_t* pv = blk + (bottom ? off_b : off_u);
					uint32_t v = NO_VERTEX;
					float4 xa[NSTEP > 0 ? NSTEP : 1];
					if (in) {
						v = __ldg(reinterpret_cast<const uint32_t*>(blk) + team);
#pragma unroll
						for (int s = 0; s < NSTEP; ++s) xa[s] = __ldg(reinterpret_cast<const float4*>(pv + s * 128));
					}
					// visited test-and-set (:596-605) as soon as the id is here, under the vector loads
					bool seen = false;
					uint8_t seenb = 0;
					if (tl == 0 && bottom && v != NO_VERTEX) {
						if (bits) seen = (atomicOr(bits + (v >> 5), 1u << (v & 31)) >> (v & 31)) & 1u;
						else seenb = vis[v];
					}
					GI_S(1);
					float sa = 0.f;
					if (in) {
#pragma unroll
						for (int s = 0; s < NSTEP; ++s) { // same order of operations as gather_dist_f32
							float t;
							t = qr[s].x - xa[s].x; sa = fmaf(t, t, sa); t = qr[s].y - xa[s].y; sa = fmaf(t, t, sa);
							t = qr[s].z - xa[s].z; sa = fmaf(t, t, sa); t = qr[s].w - xa[s].w; sa = fmaf(t, t, sa);
						}
					}
					GI_S(2);
#pragma unroll
					for (int o = 4; o >= 1; o >>= 1) sa += __shfl_xor_sync(FULL, sa, o);
					GI_S(3);
					{ // (branch-free: every lane forms the key, the team's first lane stores it)
						if (!bits && tl == 0 && bottom && v != NO_VERTEX) {
							seen = seenb == ep8;
							if (!seen) vis[v] = ep8;
						}
						const bool fresh = v != NO_VERTEX && !seen;
						const uint32_t od = f32_orderable(sa);
						const uint64_t key = fresh ? ((uint64_t)od << 32) | (bottom ? v << 1 : team) : KEY_EMPTY; // bottom: beam_key(sa, v)
						if (tl == 0 && in) {
							out[team] = key;
							if (bottom && fresh && od < job.w) pfv = v;
						}
					}
				} else {
					const uint32_t rowlen = bottom ? m0s : ms;
					const uint8_t* vecs = blk + (bottom ? p.bb_idb : p.ub_idb + ms * 16);
					for (uint32_t p0 = 0; p0 < rowlen; p0 += GI_SCORERS * 4) {
						const uint32_t pos = p0 + team;
						const bool in = pos < rowlen;
						const uint32_t v = in ? __ldg(reinterpret_cast<const uint32_t*>(blk) + pos) : NO_VERTEX;
						const float* pa = reinterpret_cast<const float*>(vecs) + (size_t)(in ? pos : 0) * d_pad + tl * 4;
						float sa = 0.f;
						if (COMP && bottom) { // 8-bit codes (rows of d_pad bytes)
							const uint8_t* pc = vecs + (size_t)(in ? pos : 0) * d_pad;
							uint32_t sc = 0;
							for (uint32_t c = tl; c < nchunk; c += 8) {
								const uint4 xc = __ldg(reinterpret_cast<const uint4*>(pc + c * 16));
								const uint4 q = *reinterpret_cast<const uint4*>(qcs + c * 16);
								uint32_t t;
								t = __vabsdiffu4(q.x, xc.x); sc = __dp4a(t, t, sc); t = __vabsdiffu4(q.y, xc.y); sc = __dp4a(t, t, sc);
								t = __vabsdiffu4(q.z, xc.z); sc = __dp4a(t, t, sc); t = __vabsdiffu4(q.w, xc.w); sc = __dp4a(t, t, sc);
							}
#pragma unroll
							for (int o = 4; o >= 1; o >>= 1) sc += __shfl_xor_sync(FULL, sc, o);
							sa = (float)(int32_t)sc;
						} else {
						for (int s0 = 0; s0 < nstep; s0 += 4) {
							float4 xa[4];
#pragma unroll
							for (int s = 0; s < 4; ++s)
								if (s0 + s < nstep) xa[s] = ldg128f(pa + (s0 + s) * 32);
#pragma unroll
							for (int s = 0; s < 4; ++s)
								if (s0 + s < nstep) { // same order of operations as gather_dist_f32
									const float4 q = *reinterpret_cast<const float4*>(qsm + (s0 + s) * 32 + tl * 4);
									float t;
									t = q.x - xa[s].x; sa = fmaf(t, t, sa); t = q.y - xa[s].y; sa = fmaf(t, t, sa);
									t = q.z - xa[s].z; sa = fmaf(t, t, sa); t = q.w - xa[s].w; sa = fmaf(t, t, sa);
								}
						}
#pragma unroll
						for (int o = 4; o >= 1; o >>= 1) sa += __shfl_xor_sync(FULL, sa, o);
						}
						if (tl == 0 && in) {
							uint64_t key = KEY_EMPTY;
							if (v != NO_VERTEX) {
								if (bottom) { // visited test-and-set (:596-605)
									bool seen;
									if (bits) seen = (atomicOr(bits + (v >> 5), 1u << (v & 31)) >> (v & 31)) & 1u;
									else {
										seen = vis[v] == ep8;
										if (!seen) vis[v] = ep8;
									}
									if (!seen) key = beam_key(sa, v);
								} else
									key = ((uint64_t)f32_orderable(sa) << 32) | pos;
							}
							out[pos] = key;
						}
					}
				}
				__syncwarp();
				if (lane == 0) GI_ARRIVE(&bar_sc[par]);
				GI_S(4);
				{ // one request per warp, with a warp-uniform address (per-lane operands make ptxas wrap the instruction in a loop)
					const unsigned pm = __ballot_sync(FULL, pfv != NO_VERTEX);
					if (pm) {
						const uint32_t v1 = __shfl_sync(FULL, pfv, __ffs(pm) - 1);
						if (lane == 0) bulk_prefetch_l2(p.blk_bottom + (size_t)v1 * p.bb_bytes, p.bb_bytes);
					}
				}
				GI_S(5);
#ifdef GI_PROFILE
				sp_jobs++;
#endif
				++h;
				if (bottom) {
					// ---- which vertex does the next hop expand? The smaller of the closest fresh neighbour (if the beam admits
					// it: then it is the smallest key admitted) and the closest unexpanded beam entry as of the last admission.
					GI_WAIT(&bar_sc[par], (h - 1) >> 1, (uint32_t)GI_SCORERS);
					GI_WAIT(&bar_adm[Jb & 1], Jb >> 1, 1u);
					const uint4 a = st[(Jb & 1) * 2], b = st[(Jb & 1) * 2 + 1];
					++Jb;
					const uint64_t r1 = ((uint64_t)a.y << 32) | a.x, r2 = ((uint64_t)a.w << 32) | a.z;
					uint64_t fk = (uint32_t)lane < m0s ? out[lane] : KEY_EMPTY;
					for (uint32_t i = 32 + lane; i < m0s; i += 32) fk = min(fk, out[i]);
					fk = warp_min_u64(fk);
					if ((uint32_t)(fk >> 32) >= b.x) fk = KEY_EMPTY; // not admitted (b.x = 0xFFFFFFFF while the beam is not full)
					const uint64_t nextk = min(fk, r1);
					if (nextk == KEY_EMPTY) { // the search is over
						// (compressed graphs: the controller reuses rk for its re-rank; it may do so once every scorer is here)
						if (COMP) asm volatile("bar.sync 1, %0;" ::"r"(GI_THREADS) : "memory");
						break;
					}
					const uint64_t second = nextk == fk ? r1 : min(fk, r2);
					const uint64_t addr = reinterpret_cast<uint64_t>(p.blk_bottom + (size_t)((uint32_t)(nextk >> 1) & 0x7FFFFFFFu) * p.bb_bytes);
					job = make_uint4((uint32_t)addr, (uint32_t)(addr >> 32), 2u, min((uint32_t)(second >> 32), b.x));
					self = true;
					GI_S(5);
				}
			}
#ifdef GI_PROFILE
			if (tid == 32 && (gi_prof_calls & 63u) == 40u)
				printf("GISCORE jobs=%d wait_go=%lld issue+visit=%lld fma=%lld reduce=%lld store+arrive=%lld prefetch=%lld\n", sp_jobs, sp[0], sp[1], sp[2], sp[3], sp[4], sp[5]);
#endif
		} else {
			// =============================== controller ===============================
			uint32_t dc = 0;
#ifdef GI_PROFILE
			long long pf[10] = { 0, 0, 0, 0, 0, 0, 0, 0, 0, 0 }, pf_t = clock64(), pf_t0 = pf_t;
			int pf_hops = 0, pf_up = 0;
#endif
			auto publish = [&](uint32_t mode, const uint8_t* blk, uint32_t bound = 0) { // job h for the scorers
				if (lane == 0) {
					const uint64_t a = reinterpret_cast<uint64_t>(blk);
					ctl[Jg & 1] = make_uint4((uint32_t)a, (uint32_t)(a >> 32), mode, bound);
					GI_ARRIVE(&bar_go[Jg & 1]);
				}
				++Jg;
			};
			// all lanes: the smallest key of job h's scores
			auto min_score = [&](uint32_t par, uint32_t rowlen) -> uint64_t {
				uint64_t k = (uint32_t)lane < rowlen ? rk[par * GS_MAXNB + lane] : KEY_EMPTY;
				for (uint32_t i = 32 + lane; i < rowlen; i += 32) k = min(k, rk[par * GS_MAXNB + i]);
				return warp_min_u64(k);
			};
			// ---- upper layers: greedy descent (:878-893), one block per step
			uint32_t ep = p.g.start, cu_ui = p.g.up_start_ui, cu_nl = p.g.up_start_nl;
			uint32_t layer = p.g.max_layer > 0 ? p.g.max_layer - 1 : 0;
			while (layer > 0 && cu_nl <= layer) --layer;
			if (layer > 0) publish(1, p.blk_upper + (size_t)(cu_ui + layer - 1) * p.ub_bytes);
			float ep_dist;
			{ // distance of the starting vertex (its vector is in L2 from the previous call on)
				float* tmpd = reinterpret_cast<float*>(ck);
				uint32_t* tmpi = reinterpret_cast<uint32_t*>(ck + 8);
				if (lane == 0) tmpi[0] = ep;
				__syncwarp();
				gather_dist_f32(p.g.base, d_pad, qsm, tmpi, tmpd, 1, lane);
				ep_dist = tmpd[0];
				dc++;
				__syncwarp();
			}
			while (layer > 0) {
				const uint8_t* blk = p.blk_upper + (size_t)(cu_ui + layer - 1) * p.ub_bytes;
				// the walk either moves to a neighbour (block unknown until this one is scored) or stays and drops a layer:
				// that block — the bottom block of the vertex at layer 1 — is asked into L2 now
				if (lane == 0) {
					if (layer >= 2) bulk_prefetch_l2(p.blk_upper + (size_t)(cu_ui + layer - 2) * p.ub_bytes, p.ub_bytes);
					else bulk_prefetch_l2(p.blk_bottom + (size_t)ep * p.bb_bytes, p.bb_bytes);
				}
				const uint4 mt = (uint32_t)lane < ms ? __ldg(reinterpret_cast<const uint4*>(blk + p.ub_idb) + lane) : make_uint4(0, 0, 0, 0); // (rows of up to 32)
				const uint32_t par = h & 1;
				GI_WAIT(&bar_sc[par], h >> 1, (uint32_t)GI_SCORERS);
				++h;
				// first occurrence of the minimum, only if strictly better (:879-889)
				uint32_t cnt = 0;
				for (uint32_t i0 = 0; i0 < ms; i0 += 32) cnt += __popc(__ballot_sync(FULL, i0 + lane < ms && rk[par * GS_MAXNB + i0 + lane] != KEY_EMPTY));
				dc += cnt;
				const uint64_t mk = min_score(par, ms);
				if ((uint32_t)(mk >> 32) < f32_orderable(ep_dist)) {
					const uint32_t besti = (uint32_t)mk;
					uint4 m;
					if (ms <= 32) {
						m.x = __shfl_sync(FULL, mt.x, besti); m.y = __shfl_sync(FULL, mt.y, besti); m.z = __shfl_sync(FULL, mt.z, besti);
					} else
						m = __ldg(reinterpret_cast<const uint4*>(blk + p.ub_idb) + besti);
					cu_ui = m.x;
					cu_nl = m.y;
					ep = m.z;
					ep_dist = f32_from_orderable((uint32_t)(mk >> 32));
				} else
					--layer;
				while (layer > 0 && cu_nl <= layer) --layer;
				if (layer > 0) publish(1, p.blk_upper + (size_t)(cu_ui + layer - 1) * p.ub_bytes);
#ifdef GI_PROFILE
				pf_up++;
#endif
			}
			GI_T(0);
			// ---- bottom layer (:495-708 / compressed :710-851). The entry vertex is the first to be expanded.
			if (COMP) { // its key in the beam is its score on the codes
				float* tmpd = reinterpret_cast<float*>(ck);
				uint32_t* tmpi = reinterpret_cast<uint32_t*>(ck + 8);
				if (lane == 0) tmpi[0] = ep;
				__syncwarp();
				gather_score_u8(p.g.codes, d_pad, qcs, tmpi, tmpd, 1, lane);
				ep_dist = tmpd[0];
				__syncwarp();
			}
			if (lane == 0) {
				if (bits) atomicOr(bits + (ep >> 5), 1u << (ep & 31));
				else vis[ep] = ep8;
				bm[0] = beam_key(ep_dist, ep) | 1ull;
			}
			__syncwarp();
			dc++; // the reference evaluates the entry vertex again on entering the bottom layer
			bool full = ef == 1;
			uint32_t thr_hi = full ? f32_orderable(ep_dist) : 0xFFFFFFFFu; // orderable distance at position ef - 1
			if (lane == 0) { // what the first decision starts from
				st[(Jb & 1) * 2] = make_uint4(0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu);
				st[(Jb & 1) * 2 + 1] = make_uint4(thr_hi, 0, 0, 0);
				GI_ARRIVE(&bar_adm[Jb & 1]);
			}
			publish(2, p.blk_bottom + (size_t)ep * p.bb_bytes, thr_hi);
			WarpList<LPL> beam;
			beam.init();
			beam.insert(beam_key(ep_dist, ep) | 1ull, lane);
			int bsz = 1;
			for (;;) {
				const uint32_t par = h & 1;
				GI_WAIT(&bar_sc[par], h >> 1, (uint32_t)GI_SCORERS);
				++h;
				GI_T(1);
				// ---- under the scorers' memory trip of the next hop (they have worked out its vertex themselves): admission
				// (:666-676) of this hop's whole row at once. Every fresh neighbour under the bound the hop started with is merged
				// into the sorted beam by rank counting — candidate c lands at #beam keys below c + #candidates below c, a beam key
				// moves up by #candidates below it (the round's candidates are staged in shared memory and read back as
				// broadcasts). Then the closest unexpanded entry — the vertex the scorers are on — is marked expanded, and the two
				// after it are what the next decision starts from.
				const uint64_t* keys = rk + par * GS_MAXNB;
				const uint64_t key0 = (uint32_t)lane < m0s ? keys[lane] : KEY_EMPTY;
				++Jb;
				GI_T(2);
#ifdef GI_PROFILE
				pf_hops++;
#endif
				const bool fullh = full;
				const uint32_t thrh = thr_hi;
				for (uint32_t i0 = 0; i0 < m0s; i0 += 32) {
					const uint32_t i = i0 + lane;
					uint64_t mykey = i0 == 0 ? key0 : (i < m0s ? keys[i] : KEY_EMPTY);
					const bool cand = mykey != KEY_EMPTY && (!fullh || (uint32_t)(mykey >> 32) < thrh);
					dc += __popc(__ballot_sync(FULL, mykey != KEY_EMPTY));
					const unsigned mk = __ballot_sync(FULL, cand);
					if (!mk) continue;
					const int ncand = __popc(mk);
					if (!cand) mykey = KEY_EMPTY;
					const int ci = cand ? __popc(mk & ((1u << lane) - 1u)) : -1; // my place among the round's candidates
					if (cand) ck[ci] = mykey;
					__syncwarp();
					int myrank = 0, lo = 0, shift[LPL];
#pragma unroll
					for (int s = 0; s < LPL; ++s) shift[s] = 0;
#pragma unroll 4
					for (int j = 0; j < ncand; ++j) {
						const uint64_t kk = ck[j];
						myrank += kk < mykey;
						int below = 0; // beam keys below candidate j (expanded bits do not decide a comparison with a fresh vertex)
#pragma unroll
						for (int s = 0; s < LPL; ++s) {
							below += __popc(__ballot_sync(FULL, beam.e[s] < kk));
							shift[s] += kk < beam.e[s];
						}
						if (j == ci) lo = below;
					}
#pragma unroll
					for (int s = 0; s < LPL; ++s) {
						const int np = s * 32 + lane + shift[s];
						if (beam.e[s] != KEY_EMPTY && np < 32 * LPL) bm[np] = beam.e[s];
					}
					if (cand && myrank + lo < 32 * LPL) bm[myrank + lo] = mykey;
					bsz = min(bsz + ncand, 32 * LPL);
					__syncwarp();
#pragma unroll
					for (int s = 0; s < LPL; ++s) beam.e[s] = s * 32 + lane < bsz ? bm[s * 32 + lane] : KEY_EMPTY;
					__syncwarp();
				}
				// (bm is the beam as of the last merge; a mark set since then sits on an entry that is not read here)
				full = bsz >= (int)ef;
				if (full) thr_hi = (uint32_t)(bm[ef - 1] >> 32);
				uint64_t r1, r2;
				{ // the three closest unexpanded entries among the first ef: the first is being expanded, the others are r1, r2
					int p0 = -1, p1 = -1, p2 = -1;
#pragma unroll
					for (int s = 0; s < LPL; ++s) {
						const int i = s * 32 + lane;
						unsigned mk = __ballot_sync(FULL, i < (int)ef && beam.e[s] != KEY_EMPTY && !(beam.e[s] & 1ull));
						if (p0 < 0 && mk) { p0 = s * 32 + __ffs(mk) - 1; mk &= mk - 1; }
						if (p0 >= 0 && p1 < 0 && mk) { p1 = s * 32 + __ffs(mk) - 1; mk &= mk - 1; }
						if (p1 >= 0 && p2 < 0 && mk) p2 = s * 32 + __ffs(mk) - 1;
					}
					if (p0 < 0) break; // nothing left to expand: the search is over (the scorers have seen the same)
#pragma unroll
					for (int s = 0; s < LPL; ++s)
						if (s * 32 + lane == p0) beam.e[s] |= 1ull;
					r1 = p1 >= 0 ? bm[p1] : KEY_EMPTY;
					r2 = p2 >= 0 ? bm[p2] : KEY_EMPTY;
					// the runner-up is the likely vertex after the one being scored: its block goes to L2
					if (r1 != KEY_EMPTY && lane == 0) bulk_prefetch_l2(p.blk_bottom + (size_t)((uint32_t)(r1 >> 1) & 0x7FFFFFFFu) * p.bb_bytes, p.bb_bytes);
				}
				if (lane == 0) { // for the decision after the next hop's scores
					st[(Jb & 1) * 2] = make_uint4((uint32_t)r1, (uint32_t)(r1 >> 32), (uint32_t)r2, (uint32_t)(r2 >> 32));
					st[(Jb & 1) * 2 + 1] = make_uint4(full ? thr_hi : 0xFFFFFFFFu, 0, 0, 0);
					GI_ARRIVE(&bar_adm[Jb & 1]);
				}
				GI_T(3);
			}
			// ---- output, as in the other forms
			uint64_t* dst = p.g.out_keys + (size_t)qi * p.g.k;
			uint32_t nres = 0;
			if (COMP) {
				asm volatile("bar.sync 1, %0;" ::"r"(GI_THREADS) : "memory"); // every scorer has read the last hop's scores
				// (:845-848) fp32 distances of the ef survivors; then — the deliberate deviation of the other forms, SURVEY 8(a)
				// a13 — re-sorted by the fp32 key before the truncation to k. The scorers are idle: rk is scratch. Sorted by rank
				// counting (every key against all others through shared memory), not by cnt insertions.
				uint32_t* sid = reinterpret_cast<uint32_t*>(rk);
				float* sd = reinterpret_cast<float*>(sid + GS_MAXNB);
				uint64_t* sorted = rk + GS_MAXNB;
				int cnt = 0;
#pragma unroll
				for (int s = 0; s < LPL; ++s) {
					const int i = s * 32 + lane;
					const bool ok = i < (int)ef && beam.e[s] != KEY_EMPTY;
					const unsigned mk = __ballot_sync(FULL, ok);
					if (ok) sid[cnt + __popc(mk & ((1u << lane) - 1))] = (uint32_t)(beam.e[s] >> 1) & 0x7FFFFFFFu;
					cnt += __popc(mk);
				}
				__syncwarp();
				for (int i = lane; i < cnt; i += 32) { // all survivors' rows are asked for at once: only the gather's first round waits for DRAM
					const char* r = reinterpret_cast<const char*>(p.g.base + (size_t)sid[i] * d_pad);
					for (uint32_t o = 0; o < d_pad * 4; o += 128) prefetch_l2(r + o);
				}
				gather_dist_f32(p.g.base, d_pad, qsm, sid, sd, cnt, lane);
				uint64_t mine[LPL];
#pragma unroll
				for (int s = 0; s < LPL; ++s) {
					const int i = s * 32 + lane;
					mine[s] = i < cnt ? make_key(sd[i], sid[i]) : KEY_EMPTY;
					if (i < cnt) bm[i] = mine[s];
				}
				__syncwarp();
				int rank[LPL];
#pragma unroll
				for (int s = 0; s < LPL; ++s) rank[s] = 0;
#pragma unroll 4
				for (int j = 0; j < cnt; ++j) {
					const uint64_t kk = bm[j];
#pragma unroll
					for (int s = 0; s < LPL; ++s) rank[s] += kk < mine[s];
				}
#pragma unroll
				for (int s = 0; s < LPL; ++s)
					if (mine[s] != KEY_EMPTY) sorted[rank[s]] = mine[s];
				__syncwarp();
#pragma unroll
				for (int s = 0; s < LPL; ++s) {
					const int i = s * 32 + lane;
					uint64_t key = KEY_EMPTY;
					if (i < (int)p.g.k) {
						key = i < cnt ? sorted[i] + p.g.id_add : KEY_EMPTY;
						dst[i] = key;
					}
					emit_decoded(p.g, qi, i, key, nres);
				}
			} else {
#pragma unroll
			for (int s = 0; s < LPL; ++s) {
				const int i = s * 32 + lane;
				uint64_t key = KEY_EMPTY;
				if (i < (int)p.g.k) {
					uint64_t e = beam.e[s];
					if (i >= (int)ef) e = KEY_EMPTY;
					key = e == KEY_EMPTY ? KEY_EMPTY : ((e >> 32) << 32) | (((uint32_t)(e >> 1) & 0x7FFFFFFFu) + p.g.id_add);
					dst[i] = key;
				}
				emit_decoded(p.g, qi, i, key, nres);
			}
			}
			if (p.g.dec_counts && lane == 0) p.g.dec_counts[qi] = nres;
			dc_total += dc;
#ifdef GI_PROFILE
			if (lane == 0 && (atomicAdd(&gi_prof_calls, 1u) & 63u) == 40u)
				printf("GIPROF total=%lld upper=%lld (steps %d) | hops=%d wait_scores=%lld decide=%lld admit+select=%lld dc=%u\n",
				       clock64() - pf_t0, pf[0], pf_up, pf_hops, pf[1], pf[2], pf[3], dc);
#endif
		}
		__syncthreads();
	}
	if (tid == 0 && !bits) p.g.epoch[slot] = epoch;
	if (tid == 0 && p.g.distcomps && dc_total) atomicAdd(p.g.distcomps, dc_total);
}

// ---------------------------------------------------------------------------------------------------------
// K5, batch form (one warp per query, graph_search_kernel) over the neighbourhood blocks. Same search, same results,
// same prefetches (the runner-up's ids a hop early, the rows of the gather's later rounds when the ids are known); what
// changes is where a hop's bytes come from: the ids and the vectors of the expanded vertex's neighbours are one contiguous
// block (rows addressed by position), 16.5 KB in ONE region of HBM instead of ~26 rows scattered over the base: +8-12 % QPS
// at every beam width (1M x 128: ef 40 5.77 -> 6.24 M QPS). MEASURED AND DROPPED: asking for whole blocks with
// cp.async.bulk.prefetch.L2 when a vertex is chosen and for the runner-up (4.27 M QPS: the visited neighbours' rows and the
// wrong guesses are paid for in DRAM bandwidth, which this form, unlike the one-query form, is short of); the first 4 KB of
// the block's vectors asked for with the ids (no change); scoring ALL rows of the block without waiting for the ids and the
// visited bytes (first eight rows in the ids' memory trip, the rest out of L2, distances of visited vertices thrown away:
// 4.44 M QPS — a fifth more bytes and FMAs cost more than the shorter chain saves).
// ---------------------------------------------------------------------------------------------------------
// gather_dist_f32 with the row numbers and the results in ONE array (io[i]: in = row of `rows`, out = distance bits)
__device__ __forceinline__ void gather_dist_f32_inplace(const float* __restrict__ rows, uint32_t d_pad, const float* qsm, uint32_t* io, int cnt, int lane) {
	const int team = lane >> 3, tl = lane & 7;
	const int nstep = d_pad >> 5;
	for (int i0 = 0; i0 < cnt; i0 += 8) {
		const int ia = i0 + team, ib = i0 + 4 + team;
		const uint32_t va = ia < cnt ? io[ia] : 0, vb = ib < cnt ? io[ib] : 0;
		const float* pa = rows + (size_t)va * d_pad + tl * 4;
		const float* pb = rows + (size_t)vb * d_pad + tl * 4;
		float sa = 0.f, sb = 0.f;
		for (int s0 = 0; s0 < nstep; s0 += 4) {
			float4 xa[4], xb[4];
#pragma unroll
			for (int s = 0; s < 4; ++s)
				if (s0 + s < nstep) { xa[s] = ldg128f(pa + (s0 + s) * 32); xb[s] = ldg128f(pb + (s0 + s) * 32); }
#pragma unroll
			for (int s = 0; s < 4; ++s)
				if (s0 + s < nstep) { // same order of operations as gather_dist_f32
					const float4 q = *reinterpret_cast<const float4*>(qsm + (s0 + s) * 32 + tl * 4);
					float t;
					t = q.x - xa[s].x; sa = fmaf(t, t, sa); t = q.y - xa[s].y; sa = fmaf(t, t, sa);
					t = q.z - xa[s].z; sa = fmaf(t, t, sa); t = q.w - xa[s].w; sa = fmaf(t, t, sa);
					t = q.x - xb[s].x; sb = fmaf(t, t, sb); t = q.y - xb[s].y; sb = fmaf(t, t, sb);
					t = q.z - xb[s].z; sb = fmaf(t, t, sb); t = q.w - xb[s].w; sb = fmaf(t, t, sb);
				}
		}
#pragma unroll
		for (int o = 4; o >= 1; o >>= 1) { sa += __shfl_xor_sync(FULL, sa, o); sb += __shfl_xor_sync(FULL, sb, o); }
		__syncwarp();
		if (tl == 0) {
			if (ia < cnt) io[ia] = __float_as_uint(sa);
			if (ib < cnt) io[ib] = __float_as_uint(sb);
		}
	}
	__syncwarp();
}

template <int LPL, int MINB>
__global__ void __launch_bounds__(32 * GS_WARPS, MINB) graph_search_blocks_kernel(const InlineParams ip) {
	extern __shared__ uint8_t gs_smem[];
	const GraphParams& p = ip.g;
	const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
	const uint32_t slot = blockIdx.x * GS_WARPS + wib, nslots = gridDim.x * GS_WARPS;
	// per-warp smem: query (d_pad floats) | nb ids (GS_MAXNB) | block rows in / distances out (GS_MAXNB)
	const size_t per_warp = (size_t)p.d_pad * 4 + GS_MAXNB * 8;
	uint8_t* wsm = gs_smem + per_warp * wib;
	float* qsm = reinterpret_cast<float*>(wsm);
	uint32_t* nb = reinterpret_cast<uint32_t*>(qsm + p.d_pad);
	uint32_t* nio = nb + GS_MAXNB;
	uint8_t* vis = p.visited + (size_t)slot * p.vis_stride;
	uint32_t epoch = p.epoch[slot];
	unsigned long long dc_total = 0;

	for (uint32_t qi = slot; qi < p.m; qi += nslots) {
		epoch++;
		if ((epoch & 0xFF) == 0) {
			epoch++;
			for (size_t i = (size_t)lane * 16; i < p.vis_stride; i += 32 * 16)
				*reinterpret_cast<uint4*>(vis + i) = make_uint4(0, 0, 0, 0);
			__syncwarp();
		}
		const uint8_t ep8 = (uint8_t)(epoch & 0xFF);
		for (uint32_t i = lane; i < p.d_pad; i += 32) qsm[i] = p.queries[(size_t)qi * p.d_pad + i];
		__syncwarp();
		uint32_t dc = 0;

		// ---- upper layers: greedy descent (src/antitopo_engine.h:861-903), as in graph_search_kernel
		uint32_t ep = p.start;
		float ep_dist;
		float* nbd = reinterpret_cast<float*>(nio);
		if (lane == 0) nb[0] = ep;
		__syncwarp();
		gather_dist_f32(p.base, p.d_pad, qsm, nb, nbd, 1, lane);
		ep_dist = nbd[0];
		dc++;
		for (uint32_t layer = p.max_layer > 0 ? p.max_layer - 1 : 0; layer > 0; --layer) {
			bool changed = true;
			while (changed) {
				changed = false;
				__syncwarp();
				const uint32_t ui = p.upper_idx[ep];
				if (ui == NO_VERTEX || p.upper_nl[ep] <= layer) break;
				const uint32_t* row = p.upper + ((size_t)ui + (layer - 1)) * p.ms;
				int cnt = 0;
				for (uint32_t j0 = 0; j0 < p.ms; j0 += 32) {
					const uint32_t v = j0 + lane < p.ms ? row[j0 + lane] : NO_VERTEX;
					const unsigned mk = __ballot_sync(FULL, v != NO_VERTEX);
					if (v != NO_VERTEX) nb[cnt + __popc(mk & ((1u << lane) - 1))] = v;
					cnt += __popc(mk);
				}
				__syncwarp();
				gather_dist_f32(p.base, p.d_pad, qsm, nb, nbd, cnt, lane);
				dc += cnt;
				float best = ep_dist;
				int besti = -1;
				for (int i = lane; i < cnt; i += 32)
					if (nbd[i] < best) { best = nbd[i]; besti = i; }
#pragma unroll
				for (int o = 16; o >= 1; o >>= 1) {
					const float ob = __shfl_xor_sync(FULL, best, o);
					const int oi = __shfl_xor_sync(FULL, besti, o);
					if (oi >= 0 && (ob < best || (ob == best && (besti < 0 || oi < besti)))) { best = ob; besti = oi; }
				}
				if (besti >= 0) { ep = nb[besti]; ep_dist = best; changed = true; }
			}
		}
		if (lane == 0) bulk_prefetch_l2(ip.blk_bottom + (size_t)ep * ip.bb_bytes, ip.bb_bytes);

		// ---- bottom layer beam search (src/antitopo_engine.h:495-708)
		WarpList<LPL> beam;
		beam.init();
		dc++;
		beam.insert(beam_key(ep_dist, ep), lane);
		if (lane == 0) vis[ep] = ep8;
		__syncwarp();
		float thr = __int_as_float(0x7f800000); // distance at position ef-1 (inf while the beam is not full)
		bool full = false;
		if (p.ef == 1) { full = true; thr = ep_dist; }
		for (;;) {
			// closest unexpanded entry among the first ef
			int pos = -1, pos2 = -1;
#pragma unroll
			for (int s = 0; s < LPL; ++s) {
				const int i = s * 32 + lane;
				unsigned mk = __ballot_sync(FULL, i < (int)p.ef && beam.e[s] != KEY_EMPTY && !(beam.e[s] & 1ull));
				if (pos < 0 && mk) { pos = s * 32 + __ffs(mk) - 1; mk &= mk - 1; }
				if (pos >= 0 && pos2 < 0 && mk) pos2 = s * 32 + __ffs(mk) - 1;
			}
			if (pos < 0) break;
			const uint64_t ck = beam.at(pos);
			const uint32_t cur = (uint32_t)(ck >> 1) & 0x7FFFFFFFu;
			const uint8_t* blk = ip.blk_bottom + (size_t)cur * ip.bb_bytes;
			const uint64_t rk2 = pos2 >= 0 ? beam.at(pos2) : KEY_EMPTY;
			// the runner-up is the likely next vertex: the line with its neighbours' ids goes to L2 a hop early
			if (rk2 != KEY_EMPTY && (uint32_t)lane * 32u < p.m0s) prefetch_l2(ip.blk_bottom + (size_t)((uint32_t)(rk2 >> 1) & 0x7FFFFFFFu) * ip.bb_bytes + lane * 128);
#pragma unroll
			for (int s = 0; s < LPL; ++s)
				if (s * 32 + lane == pos) beam.e[s] |= 1ull;
			// unvisited neighbours of cur (:595-607): ids from the block, fresh ones compacted with their row in the block
			int cnt = 0;
			__syncwarp();
			for (uint32_t j0 = 0; j0 < p.m0s; j0 += 32) {
				const uint32_t v = j0 + lane < p.m0s ? __ldg(reinterpret_cast<const uint32_t*>(blk) + j0 + lane) : NO_VERTEX;
				bool fresh = false;
				if (v != NO_VERTEX) {
					fresh = vis[v] != ep8;
					if (fresh) vis[v] = ep8;
				}
				const unsigned mk = __ballot_sync(FULL, fresh);
				if (fresh) {
					const int at = cnt + __popc(mk & ((1u << lane) - 1));
					nb[at] = v;
					nio[at] = j0 + lane;
				}
				cnt += __popc(mk);
			}
			__syncwarp();
			if (cnt == 0) continue;
			if (cnt > 8) { // rows of the gather's later rounds (8 per round) are asked for now: only the first round waits for DRAM
				for (int i = 8 + lane; i < cnt; i += 32) {
					const uint8_t* r = blk + ip.bb_idb + (size_t)nio[i] * p.d_pad * 4;
					for (uint32_t o = 0; o < p.d_pad * 4; o += 128) prefetch_l2(r + o);
				}
			}
			gather_dist_f32_inplace(reinterpret_cast<const float*>(blk + ip.bb_idb), p.d_pad, qsm, nio, cnt, lane);
			dc += cnt;
			// admission (:666-676): size < ef or d < worst
			for (int i0 = 0; i0 < cnt; i0 += 32) {
				const int i = i0 + lane;
				const float dv = i < cnt ? __uint_as_float(nio[i]) : 0.f;
				const uint32_t vv = i < cnt ? nb[i] : 0;
				unsigned mk = __ballot_sync(FULL, i < cnt && (!full || dv < thr));
				while (mk) {
					const int src = __ffs(mk) - 1;
					mk &= mk - 1;
					const float dd = __shfl_sync(FULL, dv, src);
					const uint32_t id = __shfl_sync(FULL, vv, src);
					if (!full || dd < thr) {
						beam.insert(beam_key(dd, id), lane);
						const uint64_t tk = beam.at(p.ef - 1);
						full = tk != KEY_EMPTY;
						if (full) thr = f32_from_orderable((uint32_t)(tk >> 32));
					}
				}
			}
		}
		// ---- output: first k of the ascending beam
		uint64_t* dst = p.out_keys + (size_t)qi * p.k;
#pragma unroll
		for (int s = 0; s < LPL; ++s) {
			const int i = s * 32 + lane;
			if (i < (int)p.k) {
				uint64_t e = beam.e[s];
				if (i >= (int)p.ef) e = KEY_EMPTY;
				dst[i] = e == KEY_EMPTY ? KEY_EMPTY : ((e >> 32) << 32) | (((uint32_t)(e >> 1) & 0x7FFFFFFFu) + p.id_add);
			}
		}
		dc_total += dc;
		__syncwarp();
	}
	if (lane == 0) {
		p.epoch[slot] = epoch;
		if (p.distcomps && dc_total) atomicAdd(p.distcomps, dc_total);
	}
}

// ---- building the blocks (once, at build()): one warp per (block, entry)
__global__ void build_bottom_blocks_kernel(const float* __restrict__ base, const uint32_t* __restrict__ bottom, uint8_t* __restrict__ blocks,
                                           uint32_t n, uint32_t m0s, uint32_t d_pad, uint32_t bb_bytes, uint32_t bb_idb) {
	const size_t w = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
	const uint32_t lane = threadIdx.x & 31;
	if (w >= (size_t)n * m0s) return;
	const uint32_t v = (uint32_t)(w / m0s), j = (uint32_t)(w % m0s);
	uint8_t* blk = blocks + (size_t)v * bb_bytes;
	const uint32_t nbv = bottom[(size_t)v * m0s + j];
	if (lane == 0) reinterpret_cast<uint32_t*>(blk)[j] = nbv;
	if (j == 0) // padding of the id section
		for (uint32_t i = m0s + lane; i < bb_idb / 4; i += 32) reinterpret_cast<uint32_t*>(blk)[i] = NO_VERTEX;
	float4* dst = reinterpret_cast<float4*>(blk + bb_idb + (size_t)j * d_pad * 4);
	const float4* src = reinterpret_cast<const float4*>(base + (size_t)(nbv == NO_VERTEX ? 0 : nbv) * d_pad);
	for (uint32_t i = lane; i < d_pad / 4; i += 32) dst[i] = nbv == NO_VERTEX ? make_float4(0.f, 0.f, 0.f, 0.f) : src[i];
}

// compressed graphs: block(v) = [ids | the neighbours' 8-bit codes, d_pad bytes each]
__global__ void build_bottom_code_blocks_kernel(const uint8_t* __restrict__ codes, const uint32_t* __restrict__ bottom, uint8_t* __restrict__ blocks,
                                                uint32_t n, uint32_t m0s, uint32_t d_pad, uint32_t bb_bytes, uint32_t bb_idb) {
	const size_t w = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
	const uint32_t lane = threadIdx.x & 31;
	if (w >= (size_t)n * m0s) return;
	const uint32_t v = (uint32_t)(w / m0s), j = (uint32_t)(w % m0s);
	uint8_t* blk = blocks + (size_t)v * bb_bytes;
	const uint32_t nbv = bottom[(size_t)v * m0s + j];
	if (lane == 0) reinterpret_cast<uint32_t*>(blk)[j] = nbv;
	if (j == 0) {
		for (uint32_t i = m0s + lane; i < bb_idb / 4; i += 32) reinterpret_cast<uint32_t*>(blk)[i] = NO_VERTEX;
		for (uint32_t i = bb_idb + m0s * d_pad + lane; i < bb_bytes; i += 32) blk[i] = 0; // tail padding
	}
	uint4* dst = reinterpret_cast<uint4*>(blk + bb_idb + (size_t)j * d_pad);
	const uint4* src = reinterpret_cast<const uint4*>(codes + (size_t)(nbv == NO_VERTEX ? 0 : nbv) * d_pad);
	for (uint32_t i = lane; i < d_pad / 16; i += 32) dst[i] = nbv == NO_VERTEX ? make_uint4(0, 0, 0, 0) : src[i];
}

__global__ void build_upper_blocks_kernel(const float* __restrict__ up_vec, const uint4* __restrict__ up_meta, const uint32_t* __restrict__ upper_c,
                                          uint8_t* __restrict__ blocks, uint32_t rows, uint32_t ms, uint32_t d_pad, uint32_t ub_bytes, uint32_t ub_idb) {
	const size_t w = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
	const uint32_t lane = threadIdx.x & 31;
	if (w >= (size_t)rows * ms) return;
	const uint32_t r = (uint32_t)(w / ms), j = (uint32_t)(w % ms);
	uint8_t* blk = blocks + (size_t)r * ub_bytes;
	const uint32_t c = upper_c[(size_t)r * ms + j];
	if (lane == 0) {
		reinterpret_cast<uint32_t*>(blk)[j] = c;
		reinterpret_cast<uint4*>(blk + ub_idb)[j] = c == NO_VERTEX ? make_uint4(0, 0, 0, 0) : up_meta[c];
	}
	if (j == 0)
		for (uint32_t i = ms + lane; i < ub_idb / 4; i += 32) reinterpret_cast<uint32_t*>(blk)[i] = NO_VERTEX;
	float4* dst = reinterpret_cast<float4*>(blk + ub_idb + (size_t)ms * 16 + (size_t)j * d_pad * 4);
	const float4* src = reinterpret_cast<const float4*>(up_vec + (size_t)(c == NO_VERTEX ? 0 : c) * d_pad);
	for (uint32_t i = lane; i < d_pad / 4; i += 32) dst[i] = c == NO_VERTEX ? make_float4(0.f, 0.f, 0.f, 0.f) : src[i];
}

} // namespace expann
