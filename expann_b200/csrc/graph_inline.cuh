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
//     meet on mbarriers, not bar.sync. As soon as the scores of hop h are in, the controller knows the vertex hop h+1
//     expands WITHOUT touching the beam: it is the smaller of the closest fresh neighbour just scored (if the beam admits
//     it) and the closest unexpanded beam entry — which was found at the end of the previous admission. It publishes that
//     vertex, the scorers start their loads, and only then does the controller merge hop h's candidates into the beam and
//     look up the next two unexpanded entries: that work runs under the scorers' memory trip.
//     The measured cost of a hop was instruction latency in the beam-owning warp, not memory (ncu: 233 k cycles per query,
//     0.4 long-scoreboard stalls per issue); see profiles/r2_notes.md.
#pragma once
#include "graph_search.cuh"

namespace expann {

constexpr int GI_SCORERS = 8;                   // scorer warps
constexpr int GI_THREADS = 32 * (GI_SCORERS + 1);

struct InlineParams {
	GraphParams g;
	const uint8_t* blk_bottom; // [n] blocks of bb_bytes: ids (m0s x u32, padded to bb_idb) | vectors (m0s x d_pad x f32)
	const uint8_t* blk_upper;  // [upper rows] blocks of ub_bytes: dense indices (ms x u32, padded to ub_idb) | meta (ms x uint4) | vectors
	uint32_t bb_bytes, bb_idb, ub_bytes, ub_idb;
	uint32_t bitmap_words; // visited bitmap in shared memory (0: epoch-stamped bytes in HBM)
	uint32_t prefetch;     // scorers ask the block of every neighbour that is under the beam's bound into L2 (calls with few queries)
};

__device__ __forceinline__ void bulk_prefetch_l2(const void* gsrc, uint32_t bytes) {
	asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(gsrc), "r"(bytes) : "memory");
}
__device__ __forceinline__ uint64_t warp_min_u64(uint64_t v) {
	const uint32_t hi = (uint32_t)(v >> 32), mh = __reduce_min_sync(FULL, hi);
	const uint32_t ml = __reduce_min_sync(FULL, hi == mh ? (uint32_t)v : 0xFFFFFFFFu);
	return ((uint64_t)mh << 32) | ml;
}

#ifdef GI_PROFILE
__device__ unsigned gi_prof_calls;
#define GI_T(slot) do { const long long t_ = clock64(); pf[slot] += t_ - pf_t; pf_t = t_; } while (0)
#else
#define GI_T(slot) do { } while (0)
#endif

template <int LPL>
__global__ void __launch_bounds__(GI_THREADS) graph_search_inline_kernel(const InlineParams p) {
	extern __shared__ __align__(128) uint8_t gi_smem[];
	const int tid = threadIdx.x, lane = tid & 31, wib = tid >> 5;
	const uint32_t slot = blockIdx.x;
	const uint32_t d_pad = p.g.d_pad, m0s = p.g.m0s, ms = p.g.ms, ef = p.g.ef;
	const bool qreg = d_pad <= 128;
	// shared memory: query | scores of a hop (ids, distances) x 2 | beam copy | staged candidates | ctl x 2 | barriers | bitmap
	float* qsm = reinterpret_cast<float*>(gi_smem);
	uint32_t* res_id = reinterpret_cast<uint32_t*>(qsm + d_pad); // [2][GS_MAXNB]
	float* res_d = reinterpret_cast<float*>(res_id + 2 * GS_MAXNB);
	uint64_t* bm = reinterpret_cast<uint64_t*>(res_d + 2 * GS_MAXNB); // [32 * LPL] shared-memory copy of the beam
	uint64_t* ck = bm + 256;                                          // [32] candidate keys of one admission round
	uint64_t* ctl = ck + 32;                                          // [2][2]: block pointer, mode (0 stop, 1 upper, 2 bottom)
	uint64_t* bar_go = ctl + 4;                                       // [2] controller -> scorers
	uint64_t* bar_sc = bar_go + 2;                                    // [2] scorers -> controller
	uint32_t* bits = p.bitmap_words ? reinterpret_cast<uint32_t*>(bar_sc + 2) : nullptr;
	uint8_t* vis = p.g.visited + (size_t)slot * p.g.vis_stride;
	uint32_t epoch = p.g.epoch[slot];
	unsigned long long dc_total = 0;
	uint32_t h = 0; // jobs handed to the scorers so far: job h uses buffers / barriers h & 1, phase (h >> 1) & 1

	if (tid == 0) {
		for (int b = 0; b < 2; ++b) { mbar_init(&bar_go[b], 1); mbar_init(&bar_sc[b], GI_SCORERS); }
		mbar_fence_init();
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
		for (uint32_t i = tid; i < d_pad; i += GI_THREADS) qsm[i] = p.g.queries[(size_t)qi * d_pad + i];
		__syncthreads();

		if (wib != 0) {
			// =============================== scorers ===============================
			const int team = (tid - 32) >> 3, tl = tid & 7;
			const int nstep = d_pad >> 5;
			float4 qr[4];
#pragma unroll
			for (int s = 0; s < 4; ++s)
				qr[s] = qreg && s < nstep ? *reinterpret_cast<const float4*>(qsm + s * 32 + tl * 4) : make_float4(0.f, 0.f, 0.f, 0.f);
			for (;;) {
				const uint32_t par = h & 1;
				mbar_wait(&bar_go[par], (h >> 1) & 1);
				const uint32_t mode = (uint32_t)ctl[par * 2 + 1], thr_pub = (uint32_t)(ctl[par * 2 + 1] >> 32);
				if (mode == 0) { // end of the query (the arrival keeps the two barriers' phases in step)
					if (lane == 0) mbar_arrive(&bar_sc[par]);
					++h;
					break;
				}
				const uint8_t* blk = reinterpret_cast<const uint8_t*>(ctl[par * 2]);
				const bool bottom = mode == 2;
				const uint32_t rowlen = bottom ? m0s : ms;
				const uint8_t* vecs = blk + (bottom ? p.bb_idb : p.ub_idb + ms * 16);
				uint32_t* rid = res_id + par * GS_MAXNB;
				float* rd = res_d + par * GS_MAXNB;
				for (uint32_t p0 = 0; p0 < rowlen; p0 += GI_SCORERS * 4) {
					const uint32_t pos = p0 + team;
					const bool in = pos < rowlen;
					const uint32_t v = in ? __ldg(reinterpret_cast<const uint32_t*>(blk) + pos) : NO_VERTEX;
					const float* pa = reinterpret_cast<const float*>(vecs) + (size_t)(in ? pos : 0) * d_pad + tl * 4;
					float sa = 0.f;
					if (qreg) {
						float4 xa[4];
#pragma unroll
						for (int s = 0; s < 4; ++s)
							if (s < nstep) xa[s] = ldg128f(pa + s * 32);
#pragma unroll
						for (int s = 0; s < 4; ++s)
							if (s < nstep) { // same order of operations as gather_dist_f32
								float t;
								t = qr[s].x - xa[s].x; sa = fmaf(t, t, sa); t = qr[s].y - xa[s].y; sa = fmaf(t, t, sa);
								t = qr[s].z - xa[s].z; sa = fmaf(t, t, sa); t = qr[s].w - xa[s].w; sa = fmaf(t, t, sa);
							}
					} else
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
					if (tl == 0 && in) {
						uint32_t out = v;
						if (bottom && v != NO_VERTEX) { // visited test-and-set (:596-605)
							bool seen;
							if (bits) seen = (atomicOr(bits + (v >> 5), 1u << (v & 31)) >> (v & 31)) & 1u;
							else {
								seen = vis[v] == ep8;
								if (!seen) vis[v] = ep8;
							}
							if (seen) out = NO_VERTEX;
							// a fresh neighbour under the bound enters the beam and may be expanded — at once if it is the closest:
							// its block starts its way into L2 now, before the controller has even seen the score
							else if (p.prefetch && f32_orderable(sa) < thr_pub) bulk_prefetch_l2(p.blk_bottom + (size_t)v * p.bb_bytes, p.bb_bytes);
						}
						rid[pos] = out;
						rd[pos] = sa;
					}
				}
				__syncwarp();
				if (lane == 0) mbar_arrive(&bar_sc[par]);
				++h;
			}
		} else {
			// =============================== controller ===============================
			uint32_t dc = 0;
#ifdef GI_PROFILE
			long long pf[10] = { 0, 0, 0, 0, 0, 0, 0, 0, 0, 0 }, pf_t = clock64(), pf_t0 = pf_t;
			int pf_hops = 0, pf_up = 0;
#endif
			auto publish = [&](uint32_t mode, const uint8_t* blk, uint32_t thr_pub = 0) { // job h for the scorers
				if (lane == 0) {
					ctl[(h & 1) * 2] = reinterpret_cast<uint64_t>(blk);
					ctl[(h & 1) * 2 + 1] = mode | ((uint64_t)thr_pub << 32);
					mbar_arrive(&bar_go[h & 1]);
				}
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
				const uint32_t par = h & 1;
				mbar_wait(&bar_sc[par], (h >> 1) & 1);
				++h;
				// first occurrence of the minimum, only if strictly better (:879-889)
				uint32_t bk = 0xFFFFFFFFu, bi = 0xFFFFFFFFu, cnt = 0;
				const uint32_t epk = f32_orderable(ep_dist);
				for (uint32_t i0 = 0; i0 < ms; i0 += 32) {
					const uint32_t i = i0 + lane;
					const bool ok = i < ms && res_id[par * GS_MAXNB + i] != NO_VERTEX;
					cnt += __popc(__ballot_sync(FULL, ok));
					const uint32_t k32 = ok ? f32_orderable(res_d[par * GS_MAXNB + i]) : 0xFFFFFFFFu;
					if (k32 < epk && k32 < bk) { bk = k32; bi = i; }
				}
				const uint32_t mn = __reduce_min_sync(FULL, bk);
				dc += cnt;
				if (mn != 0xFFFFFFFFu) {
					const uint32_t besti = __reduce_min_sync(FULL, bk == mn ? bi : 0xFFFFFFFFu);
					const uint4 mt = __ldg(reinterpret_cast<const uint4*>(blk + p.ub_idb) + besti);
					cu_ui = mt.x;
					cu_nl = mt.y;
					ep = mt.z;
					ep_dist = f32_from_orderable(mn);
				} else
					--layer;
				while (layer > 0 && cu_nl <= layer) --layer;
				if (layer > 0) publish(1, p.blk_upper + (size_t)(cu_ui + layer - 1) * p.ub_bytes);
#ifdef GI_PROFILE
				pf_up++;
#endif
			}
			GI_T(0);
			// ---- bottom layer (:495-708). The entry vertex is the first to be expanded.
			if (lane == 0) {
				if (bits) atomicOr(bits + (ep >> 5), 1u << (ep & 31));
				else vis[ep] = ep8;
				bm[0] = beam_key(ep_dist, ep) | 1ull;
			}
			__syncwarp();
			publish(2, p.blk_bottom + (size_t)ep * p.bb_bytes, ef == 1 ? f32_orderable(ep_dist) : 0xFFFFFFFFu);
			WarpList<LPL> beam;
			beam.init();
			dc++; // the reference evaluates the entry vertex again on entering the bottom layer
			beam.insert(beam_key(ep_dist, ep) | 1ull, lane);
			int bsz = 1;
			bool full = ef == 1;
			uint32_t thr_hi = full ? f32_orderable(ep_dist) : 0xFFFFFFFFu; // orderable distance at position ef - 1
			uint64_t r1 = KEY_EMPTY, r2 = KEY_EMPTY; // the two closest unexpanded entries among the first ef
			int pos1 = -1;
			__syncwarp();
			for (;;) {
				const uint32_t par = h & 1;
				mbar_wait(&bar_sc[par], (h >> 1) & 1);
				++h;
				GI_T(1);
				const uint32_t* rid = res_id + par * GS_MAXNB;
				const float* rd = res_d + par * GS_MAXNB;
				// ---- which vertex does the next hop expand? The smaller of the closest fresh neighbour (if the beam admits it:
				// then it is the smallest key admitted) and the closest unexpanded beam entry.
				const uint32_t vv0 = (uint32_t)lane < m0s ? rid[lane] : NO_VERTEX;
				const float dv0 = vv0 != NO_VERTEX ? rd[lane] : 0.f;
				uint64_t fk = vv0 != NO_VERTEX ? beam_key(dv0, vv0) : KEY_EMPTY;
				dc += __popc(__ballot_sync(FULL, vv0 != NO_VERTEX));
				for (uint32_t i0 = 32; i0 < m0s; i0 += 32) {
					const uint32_t i = i0 + lane;
					const uint32_t vv = i < m0s ? rid[i] : NO_VERTEX;
					if (vv != NO_VERTEX) fk = min(fk, beam_key(rd[i], vv));
					dc += __popc(__ballot_sync(FULL, vv != NO_VERTEX));
				}
				fk = warp_min_u64(fk);
				if (fk != KEY_EMPTY && full && (uint32_t)(fk >> 32) >= thr_hi) fk = KEY_EMPTY; // not admitted: then no neighbour is
				const uint64_t nextk = min(fk, r1);
				if (nextk == KEY_EMPTY) { // nothing left to expand (and nothing to admit): the search is over
					publish(0, nullptr);
					mbar_wait(&bar_sc[h & 1], (h >> 1) & 1);
					++h;
					break;
				}
				publish(2, p.blk_bottom + (size_t)((uint32_t)(nextk >> 1) & 0x7FFFFFFFu) * p.bb_bytes, thr_hi);
				{ // the runner-up of the next hop is the likely vertex after it: its block goes to L2
					const uint64_t second = nextk == fk ? r1 : min(fk, r2);
					if (second != KEY_EMPTY && lane == 0) bulk_prefetch_l2(p.blk_bottom + (size_t)((uint32_t)(second >> 1) & 0x7FFFFFFFu) * p.bb_bytes, p.bb_bytes);
				}
				GI_T(2);
#ifdef GI_PROFILE
				pf_hops++;
#endif
				// ---- under the scorers' memory trip: admission (:666-676) of the whole row at once. Every fresh neighbour under
				// the bound the hop started with is merged into the sorted beam by rank counting — candidate c lands at #beam keys
				// below c (binary search in the shared-memory copy of the beam) + #candidates below c, a beam key moves up by
				// #candidates below it (the round's candidates are staged in shared memory and read back as broadcasts). The
				// vertex just published enters (or stays in) the beam as expanded.
				if (nextk != fk) {
#pragma unroll
					for (int s = 0; s < LPL; ++s)
						if (s * 32 + lane == pos1) beam.e[s] |= 1ull;
				}
				const bool fullh = full;
				const uint32_t thrh = thr_hi;
				for (uint32_t i0 = 0; i0 < m0s; i0 += 32) {
					const uint32_t i = i0 + lane;
					const uint32_t vv = i0 == 0 ? vv0 : (i < m0s ? rid[i] : NO_VERTEX);
					const float dv = i0 == 0 ? dv0 : (vv != NO_VERTEX ? rd[i] : 0.f);
					const bool cand = vv != NO_VERTEX && (!fullh || f32_orderable(dv) < thrh);
					const unsigned mk = __ballot_sync(FULL, cand);
					if (!mk) continue;
					const int ncand = __popc(mk);
					uint64_t mykey = cand ? beam_key(dv, vv) : KEY_EMPTY;
					if (cand) ck[__popc(mk & ((1u << lane) - 1u))] = mykey;
					__syncwarp();
					int myrank = 0, shift[LPL];
#pragma unroll
					for (int s = 0; s < LPL; ++s) shift[s] = 0;
#pragma unroll 4
					for (int j = 0; j < ncand; ++j) {
						const uint64_t kk = ck[j];
						myrank += kk < mykey;
#pragma unroll
						for (int s = 0; s < LPL; ++s) shift[s] += kk < beam.e[s];
					}
					int lo = 0;
					if (cand) { // beam keys below mine (expanded bits do not decide a comparison with a fresh vertex)
						int hi = bsz;
						while (lo < hi) {
							const int mid = (lo + hi) >> 1;
							if (bm[mid] < mykey) lo = mid + 1;
							else hi = mid;
						}
						if (mykey == nextk) mykey |= 1ull;
					}
					__syncwarp();
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
				full = bsz >= (int)ef;
				if (full) thr_hi = (uint32_t)(beam.at(ef - 1) >> 32);
				{ // the two closest unexpanded entries among the first ef
					int p1 = -1, p2 = -1;
#pragma unroll
					for (int s = 0; s < LPL; ++s) {
						const int i = s * 32 + lane;
						unsigned mk = __ballot_sync(FULL, i < (int)ef && beam.e[s] != KEY_EMPTY && !(beam.e[s] & 1ull));
						if (p1 < 0 && mk) { p1 = s * 32 + __ffs(mk) - 1; mk &= mk - 1; }
						if (p1 >= 0 && p2 < 0 && mk) p2 = s * 32 + __ffs(mk) - 1;
					}
					pos1 = p1;
					r1 = p1 >= 0 ? beam.at(p1) : KEY_EMPTY;
					r2 = p2 >= 0 ? beam.at(p2) : KEY_EMPTY;
				}
				GI_T(3);
			}
			// ---- output, as in the other forms
			uint64_t* dst = p.g.out_keys + (size_t)qi * p.g.k;
			uint32_t nres = 0;
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
	if (tid == 0) {
		if (!bits) p.g.epoch[slot] = epoch;
		if (p.g.distcomps && dc_total) atomicAdd(p.g.distcomps, dc_total);
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
