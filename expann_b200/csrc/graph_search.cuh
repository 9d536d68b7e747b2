// K5 — graph beam search, one warp per query (sm_100a).
//
// Replaces antitopo_engine<T>::_query_k (reference src/antitopo_engine.h:853-928):
//   * upper layers max_layer-1 .. 1: greedy 1-NN descent from starting_vertex (:878-893; the "ortho" penalty is
//     inert on the query path because entry_points is empty, :871-875);
//   * bottom layer: best-first beam search of width ef, query_k_at_layer<true,false,false> (:495-708), or its
//     compressed form query_k_bottom_compressed (:710-851) scoring u8 codes with dist2_compressed (:38-61) and
//     re-ranking the ef survivors in fp32 (:845-848);
//   * result = first k of the ascending list (:914-919).
//
// Machine mapping
//   * beam = ONE sorted warp-resident list (WarpList) of (dist, id, expanded-bit) keys. The reference's two
//     heaps (candidates min-heap + nearest max-heap, :540-560) are equivalent to "expand the closest
//     unexpanded entry of the best-ef list, stop when there is none" (an evicted candidate has d >= worst and
//     triggers the reference's break at :588-590); admission is the reference's strict rule (:666).
//   * visited set (:208-209,574-577,596-605): exact, one byte per vertex per resident warp slot in HBM holding
//     the slot's query epoch — no clearing between queries (full clear every 255 queries).
//   * adjacency rows are fixed-stride uint32[n][M0s] (one coalesced read); neighbour vectors are gathered by
//     teams of 8 lanes (4 rows per warp step, 16-byte loads, 3-step shuffle reduce), U steps in flight.
#pragma once
#include "common.cuh"

namespace expann {

constexpr uint32_t NO_VERTEX = 0xFFFFFFFFu;
constexpr int GS_WARPS = 4;      // warps per CTA
constexpr int GS_MAXNB = 256;    // max neighbours of one vertex (bottom stride / upper stride)
constexpr uint32_t GS_HASH = 2048; // entries of a warp's visited hash in shared memory (8 KB; moves to HBM at 3/4 full)

struct GraphParams {
	const float* base;        // [n][d_pad]
	const uint8_t* codes;     // [n][d_pad] u8 (compressed mode) or null
	const uint32_t* bottom;   // [n][m0s], padded with NO_VERTEX
	const uint32_t* upper_idx; // [n] index of the vertex's first upper row, NO_VERTEX if it has no upper layers
	const uint32_t* upper_nl;  // [n] number of layers of the vertex (>= 1)
	const uint32_t* upper;    // [rows][ms]
	const float* queries;     // [m][d_pad]
	uint64_t* out_keys;       // [m][k]
	uint8_t* visited;         // [slots][n]
	uint32_t* epoch;          // [slots]
	unsigned long long* distcomps; // [1] accumulated RECORD_STATS counter
	uint32_t n, vis_stride, d_pad, m0s, ms, m, k, ef;
	uint32_t start, max_layer, id_add;
	int compressed;
	int width; // latency form: vertices expanded per hop (1 or 2)
	uint32_t hash; // warp form: entries of the per-warp visited hash in shared memory (power of two; 0 = visited bytes in HBM only)
	// latency form, upper layers in compact form (null = off): the vertices that have upper layers (~n/M of them) get a
	// dense index c; up_meta[c] = {first upper row, number of layers, global id, 0}; upper_c = the upper rows with
	// neighbours as dense indices; up_vec[c] = a copy of the vertex's vector. A step of the descent is then row ->
	// {vectors, meta} (two dependent loads, not three), all in ~n/M * (d*4 + M*4) bytes that stay L2-resident.
	const uint4* up_meta;
	const uint32_t* upper_c;
	const float* up_vec;
	uint32_t up_start_ui, up_start_nl;
	// latency form only: results also decoded in place (ids / distances / counts of expann_gpu_query), saving the
	// decode launch of a small host call; null otherwise
	uint64_t* dec_ids;
	float* dec_dists;
	uint64_t* dec_counts;
};

__device__ __forceinline__ float4 ldg128f(const float* p) { return __ldg(reinterpret_cast<const float4*>(p)); }
__device__ __forceinline__ void prefetch_l2(const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }

// fp32 squared distance of the warp's query (in smem) to `cnt` vertices listed in ids[] -> out[] (smem).
// Teams of 8 lanes: team t handles vertex i*4+t; lane j of the team loads float4 #(j + 8*s).
__device__ __forceinline__ void gather_dist_f32(const float* __restrict__ base, uint32_t d_pad, const float* qsm,
                                                const uint32_t* ids, float* out, int cnt, int lane) {
	const int team = lane >> 3, tl = lane & 7;
	const int nstep = d_pad >> 5; // float4 steps per team lane
	for (int i0 = 0; i0 < cnt; i0 += 8) {
		const int ia = i0 + team, ib = i0 + 4 + team;
		const uint32_t va = ia < cnt ? ids[ia] : NO_VERTEX, vb = ib < cnt ? ids[ib] : NO_VERTEX;
		const float* pa = base + (size_t)(va == NO_VERTEX ? 0 : va) * d_pad + tl * 4;
		const float* pb = base + (size_t)(vb == NO_VERTEX ? 0 : vb) * d_pad + tl * 4;
		float sa = 0.f, sb = 0.f;
		for (int s0 = 0; s0 < nstep; s0 += 4) {
			float4 xa[4], xb[4];
#pragma unroll
			for (int s = 0; s < 4; ++s)
				if (s0 + s < nstep) { xa[s] = ldg128f(pa + (s0 + s) * 32); xb[s] = ldg128f(pb + (s0 + s) * 32); }
#pragma unroll
			for (int s = 0; s < 4; ++s)
				if (s0 + s < nstep) {
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
		if (tl == 0) {
			if (ia < cnt) out[ia] = sa;
			if (ib < cnt) out[ib] = sb;
		}
	}
	__syncwarp();
}

// u8 scores: sum (q_i - code_i)^2 in int32, q_i = uint32(q[i]) packed as bytes in qc (smem, d_pad bytes).
// dist2_compressed (src/antitopo_engine.h:38-61); exact for q in [0,255] (the only domain it is meaningful on).
__device__ __forceinline__ void gather_score_u8(const uint8_t* __restrict__ codes, uint32_t d_pad, const uint8_t* qc,
                                                const uint32_t* ids, float* out, int cnt, int lane) {
	const int team = lane >> 3, tl = lane & 7;
	const int nchunk = d_pad >> 4; // 16-byte chunks per row
	for (int i0 = 0; i0 < cnt; i0 += 8) {
		const int ia = i0 + team, ib = i0 + 4 + team;
		const uint32_t va = ia < cnt ? ids[ia] : NO_VERTEX, vb = ib < cnt ? ids[ib] : NO_VERTEX;
		const uint8_t* pa = codes + (size_t)(va == NO_VERTEX ? 0 : va) * d_pad;
		const uint8_t* pb = codes + (size_t)(vb == NO_VERTEX ? 0 : vb) * d_pad;
		uint32_t sa = 0, sb = 0;
		for (int c = tl; c < nchunk; c += 8) {
			const uint4 xa = __ldg(reinterpret_cast<const uint4*>(pa + c * 16));
			const uint4 xb = __ldg(reinterpret_cast<const uint4*>(pb + c * 16));
			const uint4 q = *reinterpret_cast<const uint4*>(qc + c * 16);
			uint32_t t;
			t = __vabsdiffu4(q.x, xa.x); sa = __dp4a(t, t, sa); t = __vabsdiffu4(q.y, xa.y); sa = __dp4a(t, t, sa);
			t = __vabsdiffu4(q.z, xa.z); sa = __dp4a(t, t, sa); t = __vabsdiffu4(q.w, xa.w); sa = __dp4a(t, t, sa);
			t = __vabsdiffu4(q.x, xb.x); sb = __dp4a(t, t, sb); t = __vabsdiffu4(q.y, xb.y); sb = __dp4a(t, t, sb);
			t = __vabsdiffu4(q.z, xb.z); sb = __dp4a(t, t, sb); t = __vabsdiffu4(q.w, xb.w); sb = __dp4a(t, t, sb);
		}
#pragma unroll
		for (int o = 4; o >= 1; o >>= 1) { sa += __shfl_xor_sync(FULL, sa, o); sb += __shfl_xor_sync(FULL, sb, o); }
		if (tl == 0) { // the reference converts the int32 score to T=float (src/antitopo_engine.h:738-743)
			if (ia < cnt) out[ia] = (float)(int32_t)sa;
			if (ib < cnt) out[ib] = (float)(int32_t)sb;
		}
	}
	__syncwarp();
}

// beam entry: (orderable(dist) << 32) | (id << 1) | expanded
__device__ __forceinline__ uint64_t beam_key(float d, uint32_t id) { return ((uint64_t)f32_orderable(d) << 32) | ((uint64_t)id << 1); }

// MINB = resident CTAs per SM the register allocation is held to (65536 / (128 * MINB) registers per thread): the
// search is a chain of dependent DRAM round trips per hop, so resident warps, not registers, buy throughput
template <int LPL, int MINB>
__global__ void __launch_bounds__(32 * GS_WARPS, MINB) graph_search_kernel(const GraphParams p) {
	extern __shared__ uint8_t gs_smem[];
	const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
	const uint32_t slot = blockIdx.x * GS_WARPS + wib, nslots = gridDim.x * GS_WARPS;
	// per-warp smem: query (d_pad floats) | packed u8 query (d_pad bytes) | nb ids (GS_MAXNB) | nb dist (GS_MAXNB)
	const size_t per_warp = (size_t)p.d_pad * 4 + p.d_pad + GS_MAXNB * 8;
	uint8_t* wsm = gs_smem + per_warp * wib;
	float* qsm = reinterpret_cast<float*>(wsm);
	uint8_t* qc = wsm + (size_t)p.d_pad * 4;
	uint32_t* nb = reinterpret_cast<uint32_t*>(qc + p.d_pad);
	float* nbd = reinterpret_cast<float*>(nb + GS_MAXNB);
	uint8_t* vis = p.visited + (size_t)slot * p.vis_stride;
	uint32_t epoch = p.epoch[slot];
	unsigned long long dc_total = 0;
	// Visited set in SHARED memory (exact form): an open-addressing table of vertex ids per warp. A hop is
	// then adjacency row -> neighbour vectors, two dependent global round trips instead of three (row -> visited bytes ->
	// vectors): the search is latency-bound, the visited bytes were a third of every hop's chain. Exactness is kept: a query
	// that fills the table past 3/4 writes its entries into the HBM byte array and carries on there.
	// MEASURED (1M x 128, 10 000 queries, profiles/r2_graph_hash.json): identical ids and distance-evaluation counts, but
	// SLOWER than the bytes in HBM at every beam width (ef 20 / 40 / 60 / 120: 8.50 / 5.28 / 3.63 / 1.64 against 9.22 / 5.81 /
	// 4.23 / 2.10 M QPS): the bytes mostly hit L2 and their round trip overlaps across the 20 resident warps, while the
	// table costs an 8 KB clear per query and divergent compare-and-swap probing per hop. Kept behind option graph_hash = 1.
	uint32_t* htab = reinterpret_cast<uint32_t*>(gs_smem + per_warp * GS_WARPS) + (size_t)wib * p.hash;
	const uint32_t hmask = p.hash - 1, hshift = 32 - __popc(hmask);

	for (uint32_t qi = slot; qi < p.m; qi += nslots) {
		bool use_hash = p.hash != 0;
		uint32_t hcount = 0;
		if (use_hash) {
			for (uint32_t i = lane * 4; i < p.hash; i += 128) *reinterpret_cast<uint4*>(htab + i) = make_uint4(NO_VERTEX, NO_VERTEX, NO_VERTEX, NO_VERTEX);
			__syncwarp();
		}
		// true if v had not been seen by this query; marks it seen
		auto first_visit = [&](uint32_t v) -> bool {
			if (use_hash) {
				uint32_t hs = (v * 2654435761u) >> hshift;
				for (;;) {
					const uint32_t old = atomicCAS(htab + hs, NO_VERTEX, v);
					if (old == NO_VERTEX) return true;
					if (old == v) return false;
					hs = (hs + 1) & hmask;
				}
			}
			const bool fresh = vis[v] != (uint8_t)((epoch + 0u) & 0xFF);
			if (fresh) vis[v] = (uint8_t)(epoch & 0xFF);
			return fresh;
		};
		// ---- new epoch for the visited bytes (clear the slot when the 8-bit epoch wraps)
		epoch++;
		if ((epoch & 0xFF) == 0) {
			epoch++;
			for (size_t i = (size_t)lane * 16; i < p.vis_stride; i += 32 * 16)
				*reinterpret_cast<uint4*>(vis + i) = make_uint4(0, 0, 0, 0);
			__syncwarp();
		}
		const uint8_t ep8 = (uint8_t)(epoch & 0xFF);
		// ---- query -> smem (fp32 and the truncated u8 form of src/antitopo_engine.h:726-737)
		for (uint32_t i = lane; i < p.d_pad; i += 32) {
			const float v = p.queries[(size_t)qi * p.d_pad + i];
			qsm[i] = v;
			qc[i] = (uint8_t)(uint32_t)v;
		}
		__syncwarp();
		uint32_t dc = 0;

		// ---- upper layers: greedy descent (src/antitopo_engine.h:861-903)
		uint32_t ep = p.start;
		float ep_dist;
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
				// first occurrence of the minimum, only if strictly better (sequential semantics of :879-889)
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

		// ---- bottom layer beam search (src/antitopo_engine.h:495-708 / :710-851)
		WarpList<LPL> beam;
		beam.init();
		__syncwarp();
		if (lane == 0) nb[0] = ep;
		__syncwarp();
		if (p.compressed) gather_score_u8(p.codes, p.d_pad, qc, nb, nbd, 1, lane);
		else gather_dist_f32(p.base, p.d_pad, qsm, nb, nbd, 1, lane);
		dc++;
		beam.insert(beam_key(nbd[0], ep), lane);
		if (lane == 0) first_visit(ep);
		hcount = 1;
		__syncwarp();
		float thr = __int_as_float(0x7f800000); // distance at position ef-1 (inf while the beam is not full)
		bool full = false;
		if (p.ef == 1) { full = true; thr = nbd[0]; }
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
			if (pos2 >= 0) {
				// the runner-up is the likely next vertex to expand: pull its adjacency row into L2 now, one DRAM round
				// trip earlier than the hop that will need it
				const uint32_t nxt = (uint32_t)(beam.at(pos2) >> 1) & 0x7FFFFFFFu;
				if ((uint32_t)lane * 32u < p.m0s) prefetch_l2(p.bottom + (size_t)nxt * p.m0s + lane * 32);
			}
#pragma unroll
			for (int s = 0; s < LPL; ++s)
				if (s * 32 + lane == pos) beam.e[s] |= 1ull;
			const uint32_t cur = (uint32_t)(ck >> 1) & 0x7FFFFFFFu;
			// unvisited neighbours of cur (:595-607)
			const uint32_t* row = p.bottom + (size_t)cur * p.m0s;
			int cnt = 0;
			__syncwarp();
			for (uint32_t j0 = 0; j0 < p.m0s; j0 += 32) {
				const uint32_t v = j0 + lane < p.m0s ? __ldg(row + j0 + lane) : NO_VERTEX;
				bool fresh = false;
				if (v != NO_VERTEX) fresh = first_visit(v);
				const unsigned mk = __ballot_sync(FULL, fresh);
				if (fresh) nb[cnt + __popc(mk & ((1u << lane) - 1))] = v;
				cnt += __popc(mk);
			}
			__syncwarp();
			hcount += cnt;
			if (use_hash && hcount > (p.hash >> 2) * 3) {
				// the table is three quarters full: its entries move to the visited bytes in HBM, which take over for this query
				for (uint32_t i = lane; i < p.hash; i += 32) {
					const uint32_t v = htab[i];
					if (v != NO_VERTEX) vis[v] = ep8;
				}
				use_hash = false;
				__syncwarp();
			}
			if (cnt == 0) continue;
			if (cnt > 8) {
				// the gather below works through 8 vertices per round: ask for the rows of the later rounds now, so that only
				// the first round waits for DRAM
				const uint32_t row_bytes = p.compressed ? p.d_pad : p.d_pad * 4;
				const char* rows = p.compressed ? reinterpret_cast<const char*>(p.codes) : reinterpret_cast<const char*>(p.base);
				for (int i = 8 + lane; i < cnt; i += 32) {
					const char* r = rows + (size_t)nb[i] * row_bytes;
					for (uint32_t o = 0; o < row_bytes; o += 128) prefetch_l2(r + o);
				}
			}
			if (p.compressed) gather_score_u8(p.codes, p.d_pad, qc, nb, nbd, cnt, lane);
			else gather_dist_f32(p.base, p.d_pad, qsm, nb, nbd, cnt, lane);
			dc += cnt;
			// admission (:666-676): size < ef or d < worst
			for (int i0 = 0; i0 < cnt; i0 += 32) {
				const int i = i0 + lane;
				const float dv = i < cnt ? nbd[i] : 0.f;
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
		// ---- output: first k of the ascending beam; compressed: fp32 rerank of the ef survivors
		uint64_t* dst = p.out_keys + (size_t)qi * p.k;
		if (!p.compressed) {
#pragma unroll
			for (int s = 0; s < LPL; ++s) {
				const int i = s * 32 + lane;
				if (i < (int)p.k) {
					uint64_t e = beam.e[s];
					if (i >= (int)p.ef) e = KEY_EMPTY;
					dst[i] = e == KEY_EMPTY ? KEY_EMPTY : ((e >> 32) << 32) | (((uint32_t)(e >> 1) & 0x7FFFFFFFu) + p.id_add);
				}
			}
		} else {
			// (:845-848) d = dist2(all_entries[id], q) for the survivors; then — deliberate deviation, SURVEY §8(a)
			// a13 — re-sorted by the fp32 key before the truncation to k (the reference keeps compressed order).
			__syncwarp();
			int cnt = 0;
#pragma unroll
			for (int s = 0; s < LPL; ++s) {
				const int i = s * 32 + lane;
				const bool ok = i < (int)p.ef && beam.e[s] != KEY_EMPTY;
				const unsigned mk = __ballot_sync(FULL, ok);
				if (ok) nb[cnt + __popc(mk & ((1u << lane) - 1))] = (uint32_t)(beam.e[s] >> 1) & 0x7FFFFFFFu;
				cnt += __popc(mk);
			}
			__syncwarp();
			gather_dist_f32(p.base, p.d_pad, qsm, nb, nbd, cnt, lane);
			WarpList<LPL> fin;
			fin.init();
			for (int i = 0; i < cnt; ++i) fin.insert(make_key(nbd[i], nb[i]), lane);
#pragma unroll
			for (int s = 0; s < LPL; ++s) {
				const int i = s * 32 + lane;
				if (i < (int)p.k) dst[i] = fin.e[s] == KEY_EMPTY ? KEY_EMPTY : fin.e[s] + p.id_add;
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

// ---------------------------------------------------------------------------------------------------------
// K5, latency form: ONE CTA (GC_WARPS warps) per query, for calls with few queries (the reference harness's own
// loop is one query per call, src/basic_bench.h:82-126). Same search as graph_search_kernel — same visit order,
// same admission order, same per-vertex summation order, hence the same ids and distances bit for bit — but a
// hop costs two dependent memory round trips instead of three plus one per gather round:
//   * the 32 teams of 8 lanes score 32 neighbours of the expanded vertex at once (a whole row of M0 <= 32 in
//     one round; the warp form needs four);
//   * a team asks for its neighbour's vector and its visited byte together: the distance of an already visited
//     vertex is computed and thrown away (bandwidth is idle in this regime, latency is not);
//   * the beam stays in warp 0's registers; the other warps only score.
// Entries of a row keep their positions (no compaction): nb[pos] = NO_VERTEX marks "not a candidate".
// ---------------------------------------------------------------------------------------------------------
constexpr int GC_WARPS = 8;

template <bool VIS>
__device__ __forceinline__ void cta_score_row(const GraphParams& p, const uint32_t* __restrict__ row, uint32_t rowlen,
                                              bool compressed, const float* qsm, const uint8_t* qc, uint32_t* nb,
                                              float* nbd, uint8_t* vis, uint8_t ep8, const uint32_t* __restrict__ row2,
                                              uint32_t layer) {
	const int tid = threadIdx.x, team = tid >> 3, tl = tid & 7, lane = tid & 31;
	constexpr int NTEAM = GC_WARPS * 4;
	// speculation on the next hop: row2 is the adjacency row of the runner-up of the beam (in L2 since the hop that
	// found it, see below); its neighbours' vectors and visited bytes are asked into L2 while this hop is scored
	uint32_t v2 = NO_VERTEX;
	if (VIS && row2 && (uint32_t)tid < rowlen) v2 = __ldg(row2 + tid);
	for (uint32_t p0 = 0; p0 < rowlen; p0 += NTEAM) {
		const uint32_t pos = p0 + team;
		uint32_t v = NO_VERTEX;
		if (tl == 0 && pos < rowlen) v = __ldg(row + pos);
		v = __shfl_sync(FULL, v, lane & ~7);
		const bool have = v != NO_VERTEX;
		if (VIS) {
			// every neighbour may be expanded later: its own adjacency row goes to L2 now
			if (have && (uint32_t)tl * 32u < rowlen) prefetch_l2(p.bottom + (size_t)v * p.m0s + tl * 32);
			if (p0 == 0 && v2 != NO_VERTEX) {
				const uint32_t row_bytes = compressed ? p.d_pad : p.d_pad * 4;
				const char* r = (compressed ? reinterpret_cast<const char*>(p.codes) : reinterpret_cast<const char*>(p.base)) + (size_t)v2 * row_bytes;
				for (uint32_t o = 0; o < row_bytes; o += 128) prefetch_l2(r + o);
				prefetch_l2(vis + v2);
			}
		}
		uint8_t seen = ep8;
		if (VIS && have && tl == 0) seen = vis[v]; // in flight together with the vector loads below
		// upper layers: the walk moves to one of these neighbours or drops a layer at the same vertex; the index of
		// each neighbour's upper rows is fetched with its vector so that the rows themselves can be asked into L2
		uint32_t ui2 = NO_VERTEX;
		if (!VIS && have) {
			if (tl == 1) ui2 = __ldg(p.upper_idx + v);
			if (tl == 2) prefetch_l2(p.upper_nl + v);
			if (tl == 3 && layer == 1) prefetch_l2(p.bottom + (size_t)v * p.m0s);
		}
		float dist;
		if (compressed) {
			const uint8_t* pa = p.codes + (size_t)(have ? v : 0) * p.d_pad;
			const int nchunk = p.d_pad >> 4;
			uint32_t sa = 0;
			if (have)
				for (int c = tl; c < nchunk; c += 8) {
					const uint4 xa = __ldg(reinterpret_cast<const uint4*>(pa + c * 16));
					const uint4 q = *reinterpret_cast<const uint4*>(qc + c * 16);
					uint32_t t;
					t = __vabsdiffu4(q.x, xa.x); sa = __dp4a(t, t, sa); t = __vabsdiffu4(q.y, xa.y); sa = __dp4a(t, t, sa);
					t = __vabsdiffu4(q.z, xa.z); sa = __dp4a(t, t, sa); t = __vabsdiffu4(q.w, xa.w); sa = __dp4a(t, t, sa);
				}
#pragma unroll
			for (int o = 4; o >= 1; o >>= 1) sa += __shfl_xor_sync(FULL, sa, o);
			dist = (float)(int32_t)sa;
		} else {
			const float* pa = p.base + (size_t)(have ? v : 0) * p.d_pad + tl * 4;
			const int nstep = p.d_pad >> 5;
			float sa = 0.f;
			if (have)
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
			dist = sa;
		}
		if (!VIS && ui2 != NO_VERTEX) { // rows of this layer and the one below are adjacent
			const uint32_t lo = layer >= 2 ? layer - 2 : 0, bytes = (layer >= 2 ? 2 : 1) * p.ms * 4;
			const char* r = reinterpret_cast<const char*>(p.upper + ((size_t)ui2 + lo) * p.ms);
			for (uint32_t o = 0; o < bytes; o += 128) prefetch_l2(r + o);
		}
		if (tl == 0 && pos < rowlen) {
			if (VIS && have) {
				if (seen != ep8) vis[v] = ep8;
				else v = NO_VERTEX;
			}
			nb[pos] = v;
			nbd[pos] = dist;
		}
	}
	__syncthreads();
}

// Two vertices expanded per hop (GraphParams::width == 2): the closest two unexpanded beam entries are taken together,
// their rows (2 x M0 neighbours) scored in one pass with two vectors in flight per team — half as many dependent
// hops for a few more distance evaluations (the second vertex is expanded before the first one's results could
// have pushed it out of the beam). A neighbour both parents share is scored once (ids staged in shared memory and
// compared before the loads). Results: nb/nbd[j] for parent A's neighbour j, nb/nbd[M0s + j] for parent B's.
__device__ __forceinline__ void cta_score_pair(const GraphParams& p, uint32_t curA, uint32_t curB, bool compressed,
                                               const float* qsm, const uint8_t* qc, uint32_t* nb, float* nbd,
                                               uint32_t* idsrc, uint8_t* vis, uint8_t ep8,
                                               const uint32_t* __restrict__ row2) {
	const int tid = threadIdx.x, team = tid >> 3, tl = tid & 7, lane = tid & 31;
	const uint32_t m0s = p.m0s;
	uint32_t v2 = NO_VERTEX;
	if (row2 && (uint32_t)tid < m0s) v2 = __ldg(row2 + tid);
	if ((uint32_t)tid < 2 * m0s) {
		const bool second = (uint32_t)tid >= m0s;
		const uint32_t par = second ? curB : curA;
		idsrc[tid] = par == NO_VERTEX ? NO_VERTEX : __ldg(p.bottom + (size_t)par * m0s + (tid - (second ? m0s : 0)));
	}
	__syncthreads();
	const uint32_t row_bytes = compressed ? p.d_pad : p.d_pad * 4;
	const char* rows = compressed ? reinterpret_cast<const char*>(p.codes) : reinterpret_cast<const char*>(p.base);
	for (uint32_t j0 = 0; j0 < m0s; j0 += GC_WARPS * 4) {
		const uint32_t j = j0 + team;
		uint32_t va = j < m0s ? idsrc[j] : NO_VERTEX, vb = j < m0s ? idsrc[m0s + j] : NO_VERTEX;
		bool dupl = false;
		if (vb != NO_VERTEX)
			for (uint32_t i = tl; i < m0s; i += 8) dupl |= idsrc[i] == vb;
		if ((__ballot_sync(FULL, dupl) >> (lane & ~7)) & 0xFFu) vb = NO_VERTEX;
		const bool ha = va != NO_VERTEX, hb = vb != NO_VERTEX;
		if (ha && (uint32_t)tl * 32u < m0s) prefetch_l2(p.bottom + (size_t)va * m0s + tl * 32);
		if (hb && (uint32_t)tl * 32u < m0s) prefetch_l2(p.bottom + (size_t)vb * m0s + tl * 32);
		if (j0 == 0 && v2 != NO_VERTEX) {
			const char* r = rows + (size_t)v2 * row_bytes;
			for (uint32_t o = 0; o < row_bytes; o += 128) prefetch_l2(r + o);
			prefetch_l2(vis + v2);
		}
		uint8_t seenA = ep8, seenB = ep8;
		if (tl == 0) {
			if (ha) seenA = vis[va];
			if (hb) seenB = vis[vb];
		}
		float da, db;
		if (compressed) {
			const uint8_t* pa = p.codes + (size_t)(ha ? va : 0) * p.d_pad;
			const uint8_t* pb = p.codes + (size_t)(hb ? vb : 0) * p.d_pad;
			const int nchunk = p.d_pad >> 4;
			uint32_t sa = 0, sb = 0;
			for (int c = tl; c < nchunk; c += 8) {
				uint4 xa = make_uint4(0, 0, 0, 0), xb = make_uint4(0, 0, 0, 0);
				if (ha) xa = __ldg(reinterpret_cast<const uint4*>(pa + c * 16));
				if (hb) xb = __ldg(reinterpret_cast<const uint4*>(pb + c * 16));
				const uint4 q = *reinterpret_cast<const uint4*>(qc + c * 16);
				uint32_t t;
				t = __vabsdiffu4(q.x, xa.x); sa = __dp4a(t, t, sa); t = __vabsdiffu4(q.y, xa.y); sa = __dp4a(t, t, sa);
				t = __vabsdiffu4(q.z, xa.z); sa = __dp4a(t, t, sa); t = __vabsdiffu4(q.w, xa.w); sa = __dp4a(t, t, sa);
				t = __vabsdiffu4(q.x, xb.x); sb = __dp4a(t, t, sb); t = __vabsdiffu4(q.y, xb.y); sb = __dp4a(t, t, sb);
				t = __vabsdiffu4(q.z, xb.z); sb = __dp4a(t, t, sb); t = __vabsdiffu4(q.w, xb.w); sb = __dp4a(t, t, sb);
			}
#pragma unroll
			for (int o = 4; o >= 1; o >>= 1) { sa += __shfl_xor_sync(FULL, sa, o); sb += __shfl_xor_sync(FULL, sb, o); }
			da = (float)(int32_t)sa;
			db = (float)(int32_t)sb;
		} else {
			const float* pa = p.base + (size_t)(ha ? va : 0) * p.d_pad + tl * 4;
			const float* pb = p.base + (size_t)(hb ? vb : 0) * p.d_pad + tl * 4;
			const int nstep = p.d_pad >> 5;
			float sa = 0.f, sb = 0.f;
			for (int s0 = 0; s0 < nstep; s0 += 4) {
				float4 xa[4], xb[4];
#pragma unroll
				for (int s = 0; s < 4; ++s) {
					xa[s] = xb[s] = make_float4(0.f, 0.f, 0.f, 0.f);
					if (s0 + s < nstep) {
						if (ha) xa[s] = ldg128f(pa + (s0 + s) * 32);
						if (hb) xb[s] = ldg128f(pb + (s0 + s) * 32);
					}
				}
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
			da = sa;
			db = sb;
		}
		if (tl == 0 && j < m0s) {
			if (ha) { if (seenA != ep8) vis[va] = ep8; else va = NO_VERTEX; }
			if (hb) { if (seenB != ep8) vis[vb] = ep8; else vb = NO_VERTEX; }
			nb[j] = va;
			nbd[j] = da;
			nb[m0s + j] = vb;
			nbd[m0s + j] = db;
		}
	}
	__syncthreads();
}

// one step of the upper-layer descent on the compact form: scores the ms neighbours listed (as dense indices) in row,
// leaving nb[pos] = dense index, nbd[pos] = distance, mt[pos] = first upper row | layers << 32, gid[pos] = global id
__device__ __forceinline__ void cta_score_upper(const GraphParams& p, const uint32_t* __restrict__ row, uint32_t layer,
                                                const float* qsm, uint32_t* nb, float* nbd, uint64_t* mt, uint32_t* gid) {
	const int tid = threadIdx.x, team = tid >> 3, tl = tid & 7, lane = tid & 31;
	for (uint32_t p0 = 0; p0 < p.ms; p0 += GC_WARPS * 4) {
		const uint32_t pos = p0 + team;
		uint32_t c = NO_VERTEX;
		if (tl == 0 && pos < p.ms) c = __ldg(row + pos);
		c = __shfl_sync(FULL, c, lane & ~7);
		const bool have = c != NO_VERTEX;
		uint4 m = make_uint4(0, 0, 0, 0);
		if (have && tl == 1) m = __ldg(p.up_meta + c);
		const float* pa = p.up_vec + (size_t)(have ? c : 0) * p.d_pad + tl * 4;
		const int nstep = p.d_pad >> 5;
		float sa = 0.f;
		if (have)
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
		if (tl == 0 && pos < p.ms) { nb[pos] = c; nbd[pos] = sa; }
		if (tl == 1 && have) {
			mt[pos] = (uint64_t)m.x | ((uint64_t)m.y << 32);
			gid[pos] = m.z;
			if (layer == 1) prefetch_l2(p.bottom + (size_t)m.z * p.m0s);
		}
	}
	__syncthreads();
}

// decode_keys_kernel's work for result slot i of query qi (warp-wide call; key = KEY_EMPTY beyond k)
__device__ __forceinline__ void emit_decoded(const GraphParams& p, uint32_t qi, int i, uint64_t key, uint32_t& nres) {
	const bool ok = key != KEY_EMPTY;
	nres += __popc(__ballot_sync(FULL, ok));
	if (p.dec_ids && i < (int)p.k) {
		p.dec_ids[(size_t)qi * p.k + i] = ok ? (uint64_t)key_id(key) : 0xFFFFFFFFFFFFFFFFull;
		if (p.dec_dists) p.dec_dists[(size_t)qi * p.k + i] = ok ? key_dist(key) : __int_as_float(0x7f800000);
	}
}

template <int LPL>
__global__ void __launch_bounds__(32 * GC_WARPS) graph_search_cta_kernel(const GraphParams p) {
	extern __shared__ uint8_t gs_smem[];
	const int tid = threadIdx.x, lane = tid & 31, wib = tid >> 5;
	const uint32_t slot = blockIdx.x;
	float* qsm = reinterpret_cast<float*>(gs_smem);
	uint8_t* qc = gs_smem + (size_t)p.d_pad * 4;
	uint32_t* nb = reinterpret_cast<uint32_t*>(qc + p.d_pad);
	float* nbd = reinterpret_cast<float*>(nb + GS_MAXNB);
	uint32_t* ctl = reinterpret_cast<uint32_t*>(nbd + GS_MAXNB); // [0] vertex to expand (NO_VERTEX = search over), [1] runner-up, [2] its distance
	uint64_t* bm = reinterpret_cast<uint64_t*>(ctl + 4);         // [32 * LPL] shared-memory copy of the beam
	uint32_t* idsrc = reinterpret_cast<uint32_t*>(bm + 256);     // [GS_MAXNB] neighbour ids of the hop's two parents
	const uint32_t nrow = p.width == 2 ? 2 * p.m0s : p.m0s;      // result slots of a hop
	uint8_t* vis = p.visited + (size_t)slot * p.vis_stride;
	uint32_t epoch = p.epoch[slot];
	unsigned long long dc_total = 0;
	const bool comp = p.compressed != 0;

	for (uint32_t qi = slot; qi < p.m; qi += gridDim.x) {
		epoch++;
		if ((epoch & 0xFF) == 0) {
			epoch++;
			for (size_t i = (size_t)tid * 16; i < p.vis_stride; i += 32 * GC_WARPS * 16)
				*reinterpret_cast<uint4*>(vis + i) = make_uint4(0, 0, 0, 0);
		}
		const uint8_t ep8 = (uint8_t)(epoch & 0xFF);
		for (uint32_t i = tid; i < p.d_pad; i += 32 * GC_WARPS) {
			const float v = p.queries[(size_t)qi * p.d_pad + i];
			qsm[i] = v;
			qc[i] = (uint8_t)(uint32_t)v;
		}
		__syncthreads();
		uint32_t dc = 0;
#ifdef GC_PROFILE
		long long pf_t0 = clock64(), pf_sel = 0, pf_score = 0, pf_adm = 0, pf_up = 0, pf_t;
		int pf_hops = 0, pf_ins = 0, pf_upsteps = 0;
#endif

		// ---- upper layers: greedy descent, every warp tracks (ep, ep_dist) redundantly
		uint32_t ep = p.start;
		if (wib == 0) {
			if (lane == 0) nb[0] = ep;
			__syncwarp();
			gather_dist_f32(p.base, p.d_pad, qsm, nb, nbd, 1, lane);
		}
		__syncthreads();
		float ep_dist = nbd[0];
		uint32_t cu_ui = p.up_start_ui, cu_nl = p.up_start_nl; // compact form: the current vertex's upper rows
		dc++;
		__syncthreads();
		for (uint32_t layer = p.max_layer > 0 ? p.max_layer - 1 : 0; layer > 0; --layer) {
			bool changed = true;
			while (changed) {
				changed = false;
				if (p.up_meta) {
					if (cu_nl <= layer) break;
					cta_score_upper(p, p.upper_c + ((size_t)cu_ui + (layer - 1)) * p.ms, layer, qsm, nb, nbd, bm, idsrc);
				} else {
					const uint32_t ui = __ldg(p.upper_idx + ep), unl = __ldg(p.upper_nl + ep); // both in flight together
					if (ui == NO_VERTEX || unl <= layer) break;
					const uint32_t* row = p.upper + ((size_t)ui + (layer - 1)) * p.ms;
					cta_score_row<false>(p, row, p.ms, false, qsm, qc, nb, nbd, vis, ep8, nullptr, layer);
				}
#ifdef GC_PROFILE
				pf_upsteps++;
#endif
				float best = ep_dist;
				int besti = -1, cnt = 0;
				for (uint32_t i0 = 0; i0 < p.ms; i0 += 32) {
					const uint32_t i = i0 + lane;
					const bool ok = i < p.ms && nb[i] != NO_VERTEX;
					cnt += __popc(__ballot_sync(FULL, ok));
					if (ok && nbd[i] < best) { best = nbd[i]; besti = (int)i; }
				}
#pragma unroll
				for (int o = 16; o >= 1; o >>= 1) {
					const float ob = __shfl_xor_sync(FULL, best, o);
					const int oi = __shfl_xor_sync(FULL, besti, o);
					if (oi >= 0 && (ob < best || (ob == best && (besti < 0 || oi < besti)))) { best = ob; besti = oi; }
				}
				dc += cnt;
				if (besti >= 0) {
					if (p.up_meta) {
						const uint64_t mt = bm[besti];
						cu_ui = (uint32_t)mt;
						cu_nl = (uint32_t)(mt >> 32);
						ep = idsrc[besti];
					} else
						ep = nb[besti];
					ep_dist = best;
					changed = true;
				}
				__syncthreads();
			}
		}

#ifdef GC_PROFILE
		pf_up = clock64() - pf_t0;
		pf_t = clock64();
#endif
		// ---- bottom layer: warp 0 owns the beam
		WarpList<LPL> beam;
		float thr = __int_as_float(0x7f800000);
		bool full = false;
		int bsz = 0; // entries in the beam
		if (wib == 0) {
			beam.init();
			if (lane == 0) nb[0] = ep;
			__syncwarp();
			if (comp) gather_score_u8(p.codes, p.d_pad, qc, nb, nbd, 1, lane);
			else gather_dist_f32(p.base, p.d_pad, qsm, nb, nbd, 1, lane);
			dc++;
			beam.insert(beam_key(nbd[0], ep), lane);
			if (lane == 0) { vis[ep] = ep8; bm[0] = beam_key(nbd[0], ep); }
			bsz = 1;
			if (p.ef == 1) { full = true; thr = nbd[0]; }
			__syncwarp();
		}
		for (;;) {
			if (wib == 0) {
				int pos = -1, pos2 = -1, pos3 = -1;
#pragma unroll
				for (int s = 0; s < LPL; ++s) {
					const int i = s * 32 + lane;
					unsigned mk = __ballot_sync(FULL, i < (int)p.ef && beam.e[s] != KEY_EMPTY && !(beam.e[s] & 1ull));
					if (pos < 0 && mk) { pos = s * 32 + __ffs(mk) - 1; mk &= mk - 1; }
					if (pos >= 0 && pos2 < 0 && mk) { pos2 = s * 32 + __ffs(mk) - 1; mk &= mk - 1; }
					if (pos2 >= 0 && pos3 < 0 && mk) pos3 = s * 32 + __ffs(mk) - 1;
				}
				uint32_t cur = NO_VERTEX, curB = NO_VERTEX, nxt = NO_VERTEX;
				float nxt_d = 0.f;
				if (pos >= 0) { // bm = the beam as of the last merge (expanded bits apart)
					cur = (uint32_t)(bm[pos] >> 1) & 0x7FFFFFFFu;
					int posn = pos2; // runner-up: the closest entry not expanded in this hop
					if (p.width == 2 && pos2 >= 0) {
						curB = (uint32_t)(bm[pos2] >> 1) & 0x7FFFFFFFu;
						posn = pos3;
					} else
						pos2 = -1;
					if (posn >= 0) {
						const uint64_t k2 = bm[posn];
						nxt = (uint32_t)(k2 >> 1) & 0x7FFFFFFFu;
						nxt_d = f32_from_orderable((uint32_t)(k2 >> 32));
					}
#pragma unroll
					for (int s = 0; s < LPL; ++s)
						if (s * 32 + lane == pos || s * 32 + lane == pos2) beam.e[s] |= 1ull;
				}
				if (lane == 0) { ctl[0] = cur; ctl[1] = nxt; ctl[2] = __float_as_uint(nxt_d); ctl[3] = curB; }
			}
			__syncthreads();
#ifdef GC_PROFILE
			{ long long t = clock64(); pf_sel += t - pf_t; pf_t = t; pf_hops++; }
#endif
			const uint32_t cur = ctl[0], nxt = ctl[1];
			const float nxt_d = __uint_as_float(ctl[2]);
			if (cur == NO_VERTEX) break;
			if (p.width == 2)
				cta_score_pair(p, cur, ctl[3], comp, qsm, qc, nb, nbd, idsrc, vis, ep8,
				               nxt == NO_VERTEX ? nullptr : p.bottom + (size_t)nxt * p.m0s);
			else
				cta_score_row<true>(p, p.bottom + (size_t)cur * p.m0s, p.m0s, comp, qsm, qc, nb, nbd, vis, ep8,
				                    nxt == NO_VERTEX ? nullptr : p.bottom + (size_t)nxt * p.m0s, 0);
#ifdef GC_PROFILE
			{ long long t = clock64(); pf_score += t - pf_t; pf_t = t; }
#endif
			if (wib == 0) {
				// admission (:666-676) of a whole row at once: every fresh neighbour under the bound the hop started with
				// is merged into the sorted beam by rank counting (candidate c lands at #beam keys below c + #candidates
				// below c, a beam key moves up by #candidates below it) through the shared-memory copy of the beam.
				// The one-at-a-time rule of the warp form differs only when a candidate's distance EQUALS the bound it
				// would meet after earlier insertions of the same row (it is then turned away instead of ordered by id).
				for (uint32_t i0 = 0; i0 < nrow; i0 += 32) {
					const uint32_t i = i0 + lane;
					const uint32_t vv = i < nrow ? nb[i] : NO_VERTEX;
					const float dv = vv != NO_VERTEX ? nbd[i] : 0.f;
					dc += __popc(__ballot_sync(FULL, vv != NO_VERTEX));
					const bool cand = vv != NO_VERTEX && (!full || dv < thr);
					const unsigned mk = __ballot_sync(FULL, cand);
					if (!mk) continue;
					const uint64_t mykey = cand ? beam_key(dv, vv) : KEY_EMPTY;
					int myrank = 0, mybelow = 0, shift[LPL];
#pragma unroll
					for (int s = 0; s < LPL; ++s) shift[s] = 0;
					for (unsigned m2 = mk; m2; m2 &= m2 - 1) {
						const int src = __ffs(m2) - 1;
						const uint64_t kk = shfl_u64(mykey, src);
						myrank += kk < mykey;
#pragma unroll
						for (int s = 0; s < LPL; ++s) {
							const unsigned bl = __ballot_sync(FULL, beam.e[s] < kk);
							if (lane == src) mybelow += __popc(bl);
							shift[s] += kk < beam.e[s];
						}
					}
#pragma unroll
					for (int s = 0; s < LPL; ++s) {
						const int np = s * 32 + lane + shift[s];
						if (beam.e[s] != KEY_EMPTY && np < 32 * LPL) bm[np] = beam.e[s];
					}
					if (cand && myrank + mybelow < 32 * LPL) bm[myrank + mybelow] = mykey;
					bsz = min(bsz + __popc(mk), 32 * LPL);
					__syncwarp();
#pragma unroll
					for (int s = 0; s < LPL; ++s) beam.e[s] = s * 32 + lane < bsz ? bm[s * 32 + lane] : KEY_EMPTY;
					full = bsz >= (int)p.ef;
					if (full) thr = f32_from_orderable((uint32_t)(bm[p.ef - 1] >> 32));
					__syncwarp();
#ifdef GC_PROFILE
					pf_ins += __popc(mk);
#endif
				}
			} else if (wib == 1) {
				// the vertex the next hop expands is the closer of the runner-up (whose neighbourhood was asked into L2
				// while this hop was scored) and the closest fresh neighbour just scored: in the second case ask for ITS
				// neighbours' vectors and visited bytes now, while warp 0 is busy with the admission
				float best = __int_as_float(0x7f800000);
				uint32_t bv = NO_VERTEX;
				for (uint32_t i0 = 0; i0 < nrow; i0 += 32) {
					const uint32_t i = i0 + lane;
					const uint32_t vv = i < nrow ? nb[i] : NO_VERTEX;
					if (vv != NO_VERTEX && nbd[i] < best) { best = nbd[i]; bv = vv; }
				}
#pragma unroll
				for (int o = 16; o >= 1; o >>= 1) {
					const float ob = __shfl_xor_sync(FULL, best, o);
					const uint32_t ov = __shfl_xor_sync(FULL, bv, o);
					if (ob < best) { best = ob; bv = ov; }
				}
				if (bv != NO_VERTEX && (nxt == NO_VERTEX || best < nxt_d)) {
					const uint32_t* r2 = p.bottom + (size_t)bv * p.m0s;
					const uint32_t row_bytes = comp ? p.d_pad : p.d_pad * 4;
					const char* rows = comp ? reinterpret_cast<const char*>(p.codes) : reinterpret_cast<const char*>(p.base);
					for (uint32_t t = lane; t < p.m0s; t += 32) {
						const uint32_t v2 = __ldg(r2 + t);
						if (v2 != NO_VERTEX) {
							const char* r = rows + (size_t)v2 * row_bytes;
							for (uint32_t o = 0; o < row_bytes; o += 128) prefetch_l2(r + o);
							prefetch_l2(vis + v2);
						}
					}
				}
			}
#ifdef GC_PROFILE
			{ long long t = clock64(); pf_adm += t - pf_t; pf_t = t; }
#endif
			// warps 0 and 1 reach the next rendezvous only after they have read nb/nbd; nobody writes them before it
		}
		// ---- output (warp 0), as in the warp form
		if (wib == 0) {
			uint64_t* dst = p.out_keys + (size_t)qi * p.k;
			uint32_t nres = 0;
			if (!comp) {
#pragma unroll
				for (int s = 0; s < LPL; ++s) {
					const int i = s * 32 + lane;
					uint64_t key = KEY_EMPTY;
					if (i < (int)p.k) {
						uint64_t e = beam.e[s];
						if (i >= (int)p.ef) e = KEY_EMPTY;
						key = e == KEY_EMPTY ? KEY_EMPTY : ((e >> 32) << 32) | (((uint32_t)(e >> 1) & 0x7FFFFFFFu) + p.id_add);
						dst[i] = key;
					}
					emit_decoded(p, qi, i, key, nres);
				}
			} else {
				__syncwarp();
				int cnt = 0;
#pragma unroll
				for (int s = 0; s < LPL; ++s) {
					const int i = s * 32 + lane;
					const bool ok = i < (int)p.ef && beam.e[s] != KEY_EMPTY;
					const unsigned mk = __ballot_sync(FULL, ok);
					if (ok) nb[cnt + __popc(mk & ((1u << lane) - 1))] = (uint32_t)(beam.e[s] >> 1) & 0x7FFFFFFFu;
					cnt += __popc(mk);
				}
				__syncwarp();
				gather_dist_f32(p.base, p.d_pad, qsm, nb, nbd, cnt, lane);
				WarpList<LPL> fin;
				fin.init();
				for (int i = 0; i < cnt; ++i) fin.insert(make_key(nbd[i], nb[i]), lane);
#pragma unroll
				for (int s = 0; s < LPL; ++s) {
					const int i = s * 32 + lane;
					uint64_t key = KEY_EMPTY;
					if (i < (int)p.k) {
						key = fin.e[s] == KEY_EMPTY ? KEY_EMPTY : fin.e[s] + p.id_add;
						dst[i] = key;
					}
					emit_decoded(p, qi, i, key, nres);
				}
			}
			if (p.dec_counts && lane == 0) p.dec_counts[qi] = nres;
			dc_total += dc;
#ifdef GC_PROFILE
			if (lane == 0 && (epoch & 63) == 1)
				printf("GCPROF q=%u total=%lld upper=%lld (steps %d) select+sync=%lld score=%lld admit=%lld hops=%d inserts=%d dc=%u\n", qi,
				       clock64() - pf_t0, pf_up, pf_upsteps, pf_sel, pf_score, pf_adm, pf_hops, pf_ins, dc);
#endif
		}
		__syncthreads();
	}
	if (tid == 0) {
		p.epoch[slot] = epoch;
		if (p.distcomps && dc_total) atomicAdd(p.distcomps, dc_total);
	}
}

} // namespace expann
