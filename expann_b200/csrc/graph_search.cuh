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

	for (uint32_t qi = slot; qi < p.m; qi += nslots) {
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
		if (lane == 0) vis[ep] = ep8;
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
				if (v != NO_VERTEX) {
					fresh = vis[v] != ep8;
					if (fresh) vis[v] = ep8;
				}
				const unsigned mk = __ballot_sync(FULL, fresh);
				if (fresh) nb[cnt + __popc(mk & ((1u << lane) - 1))] = v;
				cnt += __popc(mk);
			}
			__syncwarp();
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

} // namespace expann
