// K4 — exact fp32 re-rank of candidate ids (one warp per query), then the final (key asc, id asc) order.
// Reference: the "final rerank" of query_k_bottom_compressed, src/antitopo_engine.h:845-848
// (d = dist2(all_entries[id], q) for every survivor). Deliberate deviation: the survivors are
// re-sorted by the fp32 key before truncation to k; the reference keeps the compressed order (:849).
// Also: quantizers (src/quantizer.h) and explicit-pair distance primitives for the parity tests.
#pragma once
#include "common.cuh"
#include "graph_search.cuh"

namespace expann {

// -(q . x) for `cnt` vertices (teams of 8 lanes, as gather_dist_f32)
__device__ __forceinline__ void gather_negdot_f32(const float* __restrict__ base, uint32_t d_pad, const float* qsm,
                                                  const uint32_t* ids, float* out, int cnt, int lane) {
	const int team = lane >> 3, tl = lane & 7;
	const int nstep = d_pad >> 5;
	for (int i0 = 0; i0 < cnt; i0 += 4) {
		const int ia = i0 + team;
		const uint32_t va = ia < cnt ? ids[ia] : NO_VERTEX;
		const float* pa = base + (size_t)(va == NO_VERTEX ? 0 : va) * d_pad + tl * 4;
		float sa = 0.f;
		for (int s = 0; s < nstep; ++s) {
			const float4 x = ldg128f(pa + s * 32);
			const float4 q = *reinterpret_cast<const float4*>(qsm + s * 32 + tl * 4);
			sa = fmaf(q.x, x.x, sa); sa = fmaf(q.y, x.y, sa); sa = fmaf(q.z, x.z, sa); sa = fmaf(q.w, x.w, sa);
		}
#pragma unroll
		for (int o = 4; o >= 1; o >>= 1) sa += __shfl_xor_sync(FULL, sa, o);
		if (tl == 0 && ia < cnt) out[ia] = -sa;
	}
	__syncwarp();
}

// Ranking key of one row with the SAME association as scan_topk_f32_kernel (scan_f32.cuh): two running sums, even
// elements in one and odd elements in the other (the packed f32x2 accumulator of the scan), added at the end. With
// this the tensor-core path (K3 filter + this re-rank) returns bit-identical keys to the SIMT scan on any data.
__device__ __forceinline__ float scan_order_key(const float* __restrict__ row, const float* qsm, uint32_t d_pad, int metric) {
	float lo = 0.f, hi = 0.f;
	for (uint32_t i = 0; i < d_pad; i += 4) {
		const float4 x = ldg128f(row + i);
		const float4 q = *reinterpret_cast<const float4*>(qsm + i);
		if (metric == 0) {
			const float d0 = q.x - x.x, d1 = q.y - x.y, d2 = q.z - x.z, d3 = q.w - x.w;
			lo = fmaf(d0, d0, lo); hi = fmaf(d1, d1, hi); lo = fmaf(d2, d2, lo); hi = fmaf(d3, d3, hi);
		} else {
			lo = fmaf(q.x, x.x, lo); hi = fmaf(q.y, x.y, hi); lo = fmaf(q.z, x.z, lo); hi = fmaf(q.w, x.w, hi);
		}
	}
	const float s = lo + hi;
	return metric == 0 ? s : -s;
}

struct RerankParams {
	const float* base;      // [n][d_pad]
	const float* queries;   // [m][d_pad]
	const uint64_t* cand;   // [m][kp] keys whose low 32 bits are local ids (KEY_EMPTY = none)
	uint64_t* out_keys;     // [m][k]
	uint32_t m, k, kp, cand_stride, d_pad, id_add; // kp entries are read from each query's cand_stride-strided list
	int metric;
	int scan_order; // 1: scan_order_key (lane per candidate) instead of the 8-lane team reduction
};

template <int LPL> __global__ void __launch_bounds__(128) rerank_kernel(const RerankParams p) {
	extern __shared__ uint8_t rr_smem[];
	const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
	const uint32_t q = blockIdx.x * 4 + wib;
	if (q >= p.m) return;
	const size_t per_warp = (size_t)p.d_pad * 4 + GS_MAXNB * 8;
	float* qsm = reinterpret_cast<float*>(rr_smem + per_warp * wib);
	uint32_t* ids = reinterpret_cast<uint32_t*>(qsm + p.d_pad);
	float* dist = reinterpret_cast<float*>(ids + GS_MAXNB);
	for (uint32_t i = lane; i < p.d_pad; i += 32) qsm[i] = p.queries[(size_t)q * p.d_pad + i];
	int cnt = 0;
	for (uint32_t j0 = 0; j0 < p.kp; j0 += 32) {
		const uint64_t key = j0 + lane < p.kp ? p.cand[(size_t)q * p.cand_stride + j0 + lane] : KEY_EMPTY;
		const bool ok = key != KEY_EMPTY;
		const unsigned mk = __ballot_sync(FULL, ok);
		if (ok) ids[cnt + __popc(mk & ((1u << lane) - 1))] = key_id(key);
		cnt += __popc(mk);
	}
	__syncwarp();
	if (p.scan_order) {
		for (int i = lane; i < cnt; i += 32) dist[i] = scan_order_key(p.base + (size_t)ids[i] * p.d_pad, qsm, p.d_pad, p.metric);
		__syncwarp();
	} else if (p.metric == 0) gather_dist_f32(p.base, p.d_pad, qsm, ids, dist, cnt, lane);
	else gather_negdot_f32(p.base, p.d_pad, qsm, ids, dist, cnt, lane);
	WarpList<LPL> fin;
	fin.init();
	for (int i = 0; i < cnt; ++i) fin.insert(make_key(dist[i], ids[i]), lane);
#pragma unroll
	for (int s = 0; s < LPL; ++s) {
		const int i = s * 32 + lane;
		if (i < (int)p.k) p.out_keys[(size_t)q * p.k + i] = fin.e[s] == KEY_EMPTY ? KEY_EMPTY : fin.e[s] + p.id_add;
	}
}

// ---- quantizers ------------------------------------------------------------------------------------------
// quantizer_simple<uint8_t>::build (src/quantizer.h:132-141): code = uint8_t(x) (truncation on [0,256)); also
// emits sum c^2 per row. One warp per row; dst row stride dc (zero padded).
static __global__ void quant_simple_u8_kernel(const float* __restrict__ src, size_t src_stride, uint32_t d, uint8_t* __restrict__ dst,
                                       uint32_t dc, int32_t* __restrict__ norm, size_t n) {
	const size_t row = (size_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
	const int lane = threadIdx.x & 31;
	if (row >= n) return;
	uint32_t s = 0;
	for (uint32_t i = lane; i < dc; i += 32) {
		uint32_t c = 0;
		if (i < d) c = (uint32_t)src[row * src_stride + i] & 0xFFu;
		dst[row * dc + i] = (uint8_t)c;
		s += c * c;
	}
#pragma unroll
	for (int o = 16; o >= 1; o >>= 1) s += __shfl_xor_sync(FULL, s, o);
	if (lane == 0 && norm) norm[row] = (int32_t)s;
}

// quantizer_ranged_q8 (src/quantizer.h:186-238): global min/max -> scale = 128/(max-min), offset = -scale*min;
// code = clamp(round(scale*x+offset), 0, 127), a negative pre-clamp value becomes 127 (size_t wrap, :196-205).
static __global__ void minmax_kernel(const float* __restrict__ src, size_t src_stride, uint32_t d, size_t n, float* mm /*[2]*/) {
	float mx = 1.17549435e-38f /* FLT_MIN, src/quantizer.h:210 */, mn = 3.402823466e+38f;
	const size_t total = n * d;
	for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
		const float v = src[(i / d) * src_stride + (i % d)];
		mx = fmaxf(mx, v);
		mn = fminf(mn, v);
	}
#pragma unroll
	for (int o = 16; o >= 1; o >>= 1) { mx = fmaxf(mx, __shfl_xor_sync(FULL, mx, o)); mn = fminf(mn, __shfl_xor_sync(FULL, mn, o)); }
	if ((threadIdx.x & 31) == 0) {
		// float atomics via the orderable map
		atomicMax(reinterpret_cast<unsigned*>(mm) + 0, f32_orderable(mx));
		atomicMin(reinterpret_cast<unsigned*>(mm) + 1, f32_orderable(mn));
	}
}
static __global__ void quant_ranged_q8_kernel(const float* __restrict__ src, size_t src_stride, uint32_t d, uint8_t* __restrict__ dst,
                                       uint32_t dc, int32_t* __restrict__ norm, size_t n, float scale, float offset) {
	const size_t row = (size_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
	const int lane = threadIdx.x & 31;
	if (row >= n) return;
	uint32_t s = 0;
	for (uint32_t i = lane; i < dc; i += 32) {
		uint32_t c = 0;
		if (i < d) {
			const float r = roundf(__fadd_rn(__fmul_rn(scale, src[row * src_stride + i]), offset));
			c = (r < 0.f || r > 127.f) ? 127u : (uint32_t)r;
		}
		dst[row * dc + i] = (uint8_t)c;
		s += c * c;
	}
#pragma unroll
	for (int o = 16; o >= 1; o >>= 1) s += __shfl_xor_sync(FULL, s, o);
	if (lane == 0 && norm) norm[row] = (int32_t)s;
}

// ---- explicit-pair distance primitives (parity tests of SURVEY §8(a) a2-a8): one warp per pair ----------
enum { PAIR_F32_L2 = 0, PAIR_F32_DOT = 1, PAIR_U8_COMPRESSED = 2, PAIR_I8 = 3, PAIR_I16 = 4 };
static __global__ void pair_distance_kernel(int kind, const void* __restrict__ a, const void* __restrict__ b, size_t npairs,
                                     uint32_t d, float* __restrict__ out_f, int32_t* __restrict__ out_i) {
	const size_t pr = (size_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
	const int lane = threadIdx.x & 31;
	if (pr >= npairs) return;
	if (kind == PAIR_F32_L2 || kind == PAIR_F32_DOT) {
		const float* x = (const float*)a + pr * d;
		const float* y = (const float*)b + pr * d;
		float s = 0.f;
		for (uint32_t i = lane; i < d; i += 32) {
			if (kind == PAIR_F32_L2) { const float t = x[i] - y[i]; s = fmaf(t, t, s); }
			else s = fmaf(x[i], y[i], s);
		}
#pragma unroll
		for (int o = 16; o >= 1; o >>= 1) s += __shfl_xor_sync(FULL, s, o);
		if (lane == 0) out_f[pr] = s;
	} else {
		uint32_t s = 0;
		for (uint32_t i = lane; i < d; i += 32) {
			int32_t t;
			if (kind == PAIR_U8_COMPRESSED) t = (int32_t)(uint32_t)((const float*)a)[pr * d + i] - (int32_t)((const uint8_t*)b)[pr * d + i];
			else if (kind == PAIR_I8) t = (int32_t)((const int8_t*)a)[pr * d + i] - (int32_t)((const int8_t*)b)[pr * d + i];
			else t = (int32_t)((const int16_t*)a)[pr * d + i] - (int32_t)((const int16_t*)b)[pr * d + i];
			s += (uint32_t)(t * t);
		}
#pragma unroll
		for (int o = 16; o >= 1; o >>= 1) s += __shfl_xor_sync(FULL, s, o);
		if (lane == 0) out_i[pr] = (int32_t)s;
	}
}

} // namespace expann
