// Operand preparation and threshold maintenance for K3 (tc_filter.cuh).
#pragma once
#include <cuda_bf16.h>
#include "common.cuh"

namespace expann {

// per-launch statistics of converted values: [0] min (orderable), [1] max (orderable), [2] non-integer/non-bf16 flag,
// [3] max row norm^2 (orderable)
struct TcStats { uint32_t vmin, vmax, inexact, nmax; };

// fp32 rows -> bf16 operand rows (one warp per row).
//   mode 1 (exact): dst[kd] = bf16(x), zero padded
//   mode 2 (split): seg = kd/3; base rows [xh | xl | xh], query rows [qh | qh | ql] with x = xh + xl (+ residual)
// also: norm2[row] = sum x^2 (fp32), value statistics for the exactness decision.
// scale (queries of the second form: -2 for L2, -1 for the inner product; exact, a power of two) multiplies every
// written operand value. xx (base rows of the second form, may be null): bf16[rows][16] holding ||x||^2 as three bf16
// terms hi+mid+lo (24 bits: exact for integer norms < 2^24), each HALVED and written to BOTH 16-byte halves of the
// 32-byte row ([h/2 m/2 l/2 0 0 0 0 0 | h/2 m/2 l/2 0 0 0 0 0]) so that the extra MMA K-step against a row of
// [1 1 1 0 0 0 0 0 | 1 1 1 0 0 0 0 0] adds ||x||^2 whatever the 32B-swizzle does to the two halves; zeros for IP.
static __global__ void tc_convert_kernel(const float* __restrict__ src, size_t src_stride, uint32_t d, size_t rows,
                                  __nv_bfloat16* __restrict__ dst, uint32_t kd, int mode, int is_query,
                                  float* __restrict__ norm2, TcStats* __restrict__ stats, float scale = 1.f,
                                  __nv_bfloat16* __restrict__ xx = nullptr, int xx_norm = 0) {
	const size_t row = (size_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
	const int lane = threadIdx.x & 31;
	if (row >= rows) return;
	const float* x = src + row * src_stride;
	__nv_bfloat16* o = dst + row * kd;
	const uint32_t seg = mode == 2 ? kd / 3 : kd;
	float s = 0.f, mn = 3.4e38f, mx = -3.4e38f;
	uint32_t bad = 0;
	for (uint32_t i = lane; i < seg; i += 32) {
		const float v = i < d ? x[i] : 0.f;
		const __nv_bfloat16 h = __float2bfloat16_rn(v);
		const float hf = __bfloat162float(h);
		if (mode == 2) {
			const __nv_bfloat16 l = __float2bfloat16_rn(v - hf);
			const __nv_bfloat16 hs = __float2bfloat16_rn(scale * hf), ls = __float2bfloat16_rn(scale * __bfloat162float(l));
			o[i] = hs;
			o[seg + i] = is_query ? hs : ls;
			o[2 * seg + i] = is_query ? ls : hs;
		} else
			o[i] = __float2bfloat16_rn(scale * hf);
		if (i < d) {
			s = fmaf(v, v, s);
			mn = fminf(mn, v);
			mx = fmaxf(mx, v);
			bad |= (hf != v) || (v != rintf(v));
		}
	}
#pragma unroll
	for (int of = 16; of >= 1; of >>= 1) {
		s += __shfl_xor_sync(FULL, s, of);
		mn = fminf(mn, __shfl_xor_sync(FULL, mn, of));
		mx = fmaxf(mx, __shfl_xor_sync(FULL, mx, of));
		bad |= __shfl_xor_sync(FULL, bad, of);
	}
	if (xx && lane < 16) {
		float t = 0.f;
		if (xx_norm) {
			const float h = __bfloat162float(__float2bfloat16_rn(s));
			const float mres = s - h, mm = __bfloat162float(__float2bfloat16_rn(mres));
			const float lres = mres - mm, ll = __bfloat162float(__float2bfloat16_rn(lres));
			const int j = lane & 7;
			t = 0.5f * (j == 0 ? h : j == 1 ? mm : j == 2 ? ll : 0.f);
		}
		xx[row * 16 + lane] = __float2bfloat16_rn(t);
	}
	if (lane == 0) {
		if (norm2) norm2[row] = s;
		if (stats) {
			// the statistics only move one way, so a (possibly stale) read that already covers this row's value makes the
			// atomic unnecessary: 10^4 rows would otherwise queue ~4 x 10^4 atomics on four addresses (~20 us)
			volatile TcStats* vs = stats;
			const uint32_t omn = f32_orderable(mn), omx = f32_orderable(mx), os = f32_orderable(s);
			if (omn < vs->vmin) atomicMin(&stats->vmin, omn);
			if (omx > vs->vmax) atomicMax(&stats->vmax, omx);
			if (bad && !vs->inexact) atomicOr(&stats->inexact, 1u);
			if (os > vs->nmax) atomicMax(&stats->nmax, os);
		}
	}
}

static __global__ void tc_fill_kernel(float* p, size_t from, size_t to, float v) {
	const size_t i = from + (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (i < to) p[i] = v;
}

// After a chunk of rows: per query, k-th smallest a' among the buffered candidates -> new threshold
//   thr = a'_k + 2*eps(q)   (every true top-k row has a' <= thr, see tc.cu), buffer compacted to a' <= thr.
// flags[0] counts queries whose buffer overflowed (result not guaranteed -> caller falls back to the SIMT scan).
// final != 0: additionally pads the buffer with KEY_EMPTY up to `pad` entries for the re-rank kernel and flags
// queries with more than `pad` survivors.
template <int LPL>
__global__ void __launch_bounds__(128) tc_select_kernel(uint64_t* __restrict__ cand, uint32_t* __restrict__ cnt, uint32_t cap,
                                                         uint32_t m, uint32_t k, const float* __restrict__ qnorm2, float eps_rel,
                                                         float eps_abs, float xnorm_max, float* __restrict__ thr,
                                                         uint32_t* __restrict__ flags, int final, uint32_t pad) {
	const int lane = threadIdx.x & 31;
	const uint32_t q = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
	if (q >= m) return;
	uint64_t* buf = cand + (size_t)q * cap;
	uint32_t c = cnt[q];
	if (c > cap) {
		if (lane == 0) atomicAdd(flags, 1u);
		c = cap;
	}
	WarpList<LPL> best;
	best.init();
	uint64_t tk = KEY_EMPTY;
	for (uint32_t i0 = 0; i0 < c; i0 += 32) {
		const uint64_t key = i0 + lane < c ? buf[i0 + lane] : KEY_EMPTY;
		unsigned mask = __ballot_sync(FULL, key < tk);
		while (mask) {
			const int src = __ffs(mask) - 1;
			mask &= mask - 1;
			const uint64_t kk = shfl_u64(key, src);
			if (kk < tk) { best.insert(kk, lane); tk = best.at(k - 1); }
		}
	}
	float t = __int_as_float(0x7f800000);
	if (tk != KEY_EMPTY) {
		const float eps = eps_rel * sqrtf(qnorm2[q]) * xnorm_max + eps_abs;
		t = key_dist(tk) + 2.f * eps;
		if (eps > 0.f) t = nextafterf(t, __int_as_float(0x7f800000)); // round the bound up
	}
	// compact in place (reads run ahead of writes)
	uint32_t w = 0;
	for (uint32_t i0 = 0; i0 < c; i0 += 32) {
		const uint64_t key = i0 + lane < c ? buf[i0 + lane] : KEY_EMPTY;
		const bool keep = key != KEY_EMPTY && key_dist(key) <= t;
		const unsigned mk = __ballot_sync(FULL, keep);
		__syncwarp();
		if (keep) buf[w + __popc(mk & ((1u << lane) - 1))] = key;
		w += __popc(mk);
	}
	if (final) {
		for (uint32_t i = w + lane; i < pad; i += 32) buf[i] = KEY_EMPTY;
		if (w > pad && lane == 0) atomicAdd(flags, 1u);
	}
	if (lane == 0) { cnt[q] = w; thr[q] = t; }
}

} // namespace expann

namespace expann {

// ---- operand format 4 (tcgen05.mma.kind::tf32): the fp32 rows themselves are the operand -----------------------------
// a float whose low 13 mantissa bits are zero is read by the tensor core without loss
__device__ __forceinline__ float tf32_trunc(float v) { return __uint_as_float(__float_as_uint(v) & 0xFFFFE000u); }
__device__ __forceinline__ float tf32_round(float v) { // round to nearest even at 10 mantissa bits
	uint32_t u = __float_as_uint(v);
	u += 0xFFFu + ((u >> 13) & 1u);
	return __uint_as_float(u & 0xFFFFE000u);
}

// base rows: ||x||^2 (fp32) and its three tf32 terms for the norm step of the GEMM, as a 32-byte row of 8 floats
// [h/2 m/2 l/2 0 | h/2 m/2 l/2 0] (both 16-byte halves alike, see tc_convert_kernel); value statistics like tc_convert.
static __global__ void tc_norm_tf32_kernel(const float* __restrict__ src, size_t src_stride, uint32_t d, size_t rows,
                                           float* __restrict__ norm2, float* __restrict__ xx, int xx_norm, TcStats* __restrict__ stats) {
	const size_t row = (size_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
	const int lane = threadIdx.x & 31;
	if (row >= rows) return;
	const float* x = src + row * src_stride;
	float s = 0.f, mn = 3.4e38f, mx = -3.4e38f;
	for (uint32_t i = lane; i < d; i += 32) {
		const float v = x[i];
		s = fmaf(v, v, s);
		mn = fminf(mn, v);
		mx = fmaxf(mx, v);
	}
#pragma unroll
	for (int of = 16; of >= 1; of >>= 1) {
		s += __shfl_xor_sync(FULL, s, of);
		mn = fminf(mn, __shfl_xor_sync(FULL, mn, of));
		mx = fmaxf(mx, __shfl_xor_sync(FULL, mx, of));
	}
	if (lane < 8) {
		float t = 0.f;
		if (xx_norm) {
			const float h = tf32_trunc(s), r1 = s - h, mm = tf32_trunc(r1), ll = r1 - mm; // 11 + 11 + 2 bits: exact
			const int j = lane & 3;
			t = 0.5f * (j == 0 ? h : j == 1 ? mm : j == 2 ? ll : 0.f);
		}
		xx[row * 8 + lane] = t;
	}
	if (lane == 0) {
		if (norm2) norm2[row] = s;
		if (stats) {
			volatile TcStats* vs = stats;
			const uint32_t omn = f32_orderable(mn), omx = f32_orderable(mx), os = f32_orderable(s);
			if (omn < vs->vmin) atomicMin(&stats->vmin, omn);
			if (omx > vs->vmax) atomicMax(&stats->vmax, omx);
			if (os > vs->nmax) atomicMax(&stats->nmax, os);
		}
	}
}

// queries: dst[row][d_pad] = tf32(scale * q) rounded to nearest (zero padded), norm2[row] = ||q||^2
static __global__ void tc_query_tf32_kernel(const float* __restrict__ src, size_t src_stride, uint32_t d, uint32_t d_pad, size_t rows,
                                            float* __restrict__ dst, float scale, float* __restrict__ norm2) {
	const size_t row = (size_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
	const int lane = threadIdx.x & 31;
	if (row >= rows) return;
	float s = 0.f;
	for (uint32_t i = lane; i < d_pad; i += 32) {
		const float v = i < d ? src[row * src_stride + i] : 0.f;
		s = fmaf(v, v, s);
		dst[row * d_pad + i] = tf32_round(scale * v);
	}
#pragma unroll
	for (int of = 16; of >= 1; of >>= 1) s += __shfl_xor_sync(FULL, s, of);
	if (lane == 0 && norm2) norm2[row] = s;
}

} // namespace expann
