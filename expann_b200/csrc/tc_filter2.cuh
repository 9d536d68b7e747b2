// K3 (second form) — single-launch tensor-core candidate filter for k <= 16 (sm_100a: tcgen05.mma + TMEM + TMA).
//
// Same arithmetic as tc_filter.cuh:  a'(q,x) = ||x||^2 + scale * q.x  on bf16 operands with fp32 accumulation in TMEM
// (a' = dist2(q,x) - ||q||^2, src/vec.h:139-142; scale = -2, or -1 with ||x||^2 := 0 for the inner product), but the
// filter threshold is maintained INSIDE the kernel instead of by a host loop over row chunks:
//   * one CTA owns NQB blocks of 128 queries (A operand resident in shared memory) and one contiguous row range
//     ("split"); B streams through an NSTAGE ring of [128 rows][64 bf16] 128B-swizzled TMA boxes that both query
//     blocks consume, so a row tile is fetched from L2 once per 128*NQB queries (the L2->SM path, ~42 B/clk/SM, is
//     the limit next to the tensor pipe at 128 queries per fetch);
//   * TMEM holds 2 x NQB accumulators of 128 columns: the MMA warp fills one set while 4*NQB epilogue warps drain
//     the other. An epilogue thread owns one query (TMEM lane): it sees every row of the split for that query, keeps
//     the 16 smallest a' seen in registers (sorted; min/max insertion network, entered only when a' <= thr) and hence
//     its own bound  thr = (k-th smallest a' seen) + 2*eps.  Bounds are shared between the splits of a query through
//     gthr[q] (atomicMin on the orderable encoding): the k-th smallest of ANY subset of rows is >= the final k-th
//     smallest, so min over threads is still a valid bound — later waves start with the bound of the earlier ones;
//   * (a', row) is appended to the private buffer of (query, split) when a' <= thr: no atomics on the append path.
// tc_select2_kernel then takes, per query, the exact k-th smallest a' over all its buffers, keeps a' <= A_k + 2*eps
// and hands those rows to the fp32 re-rank (K4). See tc.cu for why this is exact.
#pragma once
#include "tc_filter.cuh"

namespace expann {

constexpr int TC2_TN = 128;                         // rows per tile (UMMA N)
constexpr int TC2_B_STAGE_BYTES = TC2_TN * TC_BK * 2; // 16 KB
constexpr int TC2_KK = 16;                          // per-thread sorted values (k <= 16)

struct Tc2Params {
	const float* xn;     // [ceil(n/128)*128 + 128] ||x||^2, NaN beyond n
	const float* qn;     // [m] ||q||^2 (error bound of the split format)
	uint32_t* gthr;      // [m] shared bound per query, f32_orderable encoding, initialised to orderable(+inf)
	uint64_t* cand;      // [m][rsplit][cap] keys (orderable(a') << 32 | local row)
	uint32_t* cnt;       // [m][rsplit] appends (may exceed cap: overflow)
	uint32_t cap, m, n, kd, k;
	uint32_t ntiles, tiles_per_cta, rsplit;
	float scale, eps_rel, eps_abs, xnorm_max;
};

__device__ __forceinline__ uint32_t tc2_instr_desc() { // kind::f16, bf16 x bf16 -> fp32, K-major A and B, M = 128, N = 128
	return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(TC2_TN >> 3) << 17) | ((uint32_t)(TC_BM >> 4) << 24);
}
__device__ __forceinline__ float min3f(float a, float b, float c) {
	float r;
	asm("min.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c));
	return r;
}
__device__ __forceinline__ float4 lds128f(const float* p) {
	float4 v;
	asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(smem_u32(p)));
	return v;
}

template <int NQB> struct Tc2Cfg {
	static constexpr int EPI_WARPS = 4 * NQB;
	static constexpr int THREADS = 32 * (EPI_WARPS + 2);
	static constexpr int TMEM_COLS = 2 * NQB * TC2_TN; // 256 or 512
	static size_t smem_bytes(uint32_t kd, int nstage) {
		return 1024 + (size_t)NQB * (kd / TC_BK) * TC_A_KB_BYTES + (size_t)nstage * TC2_B_STAGE_BYTES + (size_t)EPI_WARPS * TC2_TN * 4 +
		       (size_t)(1 + 2 * nstage + 4) * 8 + 16;
	}
};

template <int NQB, int NSTAGE>
__global__ void __launch_bounds__(32 * (4 * NQB + 2), 1)
tc_filter2_kernel(const __grid_constant__ CUtensorMap tmap_q, const __grid_constant__ CUtensorMap tmap_x, const Tc2Params p) {
	using Cfg = Tc2Cfg<NQB>;
	extern __shared__ uint8_t tc_smem_raw[];
	uint8_t* smem = (uint8_t*)(((uintptr_t)tc_smem_raw + 1023) & ~(uintptr_t)1023);
	const uint32_t nkb = p.kd / TC_BK;
	uint8_t* smA = smem;                                            // [NQB][nkb][128 q][64 bf16]
	uint8_t* smB = smA + (size_t)NQB * nkb * TC_A_KB_BYTES;         // [NSTAGE][128 rows][64 bf16]
	float* xs_all = reinterpret_cast<float*>(smB + (size_t)NSTAGE * TC2_B_STAGE_BYTES); // [EPI_WARPS][128]
	uint64_t* bars = reinterpret_cast<uint64_t*>(xs_all + Cfg::EPI_WARPS * TC2_TN);
	uint64_t* a_full = bars;
	uint64_t* full = bars + 1;
	uint64_t* empty = full + NSTAGE;
	uint64_t* tm_full = empty + NSTAGE;
	uint64_t* tm_empty = tm_full + 2;
	uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tm_empty + 2);

	const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
	const uint32_t q0 = blockIdx.x * (TC_BM * NQB);
	const uint32_t split = blockIdx.y;
	const uint32_t t_begin = split * p.tiles_per_cta;
	uint32_t t_end = t_begin + p.tiles_per_cta;
	if (t_end > p.ntiles) t_end = p.ntiles;
	const uint32_t my_tiles = t_begin < t_end ? t_end - t_begin : 0;

	if (tid == 0) {
		mbar_init(a_full, 1);
		for (int s = 0; s < NSTAGE; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
		for (int s = 0; s < 2; ++s) { mbar_init(&tm_full[s], 1); mbar_init(&tm_empty[s], Cfg::EPI_WARPS); }
		mbar_fence_init();
	}
	if (warp == Cfg::EPI_WARPS + 1) {
		asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "n"(Cfg::TMEM_COLS) : "memory");
		asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
	}
	tc_fence_before();
	__syncthreads();
	tc_fence_after();
	const uint32_t tmem_base = *tmem_slot;

	if (warp == Cfg::EPI_WARPS) {
		// ===================== TMA producer (one thread) =====================
		if (lane == 0 && my_tiles) {
			tma_prefetch_desc(&tmap_q);
			tma_prefetch_desc(&tmap_x);
			mbar_expect_tx(a_full, NQB * nkb * TC_A_KB_BYTES);
			for (int qb = 0; qb < NQB; ++qb)
				for (uint32_t kb = 0; kb < nkb; ++kb)
					tma_load_2d(smA + ((size_t)qb * nkb + kb) * TC_A_KB_BYTES, &tmap_q, (int)(kb * TC_BK), (int)(q0 + qb * TC_BM), a_full);
			uint32_t it = 0;
			for (uint32_t lt = 0; lt < my_tiles; ++lt) {
				const uint32_t row = (t_begin + lt) * TC2_TN;
				for (uint32_t kb = 0; kb < nkb; ++kb, ++it) {
					const uint32_t s = it % NSTAGE, use = it / NSTAGE;
					if (use > 0) mbar_wait(&empty[s], (use - 1) & 1);
					mbar_expect_tx(&full[s], TC2_B_STAGE_BYTES);
					tma_load_2d(smB + (size_t)s * TC2_B_STAGE_BYTES, &tmap_x, (int)(kb * TC_BK), (int)row, &full[s]);
				}
			}
		}
	} else if (warp == Cfg::EPI_WARPS + 1) {
		// ===================== MMA issuer (one thread) =====================
		if (lane == 0 && my_tiles) {
			const uint32_t idesc = tc2_instr_desc();
			const uint32_t a_addr = smem_u32(smA), b_addr = smem_u32(smB);
			mbar_wait(a_full, 0);
			tc_fence_after();
			uint32_t it = 0;
			for (uint32_t lt = 0; lt < my_tiles; ++lt) {
				const uint32_t acc = lt & 1, use = lt >> 1;
				if (use > 0) { mbar_wait(&tm_empty[acc], (use - 1) & 1); tc_fence_after(); }
				for (uint32_t kb = 0; kb < nkb; ++kb, ++it) {
					const uint32_t s = it % NSTAGE;
					mbar_wait(&full[s], (it / NSTAGE) & 1);
					tc_fence_after();
#pragma unroll
					for (int qb = 0; qb < NQB; ++qb) {
						const uint32_t d_tmem = tmem_base + (acc * NQB + qb) * TC2_TN;
#pragma unroll
						for (int k = 0; k < TC_BK / TC_UMMA_K; ++k) {
							const uint64_t ad = tc_smem_desc(a_addr + (qb * nkb + kb) * TC_A_KB_BYTES + k * TC_UMMA_K * 2);
							const uint64_t bd = tc_smem_desc(b_addr + s * TC2_B_STAGE_BYTES + k * TC_UMMA_K * 2);
							tc_mma(d_tmem, ad, bd, idesc, (kb | (uint32_t)k) != 0 ? 1u : 0u);
						}
					}
					tc_commit(&empty[s]); // stage reusable once these MMAs have read it
				}
				tc_commit(&tm_full[acc]); // both query blocks' accumulators of this tile complete
			}
		}
	} else {
		// ===================== epilogue: one thread = one query =====================
		const int qb = warp >> 2, quad = warp & 3;
		const uint32_t qg = q0 + qb * TC_BM + quad * 32 + lane;
		const bool qvalid = qg < p.m;
		const float INF = __int_as_float(0x7f800000);
		float slack = 0.f;
		if (qvalid && (p.eps_rel > 0.f || p.eps_abs > 0.f)) slack = 2.f * (p.eps_rel * sqrtf(p.qn[qg]) * p.xnorm_max + p.eps_abs);
		float v[TC2_KK]; // ascending; the first 16-k slots hold -inf for good, so v[15] is the k-th smallest a' seen
#pragma unroll
		for (int j = 0; j < TC2_KK; ++j) v[j] = j < TC2_KK - (int)p.k ? -INF : INF;
		float thr = qvalid ? INF : -INF, pub = thr; // invalid lanes never append and never publish
		uint32_t cnt = 0;
		uint64_t* cbuf = p.cand + ((size_t)(qvalid ? qg : 0) * p.rsplit + split) * p.cap;
		float* xs = xs_all + warp * TC2_TN;
		float4 xnext = make_float4(0.f, 0.f, 0.f, 0.f);
		if (my_tiles) xnext = __ldg(reinterpret_cast<const float4*>(p.xn + (size_t)t_begin * TC2_TN) + lane);
		for (uint32_t lt = 0; lt < my_tiles; ++lt) {
			const uint32_t acc = lt & 1;
			const uint32_t row0 = (t_begin + lt) * TC2_TN;
			if (qvalid) thr = fminf(thr, f32_from_orderable(__ldcg(p.gthr + qg)));
			__syncwarp();
			reinterpret_cast<float4*>(xs)[lane] = xnext;
			__syncwarp();
			if (lt + 1 < my_tiles) xnext = __ldg(reinterpret_cast<const float4*>(p.xn + (size_t)(row0 + TC2_TN)) + lane);
			mbar_wait(&tm_full[acc], (lt >> 1) & 1);
			tc_fence_after();
			const uint32_t tbase = tmem_base + ((uint32_t)(quad * 32) << 16) + (acc * NQB + qb) * TC2_TN;
#pragma unroll 1
			for (int c = 0; c < TC2_TN / 32; ++c) {
				uint32_t d[32];
				tc_ld32(tbase + c * 32, d);
				float a[32];
#pragma unroll
				for (int j4 = 0; j4 < 8; ++j4) {
					const float4 x4 = lds128f(xs + c * 32 + j4 * 4);
					a[j4 * 4 + 0] = fmaf(p.scale, __uint_as_float(d[j4 * 4 + 0]), x4.x);
					a[j4 * 4 + 1] = fmaf(p.scale, __uint_as_float(d[j4 * 4 + 1]), x4.y);
					a[j4 * 4 + 2] = fmaf(p.scale, __uint_as_float(d[j4 * 4 + 2]), x4.z);
					a[j4 * 4 + 3] = fmaf(p.scale, __uint_as_float(d[j4 * 4 + 3]), x4.w);
				}
				// min over the 32 values (NaN = padded row: ignored by min and by the compare below)
				float mn[11];
#pragma unroll
				for (int j = 0; j < 10; ++j) mn[j] = min3f(a[3 * j], a[3 * j + 1], a[3 * j + 2]);
				mn[10] = fminf(a[30], a[31]);
				const float m0 = min3f(mn[0], mn[1], mn[2]), m1 = min3f(mn[3], mn[4], mn[5]), m2 = min3f(mn[6], mn[7], mn[8]);
				const float mall = min3f(min3f(m0, m1, m2), mn[9], mn[10]);
				if (mall <= thr) {
#pragma unroll
					for (int j = 0; j < 32; ++j) {
						const float aj = a[j];
						if (aj <= thr) {
							if (cnt < p.cap) cbuf[cnt] = make_key(aj, row0 + c * 32 + j);
							++cnt;
#pragma unroll
							for (int i = TC2_KK - 1; i >= 1; --i) v[i] = fminf(v[i], fmaxf(v[i - 1], aj));
							v[0] = fminf(v[0], aj);
							thr = fminf(thr, __fadd_ru(v[TC2_KK - 1], slack));
						}
					}
				}
			}
			tc_fence_before();
			__syncwarp();
			if (lane == 0) mbar_arrive(&tm_empty[acc]);
			if (thr < pub) { atomicMin(p.gthr + qg, f32_orderable(thr)); pub = thr; }
		}
		if (qvalid) p.cnt[(size_t)qg * p.rsplit + split] = cnt;
	}
	tc_fence_before();
	__syncthreads();
	if (warp == Cfg::EPI_WARPS + 1) {
		tc_fence_after();
		asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(Cfg::TMEM_COLS) : "memory");
	}
}

// Per query: exact k-th smallest key over the buffers of all its splits -> t = a'_k + 2*eps; survivors (a' <= t) are
// written to out[q][pad] (KEY_EMPTY padded) for the fp32 re-rank. flags[0] counts queries whose result is not
// guaranteed (a buffer overflowed, or more than `pad` survivors): the caller redoes the batch with the SIMT scan.
template <int LPL>
__global__ void __launch_bounds__(128) tc_select2_kernel(const uint64_t* __restrict__ cand, const uint32_t* __restrict__ cnt, uint32_t cap,
                                                          uint32_t rsplit, uint32_t m, uint32_t k, const float* __restrict__ qnorm2,
                                                          float eps_rel, float eps_abs, float xnorm_max, uint64_t* __restrict__ out,
                                                          uint32_t pad, uint32_t* __restrict__ flags) {
	const int lane = threadIdx.x & 31;
	const uint32_t q = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
	if (q >= m) return;
	const uint64_t* qbuf = cand + (size_t)q * rsplit * cap;
	const uint32_t* qcnt = cnt + (size_t)q * rsplit;
	WarpList<LPL> best;
	best.init();
	uint64_t tk = KEY_EMPTY;
	bool over = false;
	for (uint32_t s = 0; s < rsplit; ++s) {
		uint32_t c = qcnt[s];
		if (c > cap) { over = true; c = cap; }
		const uint64_t* buf = qbuf + (size_t)s * cap;
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
	}
	float t = __int_as_float(0x7f800000);
	if (tk != KEY_EMPTY) {
		const float eps = eps_rel * sqrtf(qnorm2[q]) * xnorm_max + eps_abs;
		t = __fadd_ru(key_dist(tk), 2.f * eps);
	}
	uint64_t* o = out + (size_t)q * pad;
	uint32_t w = 0;
	for (uint32_t s = 0; s < rsplit; ++s) {
		const uint32_t c = min(qcnt[s], cap);
		const uint64_t* buf = qbuf + (size_t)s * cap;
		for (uint32_t i0 = 0; i0 < c; i0 += 32) {
			const uint64_t key = i0 + lane < c ? buf[i0 + lane] : KEY_EMPTY;
			const bool keep = key != KEY_EMPTY && key_dist(key) <= t;
			const unsigned mk = __ballot_sync(FULL, keep);
			const uint32_t pos = w + __popc(mk & ((1u << lane) - 1));
			if (keep && pos < pad) o[pos] = key;
			w += __popc(mk);
		}
	}
	for (uint32_t i = w + lane; i < pad; i += 32) o[i] = KEY_EMPTY;
	if ((over || w > pad) && lane == 0) atomicAdd(flags, 1u);
}

} // namespace expann
