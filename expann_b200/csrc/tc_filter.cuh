// K3 — tensor-core candidate filter for large query batches (sm_100a: tcgen05.mma + TMEM + TMA).
//
// For a block of 128 queries and a stream of 256-row tiles it evaluates, on the 5th-gen tensor cores,
//     a'(q,x) = ||x||^2 - 2 q.x          ( = dist2(q,x) - ||q||^2, same ranking per query; src/vec.h:139-142 )
// with bf16 operands and fp32 accumulation in TMEM, and appends (a', id) to the query's candidate buffer whenever
// a' <= thr[q]. thr[q] is an upper bound that every true top-k row must satisfy (see tc.cu: it is the k-th
// smallest a' seen so far plus twice the arithmetic error bound), so the buffers always contain the exact top-k;
// K4 (rerank.cuh) then recomputes the survivors in exact fp32 diff form — "candidates are always reranked in fp32".
// Operands: exact mode = bf16(x), bf16(q) for bf16-exact integer data (u8-valued SIFT: products and partial sums are
// exact, error bound 0); split mode = K-concatenated [qh|qh|ql].[xh|xl|xh] with x = xh + xl (two bf16 terms).
//
// CTA = 6 warps: warps 0-3 epilogue (TMEM lane quadrant = warp id), warp 4 TMA producer, warp 5 MMA issuer.
// A (queries) stays resident in shared memory; B (rows) streams through an NSTAGE ring of [256 rows][64 bf16]
// 128B-swizzled boxes; two 256-column fp32 accumulators in TMEM double-buffer MMA against the epilogue.
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include "common.cuh"

namespace expann {

constexpr int TC_BM = 128, TC_BN = 256, TC_BK = 64, TC_UMMA_K = 16;
constexpr int TC_A_KB_BYTES = TC_BM * TC_BK * 2;  // 16 KB per 64-wide k block of the query tile
constexpr int TC_B_STAGE_BYTES = TC_BN * TC_BK * 2; // 32 KB
constexpr int TC_THREADS = 192;

struct TcParams {
	const float* xn;   // [ceil(n/256)*256] ||x||^2 in fp32, NaN beyond n
	const float* thr;  // [m] filter threshold on a'
	uint64_t* cand;    // [m][cap] keys (orderable(a') << 32 | local row id)
	uint32_t* cnt;     // [m] number of appends (may exceed cap: overflow)
	uint32_t cap, m, n, kd;            // kd = bf16 elements per operand row (multiple of 64)
	uint32_t tile0, ntiles, tiles_per_cta; // row tiles [tile0, tile0+ntiles) of this launch
	float scale;       // a' = xn + scale * dot  (-2 for L2; -1 with xn = 0 for inner product)
};

__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(uint64_t* bar) {
	asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// K-major, 128B-swizzled operand tile whose rows are 128 bytes apart: SBO = 8 rows * 128 B, version 1 (sm_100)
__device__ __forceinline__ uint64_t tc_smem_desc(uint32_t saddr) {
	uint64_t d = (uint64_t)((saddr >> 4) & 0x3FFF);
	d |= (uint64_t)(1024 >> 4) << 32;
	d |= (uint64_t)1 << 46;
	d |= (uint64_t)2 << 61;
	return d;
}
// kind::f16, A = B = bf16 (K-major), D = fp32, M = 128, N = 256
__device__ __forceinline__ uint32_t tc_instr_desc() {
	return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(TC_BN >> 3) << 17) | ((uint32_t)(TC_BM >> 4) << 24);
}
__device__ __forceinline__ void tc_mma(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
	asm volatile(
	    "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
	    "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
	    "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
	    : "memory");
}
__device__ __forceinline__ void tc_ld32(uint32_t taddr, uint32_t (&v)[32]) {
	asm volatile(
	    "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
	    "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
	    : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]),
	      "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]), "=r"(v[17]), "=r"(v[18]),
	      "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]),
	      "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
	    : "r"(taddr)
	    : "memory");
	asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

template <int NSTAGE>
__global__ void __launch_bounds__(TC_THREADS, 1)
tc_filter_kernel(const __grid_constant__ CUtensorMap tmap_q, const __grid_constant__ CUtensorMap tmap_x, const TcParams p) {
	extern __shared__ uint8_t tc_smem_raw[];
	uint8_t* smem = (uint8_t*)(((uintptr_t)tc_smem_raw + 1023) & ~(uintptr_t)1023);
	const uint32_t nkb = p.kd / TC_BK;
	uint8_t* smA = smem;
	uint8_t* smB = smem + (size_t)nkb * TC_A_KB_BYTES;
	uint64_t* bars = reinterpret_cast<uint64_t*>(smB + (size_t)NSTAGE * TC_B_STAGE_BYTES);
	uint64_t* a_full = bars;
	uint64_t* full = bars + 1;
	uint64_t* empty = full + NSTAGE;
	uint64_t* tm_full = empty + NSTAGE;
	uint64_t* tm_empty = tm_full + 2;
	uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tm_empty + 2);

	const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
	const uint32_t q0 = blockIdx.x * TC_BM;
	const uint32_t t_begin = blockIdx.y * p.tiles_per_cta;
	uint32_t t_end = t_begin + p.tiles_per_cta;
	if (t_end > p.ntiles) t_end = p.ntiles;
	const uint32_t my_tiles = t_begin < t_end ? t_end - t_begin : 0;

	if (tid == 0) {
		mbar_init(a_full, 1);
		for (int s = 0; s < NSTAGE; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
		for (int s = 0; s < 2; ++s) { mbar_init(&tm_full[s], 1); mbar_init(&tm_empty[s], 4); }
		mbar_fence_init();
	}
	if (warp == 5) { // TMEM: all 512 columns (two 256-column fp32 accumulators)
		asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(tmem_slot)) : "memory");
		asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
	}
	tc_fence_before();
	__syncthreads();
	tc_fence_after();
	const uint32_t tmem_base = *tmem_slot;

	if (warp == 4) {
		// ===================== TMA producer =====================
		if (lane == 0 && my_tiles) {
			tma_prefetch_desc(&tmap_q);
			tma_prefetch_desc(&tmap_x);
			mbar_expect_tx(a_full, nkb * TC_A_KB_BYTES);
			for (uint32_t kb = 0; kb < nkb; ++kb) tma_load_2d(smA + (size_t)kb * TC_A_KB_BYTES, &tmap_q, (int)(kb * TC_BK), (int)q0, a_full);
			uint32_t it = 0;
			for (uint32_t lt = 0; lt < my_tiles; ++lt) {
				const uint32_t row = (p.tile0 + t_begin + lt) * TC_BN;
				for (uint32_t kb = 0; kb < nkb; ++kb, ++it) {
					const uint32_t s = it % NSTAGE, use = it / NSTAGE;
					if (use > 0) mbar_wait(&empty[s], (use - 1) & 1);
					mbar_expect_tx(&full[s], TC_B_STAGE_BYTES);
					tma_load_2d(smB + (size_t)s * TC_B_STAGE_BYTES, &tmap_x, (int)(kb * TC_BK), (int)row, &full[s]);
				}
			}
		}
	} else if (warp == 5) {
		// ===================== MMA issuer (one thread) =====================
		if (lane == 0 && my_tiles) {
			const uint32_t idesc = tc_instr_desc();
			const uint32_t a_addr = smem_u32(smA), b_addr = smem_u32(smB);
			mbar_wait(a_full, 0);
			tc_fence_after();
			uint32_t it = 0;
			for (uint32_t lt = 0; lt < my_tiles; ++lt) {
				const uint32_t acc = lt & 1, use = lt >> 1;
				if (use > 0) { mbar_wait(&tm_empty[acc], (use - 1) & 1); tc_fence_after(); }
				const uint32_t d_tmem = tmem_base + acc * TC_BN;
				for (uint32_t kb = 0; kb < nkb; ++kb, ++it) {
					const uint32_t s = it % NSTAGE;
					mbar_wait(&full[s], (it / NSTAGE) & 1);
					tc_fence_after();
#pragma unroll
					for (int k = 0; k < TC_BK / TC_UMMA_K; ++k) {
						const uint64_t ad = tc_smem_desc(a_addr + kb * TC_A_KB_BYTES + k * TC_UMMA_K * 2);
						const uint64_t bd = tc_smem_desc(b_addr + s * TC_B_STAGE_BYTES + k * TC_UMMA_K * 2);
						tc_mma(d_tmem, ad, bd, idesc, (kb | (uint32_t)k) != 0 ? 1u : 0u);
					}
					tc_commit(&empty[s]); // stage reusable once these MMAs have read it
				}
				tc_commit(&tm_full[acc]); // accumulator complete
			}
		}
	} else {
		// ===================== epilogue: TMEM -> a' -> threshold filter -> append =====================
		const uint32_t qg = q0 + warp * 32 + lane;
		const bool qvalid = qg < p.m;
		const float thr = qvalid ? p.thr[qg] : -__int_as_float(0x7f800000);
		uint64_t* cbuf = p.cand + (size_t)(qvalid ? qg : 0) * p.cap;
		for (uint32_t lt = 0; lt < my_tiles; ++lt) {
			const uint32_t acc = lt & 1;
			mbar_wait(&tm_full[acc], (lt >> 1) & 1);
			tc_fence_after();
			const uint32_t row0 = (p.tile0 + t_begin + lt) * TC_BN;
			const float4* xn4 = reinterpret_cast<const float4*>(p.xn + row0);
#pragma unroll 1
			for (int c = 0; c < TC_BN / 32; ++c) {
				uint32_t v[32];
				tc_ld32(tmem_base + ((uint32_t)(warp * 32) << 16) + acc * TC_BN + c * 32, v);
#pragma unroll
				for (int j4 = 0; j4 < 8; ++j4) {
					const float4 x4 = __ldg(xn4 + c * 8 + j4);
					const float a0 = fmaf(p.scale, __uint_as_float(v[j4 * 4 + 0]), x4.x);
					const float a1 = fmaf(p.scale, __uint_as_float(v[j4 * 4 + 1]), x4.y);
					const float a2 = fmaf(p.scale, __uint_as_float(v[j4 * 4 + 2]), x4.z);
					const float a3 = fmaf(p.scale, __uint_as_float(v[j4 * 4 + 3]), x4.w);
					if ((a0 <= thr) | (a1 <= thr) | (a2 <= thr) | (a3 <= thr)) {
						const float av[4] = { a0, a1, a2, a3 };
#pragma unroll
						for (int e = 0; e < 4; ++e)
							if (av[e] <= thr) {
								const uint32_t pos = atomicAdd(p.cnt + qg, 1u);
								if (pos < p.cap) cbuf[pos] = make_key(av[e], row0 + c * 32 + j4 * 4 + e);
							}
					}
				}
			}
			tc_fence_before();
			__syncwarp();
			if (lane == 0) mbar_arrive(&tm_empty[acc]);
		}
	}
	tc_fence_before();
	__syncthreads();
	if (warp == 5) {
		tc_fence_after();
		asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem_base) : "memory");
	}
}

} // namespace expann
