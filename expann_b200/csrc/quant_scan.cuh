// K6 — exhaustive scan over 8-bit codes with fused top-k' (sm_100a). The "quantizer compressed scan" of
// BASELINE C5 / SURVEY §8(a) a15-a16: quantizer_simple<uint8_t> codes (src/quantizer.h:125-150) scored with
// dist2_compressed (src/antitopo_engine.h:38-61), or quantizer_ranged_q8 codes (src/quantizer.h:186-238) scored
// with the sign-CORRECTED int8 distance (the reference's distance_compare_avx512f_i64, src/distance.h:29-53, is
// wrong for negative differences; deliberate deviation). Survivors are re-ranked in fp32 by K4 (rerank.cuh).
//
// Integer arithmetic is exact and order-free:  sum (q-c)^2 = sum q^2 - 2 sum q*c + sum c^2  with q, c in [0,255],
// so the hot loop is ONE dp4a per 4 dimensions per (query,row) and the result is bit-identical to the reference.
// Same machine mapping as K1 (scan_f32.cuh): TMA 128B-swizzled boxes of [256 rows][128 codes] through a 3-stage
// mbarrier ring, lane-per-row conflict-free LDS.128, broadcast query words, per-warp register top-k lists.
#pragma once
#include <cuda.h>
#include "common.cuh"
#include "scan_f32.cuh"

namespace expann {

struct QScanParams {
	const uint8_t* qcodes;   // [m][dc_pad] query codes (u8), zero padded
	const int32_t* qnorm;    // [m] sum q^2
	const int32_t* cnorm;    // [n] sum c^2
	uint64_t* partial;       // [m][nsplit][k] keys (biased int32 score << 32 | local id)
	uint32_t m, k, dc_pad, n, rows_per_split, nsplit;
	// device-side gate (the tensor-core code filter's redo): when set, the launch does nothing unless *gate != 0; a launch
	// that does run counts itself in *gate_count
	const uint32_t* gate;
	uint32_t* gate_count;
};

template <int QT, int TR, int NWARP, int STAGES, int METRIC, int LPL>
__global__ void __launch_bounds__(32 * (NWARP + 1), 1)
scan_topk_u8_kernel(const __grid_constant__ CUtensorMap tmap, const QScanParams p) {
	using Cfg = ScanCfg<QT, TR, NWARP, STAGES, METRIC, LPL>;
	extern __shared__ uint8_t smem_raw[];
	if (p.gate != nullptr) { // uniform over the grid: nobody has touched a barrier yet
		if (*reinterpret_cast<const volatile uint32_t*>(p.gate) == 0u) return;
		if (blockIdx.x == 0 && blockIdx.y == 0 && threadIdx.x == 0) atomicAdd(p.gate_count, 1u);
	}
	uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
	size_t stage_area = (size_t)STAGES * Cfg::STAGE_BYTES;
	{
		size_t merge = (size_t)NWARP * QT * Cfg::KMAX * 8;
		if (merge > stage_area) stage_area = merge;
	}
	uint8_t* qsm = smem + stage_area;
	uint64_t* full = (uint64_t*)(smem + stage_area + (((size_t)QT * p.dc_pad + 15) & ~(size_t)15));
	uint64_t* empty = full + STAGES;

	const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
	const uint32_t q0 = blockIdx.x * QT, split = blockIdx.y;
	const uint32_t row_begin = split * p.rows_per_split;
	uint32_t row_end = row_begin + p.rows_per_split;
	if (row_end > p.n) row_end = p.n;
	const uint32_t ngroups = row_begin < row_end ? (row_end - row_begin + Cfg::ROWS - 1) / Cfg::ROWS : 0;
	const uint32_t nkc = p.dc_pad / 128;
	const uint32_t total = ngroups * nkc;

	for (uint32_t i = tid; i < QT * p.dc_pad / 16; i += Cfg::THREADS) {
		const uint32_t q = (i * 16) / p.dc_pad;
		uint4 v = make_uint4(0, 0, 0, 0);
		if (q0 + q < p.m) v = reinterpret_cast<const uint4*>(p.qcodes + (size_t)q0 * p.dc_pad)[i];
		reinterpret_cast<uint4*>(qsm)[i] = v;
	}
	if (tid == 0) {
		for (int s = 0; s < STAGES; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], NWARP); }
		mbar_fence_init();
	}
	__syncthreads();

	if (warp == NWARP) {
		if (lane == 0) {
			tma_prefetch_desc(&tmap);
			for (uint32_t it = 0; it < total; ++it) {
				const uint32_t s = it % STAGES, use = it / STAGES;
				if (use > 0) mbar_wait(&empty[s], (use - 1) & 1);
				const uint32_t g = it / nkc, kc = it - g * nkc;
				mbar_expect_tx(&full[s], Cfg::STAGE_BYTES);
#pragma unroll
				for (int h = 0; h < Cfg::NBOX; ++h)
					tma_load_2d(smem + (size_t)s * Cfg::STAGE_BYTES + (size_t)h * Cfg::BOXR * 128, &tmap, (int)(kc * 128),
					            (int)(row_begin + g * Cfg::ROWS + h * Cfg::BOXR), &full[s]);
			}
		}
		return;
	}

	WarpList<LPL> list[QT];
	int32_t thr[QT], qn[QT];
#pragma unroll
	for (int q = 0; q < QT; ++q) {
		list[q].init();
		thr[q] = 0x7fffffff;
		qn[q] = q0 + q < p.m ? p.qnorm[q0 + q] : 0;
	}
	const uint32_t smem_base = smem_u32(smem), q_base = smem_u32(qsm);
	uint32_t rowoff[TR];
#pragma unroll
	for (int r = 0; r < TR; ++r) rowoff[r] = (uint32_t)(warp * 32 * TR + r * 32 + lane) * 128 + ((lane & 7) << 4);

	uint32_t it = 0;
	for (uint32_t g = 0; g < ngroups; ++g) {
		int32_t acc[QT][TR];
#pragma unroll
		for (int q = 0; q < QT; ++q)
#pragma unroll
			for (int r = 0; r < TR; ++r) acc[q][r] = 0;
		const uint32_t grow = row_begin + g * Cfg::ROWS + warp * 32 * TR + lane;
		int32_t cn[TR];
#pragma unroll
		for (int r = 0; r < TR; ++r) cn[r] = (METRIC == METRIC_L2 && grow + r * 32 < row_end) ? __ldg(p.cnorm + grow + r * 32) : 0;
		for (uint32_t kc = 0; kc < nkc; ++kc, ++it) {
			const uint32_t s = it % STAGES;
			mbar_wait(&full[s], (it / STAGES) & 1);
			const uint32_t sb = smem_base + s * Cfg::STAGE_BYTES;
			const uint32_t qb = q_base + kc * 128;
#pragma unroll
			for (int c = 0; c < 8; ++c) {
				ulonglong2 xv[TR];
#pragma unroll
				for (int r = 0; r < TR; ++r) xv[r] = lds128(sb + (rowoff[r] ^ (c << 4)));
#pragma unroll
				for (int q = 0; q < QT; ++q) {
					const ulonglong2 qv = lds128(qb + q * p.dc_pad + c * 16);
#pragma unroll
					for (int r = 0; r < TR; ++r) {
						int32_t a = acc[q][r];
						a = __dp4a((uint32_t)qv.x, (uint32_t)xv[r].x, (uint32_t)a);
						a = __dp4a((uint32_t)(qv.x >> 32), (uint32_t)(xv[r].x >> 32), (uint32_t)a);
						a = __dp4a((uint32_t)qv.y, (uint32_t)xv[r].y, (uint32_t)a);
						a = __dp4a((uint32_t)(qv.y >> 32), (uint32_t)(xv[r].y >> 32), (uint32_t)a);
						acc[q][r] = a;
					}
				}
			}
			__syncwarp();
			if (lane == 0) mbar_arrive(&empty[s]);
		}
		int32_t sc[QT][TR];
		bool any = false;
#pragma unroll
		for (int q = 0; q < QT; ++q)
#pragma unroll
			for (int r = 0; r < TR; ++r) {
				// wrapping 32-bit arithmetic like the reference's epi32 lanes
				const int32_t s = METRIC == METRIC_L2 ? (int32_t)((uint32_t)qn[q] + (uint32_t)cn[r] - 2u * (uint32_t)acc[q][r])
				                                      : (int32_t)(0u - (uint32_t)acc[q][r]);
				sc[q][r] = s;
				any |= (s <= thr[q]) && (grow + r * 32 < row_end);
			}
		if (__any_sync(FULL, any)) {
#pragma unroll
			for (int q = 0; q < QT; ++q) {
#pragma unroll
				for (int r = 0; r < TR; ++r) {
					unsigned mask = __ballot_sync(FULL, (sc[q][r] <= thr[q]) && (grow + r * 32 < row_end));
					if (mask) {
						uint64_t tk = list[q].at(p.k - 1);
						while (mask) {
							const int src = __ffs(mask) - 1;
							mask &= mask - 1;
							const int32_t sv = __shfl_sync(FULL, sc[q][r], src);
							const uint64_t key = make_key_i32(sv, grow - lane + src + r * 32);
							if (key < tk) {
								list[q].insert(key, lane);
								tk = list[q].at(p.k - 1);
							}
						}
						thr[q] = tk == KEY_EMPTY ? 0x7fffffff : key_score_i32(tk);
					}
				}
			}
		}
	}

	asm volatile("bar.sync 1, %0;" ::"n"(32 * NWARP) : "memory");
	uint64_t* msm = (uint64_t*)smem;
#pragma unroll
	for (int q = 0; q < QT; ++q)
#pragma unroll
		for (int s = 0; s < LPL; ++s) msm[((size_t)warp * QT + q) * Cfg::KMAX + s * 32 + lane] = list[q].e[s];
	asm volatile("bar.sync 1, %0;" ::"n"(32 * NWARP) : "memory");
	for (int q = warp; q < QT; q += NWARP) {
		if (q0 + q >= p.m) continue;
		WarpList<LPL> out;
		out.init();
		uint64_t tk = KEY_EMPTY;
		for (int w = 0; w < NWARP; ++w) {
#pragma unroll
			for (int s = 0; s < LPL; ++s) {
				const uint64_t key = msm[((size_t)w * QT + q) * Cfg::KMAX + s * 32 + lane];
				unsigned mask = __ballot_sync(FULL, key < tk);
				while (mask) {
					const int src = __ffs(mask) - 1;
					mask &= mask - 1;
					const uint64_t kk = shfl_u64(key, src);
					if (kk < tk) {
						out.insert(kk, lane);
						tk = out.at(p.k - 1);
					}
				}
			}
		}
		uint64_t* dst = p.partial + ((size_t)(q0 + q) * p.nsplit + split) * p.k;
#pragma unroll
		for (int s = 0; s < LPL; ++s)
			if (s * 32 + lane < (int)p.k) dst[s * 32 + lane] = out.e[s];
	}
}

} // namespace expann
