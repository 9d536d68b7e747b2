// K1/K2 — exhaustive fp32 scan with fused top-k (sm_100a).
//
// Replaces the hot loop of brute_force_engine<T>::_query_k (reference src/brute_force_engine.h:29-46):
//     for i in 0..n-1: d = dist2(v, all_entries[i]); keep the k smallest (d, i)
// with dist2 = sum (a_i-b_i)^2 (src/vec.h:139-142; AVX form src/distance.h:86-148: sub_ps then fmadd_ps per
// element — the same two roundings per element are used here) and, for the inner-product metric, key = -dot
// (src/vec.h:109-111, src/distance.h:181-190; no reference engine ranks by it — "parity unpinned").
//
// Mapping to the machine
//   * base rows are streamed HBM -> shared memory by TMA (cp.async.bulk.tensor.2d, 128B-swizzled boxes of
//     [BOXR rows][32 floats]) through a STAGES-deep mbarrier ring filled by one producer thread; the NWARP
//     consumer warps never issue a global load for the base set.
//   * one CTA holds QT queries in shared memory (broadcast LDS.128) and scans one contiguous row range
//     ("split"); grid = (query tiles, splits) with the query tile as the fast index so that CTAs resident at
//     the same time walk the same rows (the base is then read from HBM ~once per resident wave, the rest
//     are L2 hits).
//   * lane l of consumer warp w owns TR rows of every 32*TR*NWARP-row group and keeps QT*TR running sums in
//     registers; the swizzle makes the lane-per-row LDS.128 bank-conflict free. Packed FADD2/FFMA2
//     (sub.f32x2 / fma.rn.f32x2) halve the issue slots per element.
//   * top-k: ONE sorted WarpList per (CTA, query), owned by consumer warp q (QT <= NWARP), and a CTA-wide bound
//     thr[q] (its k-th distance) in shared memory. After every row group a lane posts the rows with d <= thr[q]
//     into query q's shared-memory mailbox; two named barriers bracket the owners' drain of the mailboxes, which
//     tightens thr[q] for all warps. One list per CTA instead of one per warp divides the warm-up insertions
//     (k * ln(rows/k) per list) by ~NWARP — at 8 queries per pass they were half of all issued instructions.
//     A full mailbox is handled by extra post/drain rounds (far-to-near sorted data). The split results are
//     merged by merge_topk (K7).
#pragma once
#include <cuda.h>
#include "common.cuh"

namespace expann {

enum { METRIC_L2 = 0, METRIC_IP = 1 };

struct ScanParams {
	const float* queries; // [m][d_pad] device, zero padded
	uint64_t* partial;    // [m][nsplit][k] keys (local row ids)
	uint32_t m, k, d_pad;
	uint32_t n;              // rows in this shard
	uint32_t rows_per_split; // multiple of the stage row count
	uint32_t nsplit;
	// fused split merge (nsplit > 1 or id_add != 0): the owner warp that writes the LAST list of a query (ticket on
	// counters[q], self-resetting) merges the nsplit lists into final_out[q][k] with id_add added to the ids
	uint64_t* final_out;
	uint32_t* counters;
	uint32_t id_add;
	// device-side gate (K3's fallback): when set, the launch does nothing unless *gate != 0; a launch that does run
	// counts itself in *gate_count. Lets the tensor-core path queue its SIMT redo without a host round trip.
	const uint32_t* gate;
	uint32_t* gate_count;
};

__device__ __forceinline__ uint64_t pack2(float lo, float hi) {
	return ((uint64_t)__float_as_uint(hi) << 32) | __float_as_uint(lo);
}
__device__ __forceinline__ uint64_t sub2(uint64_t a, uint64_t b) {
	uint64_t r;
	asm("sub.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
	return r;
}
__device__ __forceinline__ uint64_t fma2(uint64_t a, uint64_t b, uint64_t c) {
	uint64_t r;
	asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
	return r;
}
__device__ __forceinline__ ulonglong2 lds128(uint32_t addr) {
	ulonglong2 v;
	asm volatile("ld.shared.v2.b64 {%0, %1}, [%2];" : "=l"(v.x), "=l"(v.y) : "r"(addr));
	return v;
}
__device__ __forceinline__ uint32_t lds_u32(const void* p) {
	uint32_t v;
	asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(smem_u32(p)) : "memory");
	return v;
}
__device__ __forceinline__ void sts_u32(void* p, uint32_t v) { asm volatile("st.shared.u32 [%0], %1;" ::"r"(smem_u32(p)), "r"(v) : "memory"); }
__device__ __forceinline__ uint64_t lds_u64(const void* p) {
	uint64_t v;
	asm volatile("ld.shared.u64 %0, [%1];" : "=l"(v) : "r"(smem_u32(p)) : "memory");
	return v;
}
__device__ __forceinline__ void sts_u64(void* p, uint64_t v) { asm volatile("st.shared.u64 [%0], %1;" ::"r"(smem_u32(p)), "l"(v) : "memory"); }
__device__ __forceinline__ uint32_t atom_inc_shared(void* p) {
	uint32_t v;
	asm volatile("atom.shared.add.u32 %0, [%1], 1;" : "=r"(v) : "r"(smem_u32(p)) : "memory");
	return v;
}
__device__ __forceinline__ float sum2(uint64_t a) {
	return __uint_as_float((uint32_t)a) + __uint_as_float((uint32_t)(a >> 32));
}

// Rendezvous of the consumer warps on an mbarrier (one arrival per warp) instead of bar.sync: measured on B200, a
// bar.sync issued while TMA boxes are in flight into this CTA's shared memory costs about one DRAM round trip (the
// one-query scan spent 56 % of its time in the two per-group bar.syncs), the mbarrier rendezvous does not.
__device__ __forceinline__ void consumer_rendezvous(uint64_t* bar, uint32_t n, int lane) {
	__syncwarp();
	if (lane == 0) mbar_arrive(bar);
	mbar_wait(bar, n & 1);
}

template <int QT, int TR, int NWARP, int STAGES, int METRIC, int LPL>
struct ScanCfg {
	static constexpr int ROWS = 32 * TR * NWARP;          // rows per stage
	static constexpr int BOXR = ROWS > 256 ? 256 : ROWS;  // TMA box rows (<= 256)
	static constexpr int NBOX = ROWS / BOXR;
	static constexpr int STAGE_BYTES = ROWS * 128;
	static constexpr int THREADS = 32 * (NWARP + 1);
	static constexpr int KMAX = 32 * LPL;
	static constexpr int MB_CAP = 128; // mailbox entries per query and round
	// dynamic smem: [align 1024] stages | queries (QT*d_pad*4) | mailboxes (QT*MB_CAP*8) | counts, bounds, flags | barriers
	static size_t smem_bytes(uint32_t d_pad) {
		return 1024 + (size_t)STAGES * STAGE_BYTES + (size_t)QT * d_pad * 4 + (size_t)QT * MB_CAP * 8 + (size_t)QT * 8 + 16 +
		       (2 * STAGES + 1) * 8 + 64;
	}
};

template <int QT, int TR, int NWARP, int STAGES, int METRIC, int LPL>
__global__ void __launch_bounds__(32 * (NWARP + 1), 1)
scan_topk_f32_kernel(const __grid_constant__ CUtensorMap tmap, const ScanParams p) {
	using Cfg = ScanCfg<QT, TR, NWARP, STAGES, METRIC, LPL>;
	if (p.gate != nullptr) { // uniform over the grid: nobody has touched a barrier yet
		if (*reinterpret_cast<const volatile uint32_t*>(p.gate) == 0u) return;
		if (blockIdx.x == 0 && blockIdx.y == 0 && threadIdx.x == 0) atomicAdd(p.gate_count, 1u);
	}
	extern __shared__ uint8_t smem_raw[];
	uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
	static_assert(QT <= NWARP, "one owner warp per query of the tile");
	const size_t stage_area = (size_t)STAGES * Cfg::STAGE_BYTES;
	float* qsm = (float*)(smem + stage_area);
	uint64_t* mbox = (uint64_t*)(smem + stage_area + (size_t)QT * p.d_pad * 4); // [QT][MB_CAP]
	uint32_t* mb_cnt = (uint32_t*)(mbox + QT * Cfg::MB_CAP);                     // [QT]
	float* thr_s = (float*)(mb_cnt + QT);                                        // [QT]
	uint32_t* ovf_s = (uint32_t*)(thr_s + QT);                                   // [2] mailbox-overflow flag per round parity
	uint64_t* full = (uint64_t*)(ovf_s + 4);
	uint64_t* empty = full + STAGES;
	uint64_t* csync = empty + STAGES; // consumer-warp rendezvous (see cta_rendezvous)

	const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
	const uint32_t q0 = blockIdx.x * QT;
	const uint32_t split = blockIdx.y;
	const uint32_t row_begin = split * p.rows_per_split;
	uint32_t row_end = row_begin + p.rows_per_split;
	if (row_end > p.n) row_end = p.n;
	const uint32_t ngroups = row_begin < row_end ? (row_end - row_begin + Cfg::ROWS - 1) / Cfg::ROWS : 0;
	const uint32_t nkc = p.d_pad / 32;
	const uint32_t total = ngroups * nkc;

	// ---- prologue: queries -> smem (zero rows for q >= m), barriers
	for (uint32_t i = tid; i < QT * p.d_pad / 4; i += Cfg::THREADS) {
		uint32_t q = (i * 4) / p.d_pad;
		float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
		if (q0 + q < p.m) v = reinterpret_cast<const float4*>(p.queries + (size_t)q0 * p.d_pad)[i];
		reinterpret_cast<float4*>(qsm)[i] = v;
	}
	if (tid < QT) { mb_cnt[tid] = 0; thr_s[tid] = __int_as_float(0x7f800000); }
	if (tid == 0) {
		ovf_s[0] = ovf_s[1] = 0;
		for (int s = 0; s < STAGES; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], NWARP); }
		mbar_init(csync, NWARP);
		mbar_fence_init();
	}
	__syncthreads();

	if (warp == NWARP) {
		// ===================== TMA producer (one thread) =====================
		if (lane == 0) {
			tma_prefetch_desc(&tmap);
			for (uint32_t it = 0; it < total; ++it) {
				const uint32_t s = it % STAGES, use = it / STAGES;
				if (use > 0) mbar_wait(&empty[s], (use - 1) & 1);
				const uint32_t g = it / nkc, kc = it - g * nkc;
				mbar_expect_tx(&full[s], Cfg::STAGE_BYTES);
#pragma unroll
				for (int h = 0; h < Cfg::NBOX; ++h)
					tma_load_2d(smem + (size_t)s * Cfg::STAGE_BYTES + (size_t)h * Cfg::BOXR * 128, &tmap, (int)(kc * 32),
					            (int)(row_begin + g * Cfg::ROWS + h * Cfg::BOXR), &full[s]);
			}
		}
		return;
	}

	// ===================== consumers =====================
	WarpList<LPL> list; // the CTA's list of query `warp` (owner warps: warp < QT)
	list.init();
	uint64_t tk = KEY_EMPTY;
	uint32_t round = 0, nsync = 0;
	// lane-per-row address inside a stage with the 128B-swizzle folded in: chunk c of the row is at
	// rowoff ^ (c << 4) because rowoff already carries (lane & 7) << 4 in its low 7 bits.
	const uint32_t smem_base = smem_u32(smem), q_base = smem_u32(qsm);
	uint32_t rowoff[TR];
#pragma unroll
	for (int r = 0; r < TR; ++r) rowoff[r] = (uint32_t)(warp * 32 * TR + r * 32 + lane) * 128 + ((lane & 7) << 4);

	uint32_t it = 0;
	for (uint32_t g = 0; g < ngroups; ++g) {
		uint64_t acc[QT][TR];
#pragma unroll
		for (int q = 0; q < QT; ++q)
#pragma unroll
			for (int r = 0; r < TR; ++r) acc[q][r] = 0ull;
		for (uint32_t kc = 0; kc < nkc; ++kc, ++it) {
			const uint32_t s = it % STAGES;
			mbar_wait(&full[s], (it / STAGES) & 1);
			const uint32_t sb = smem_base + s * Cfg::STAGE_BYTES;
			const uint32_t qb = q_base + kc * 128;
#pragma unroll
			for (int c = 0; c < 8; ++c) {
				ulonglong2 xv[TR];
#pragma unroll
				for (int r = 0; r < TR; ++r)
					xv[r] = lds128(sb + (rowoff[r] ^ (c << 4)));
#pragma unroll
				for (int q = 0; q < QT; ++q) {
					const ulonglong2 qv = lds128(qb + q * p.d_pad * 4 + c * 16);
#pragma unroll
					for (int r = 0; r < TR; ++r) {
						if (METRIC == METRIC_L2) {
							uint64_t d0 = sub2(qv.x, xv[r].x), d1 = sub2(qv.y, xv[r].y);
							acc[q][r] = fma2(d0, d0, acc[q][r]);
							acc[q][r] = fma2(d1, d1, acc[q][r]);
						} else {
							acc[q][r] = fma2(qv.x, xv[r].x, acc[q][r]);
							acc[q][r] = fma2(qv.y, xv[r].y, acc[q][r]);
						}
					}
				}
			}
			__syncwarp();
			if (lane == 0) mbar_arrive(&empty[s]);
		}
		// ---- distances of this row group are complete: post candidates, owners drain, bounds tighten
		const uint32_t grow = row_begin + g * Cfg::ROWS + warp * 32 * TR + lane;
		float dist[QT][TR];
		uint32_t pend = 0; // bit q*TR+r: (query q, my row r) still has to be posted
#pragma unroll
		for (int q = 0; q < QT; ++q) {
			const float tq = __uint_as_float(lds_u32(thr_s + q));
#pragma unroll
			for (int r = 0; r < TR; ++r) {
				float d = sum2(acc[q][r]);
				if (METRIC == METRIC_IP) d = -d;
				dist[q][r] = d;
				if (d <= tq && grow + r * 32 < row_end) pend |= 1u << (q * TR + r);
			}
		}
		for (;;) {
			uint32_t* ovf = ovf_s + (round & 1);
			if (pend) {
#pragma unroll
				for (int q = 0; q < QT; ++q)
#pragma unroll
					for (int r = 0; r < TR; ++r)
						if (pend & (1u << (q * TR + r))) {
							const uint32_t slot = atom_inc_shared(mb_cnt + q);
							if (slot < (uint32_t)Cfg::MB_CAP) {
								sts_u64(mbox + q * Cfg::MB_CAP + slot, make_key(dist[q][r], grow + r * 32));
								pend &= ~(1u << (q * TR + r));
							} else
								sts_u32(ovf, 1u);
						}
			}
			consumer_rendezvous(csync, nsync++, lane); // all posts of this round are in
			const uint32_t again = lds_u32(ovf);
			if (tid == 0) sts_u32(ovf_s + ((round + 1) & 1), 0u);
			if (warp < QT) {
				const uint32_t c = min(lds_u32(mb_cnt + warp), (uint32_t)Cfg::MB_CAP);
				if (c) {
					for (uint32_t i0 = 0; i0 < c; i0 += 32) {
						const uint64_t key = i0 + lane < c ? lds_u64(mbox + warp * Cfg::MB_CAP + i0 + lane) : KEY_EMPTY;
						unsigned mask = __ballot_sync(FULL, key < tk);
						while (mask) {
							const int src = __ffs(mask) - 1;
							mask &= mask - 1;
							const uint64_t kk = shfl_u64(key, src);
							if (kk < tk) {
								list.insert(kk, lane);
								tk = list.at(p.k - 1);
							}
						}
					}
					__syncwarp();
					if (lane == 0) {
						sts_u32(mb_cnt + warp, 0u);
						sts_u32(thr_s + warp, tk == KEY_EMPTY ? 0x7f800000u : __float_as_uint(key_dist(tk)));
					}
				}
			}
			consumer_rendezvous(csync, nsync++, lane); // bounds and empty mailboxes visible
			++round;
			if (!again) break;
			// a mailbox was full: rows not yet posted are checked against the tightened bounds and posted again
#pragma unroll
			for (int q = 0; q < QT; ++q) {
				const float tq = __uint_as_float(lds_u32(thr_s + q));
#pragma unroll
				for (int r = 0; r < TR; ++r)
					if (!(dist[q][r] <= tq)) pend &= ~(1u << (q * TR + r));
			}
		}
	}

	// ---- the owner warp of each query writes the CTA's list for this split; the last writer of a query merges
	if (warp < QT && q0 + warp < p.m) {
		const uint32_t qg = q0 + warp;
		uint64_t* dst = p.partial + ((size_t)qg * p.nsplit + split) * p.k;
#pragma unroll
		for (int s = 0; s < LPL; ++s)
			if (s * 32 + lane < (int)p.k) dst[s * 32 + lane] = list.e[s];
		if (p.final_out) {
			__threadfence();
			__syncwarp();
			uint32_t ticket = 0;
			if (lane == 0) ticket = atomicAdd(p.counters + qg, 1u);
			ticket = __shfl_sync(FULL, ticket, 0);
			if (ticket == p.nsplit - 1) {
				__threadfence();
				const uint64_t* src = p.partial + (size_t)qg * p.nsplit * p.k;
				const uint32_t total = p.nsplit * p.k;
				WarpList<LPL> out;
				out.init();
				uint64_t ok = KEY_EMPTY;
				for (uint32_t i0 = 0; i0 < total; i0 += 256) {
					uint64_t kk[8]; // 8 independent loads in flight per lane
#pragma unroll
					for (int u = 0; u < 8; ++u) {
						const uint32_t i = i0 + u * 32 + lane;
						kk[u] = i < total ? __ldcg(src + i) : KEY_EMPTY;
					}
#pragma unroll
					for (int u = 0; u < 8; ++u) {
						unsigned mask = __ballot_sync(FULL, kk[u] < ok);
						while (mask) {
							const int sl = __ffs(mask) - 1;
							mask &= mask - 1;
							const uint64_t c = shfl_u64(kk[u], sl);
							if (c < ok) { out.insert(c, lane); ok = out.at(p.k - 1); }
						}
					}
				}
#pragma unroll
				for (int s = 0; s < LPL; ++s)
					if (s * 32 + lane < (int)p.k) p.final_out[(size_t)qg * p.k + s * 32 + lane] = out.e[s] == KEY_EMPTY ? KEY_EMPTY : out.e[s] + p.id_add;
				if (lane == 0) p.counters[qg] = 0;
			}
		}
	}
}

} // namespace expann
