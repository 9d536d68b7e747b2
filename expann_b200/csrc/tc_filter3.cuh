// K3 (pair form) — the second-form candidate filter (tc_filter2.cuh) on a CLUSTER OF TWO CTAs with
// tcgen05.mma.cta_group::2: one MMA covers 256 queries (128 per CTA) x 256 rows (each CTA stages 128 of them).
//
// Why: with cta_group::1, M = 128, N = 128 both operands of every 64-clk MMA come from the SM's shared memory:
// 4 KB of A + 4 KB of B = 128 B/clk, the whole shared-memory bandwidth; measured, the single-CTA kernel runs its MMA
// side at 62 % of the tensor pipe even with an empty epilogue (profiles/r1_notes.md). In the pair form each SM
// supplies 4 KB of A + 4 KB of B per 128-clk instruction = 64 B/clk, at the same L2 traffic (a row tile is still
// fetched once per 256 queries).
//
// Protocol (rank 0 = leader issues every MMA):
//   * both CTAs load their query block (A) and their half of every B k-block (+ the half's norm terms) with
//     cp.async.bulk.tensor.cta_group::2 into their OWN shared memory, completing bytes on the LEADER's barriers;
//   * the leader's MMA warp waits its full[s] (expects both halves), issues tcgen05.mma.cta_group::2 and commits with
//     .multicast::cluster to empty[s] / tm_full[acc] of BOTH CTAs;
//   * each CTA's TMEM holds the accumulator rows of its own 128 queries: 256 columns x 2 buffers = 512 columns;
//     8 epilogue warps per CTA = (lane quadrant) x (column half): a thread owns (query, half of the tile's rows) and
//     is one "split" of that query as far as bounds, buffers and the select kernels are concerned
//     (buffers per query = 2 x row splits); all 16 epilogue warps of the pair arrive on the leader's tm_empty[acc].
//   * the epilogue thread reads its 128 accumulator columns into registers in one go and the warp hands the accumulator
//     back BEFORE looking at them; chunks whose minimum passes the bound are handled from those registers;
//   * ASTREAM: no resident query tile -- every ring slot carries [A k-block | B k-block | norm terms] (any row length);
//   * seeding (Tc2Params::seed_in / seed_out): a first launch over the leading tiles leaves each stream's sorted list,
//     the main launch starts from their union instead of +inf.
// Measured (B200): bf16-exact 10 000 x 1M x 128 in 1.80 ms (84-86 % of the cuBLAS bf16 burst peak, tensor pipe 82 %
// active with the clock at the power limit); two-term split format 92-103 % of that peak (MMA-bound).
#pragma once
#include "tc_filter2.cuh"

namespace expann {

constexpr int TC3_TN = 256; // rows per tile of the pair (UMMA N), 128 staged by each CTA

__device__ __forceinline__ uint32_t cluster_ctarank() {
	uint32_t r;
	asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
	return r;
}
__device__ __forceinline__ void cluster_sync_all() {
	asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
	asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// shared::cluster address of `p` (a shared::cta pointer of this CTA) in CTA `rank` of the cluster
__device__ __forceinline__ uint32_t mapa_u32(const void* p, uint32_t rank) {
	uint32_t r;
	asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(smem_u32(p)), "r"(rank));
	return r;
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr) {
	asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
// 2-SM TMA load: data into this CTA's shared memory, bytes completed on the barrier at shared::cluster address `mbar`
__device__ __forceinline__ void tma_load_2d_2sm(void* smem_dst, const void* tmap, int c0, int c1, uint32_t mbar_cluster_addr) {
	asm volatile(
	    "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::
	        "r"(smem_u32(smem_dst)),
	    "l"(tmap), "r"(mbar_cluster_addr), "r"(c0), "r"(c1)
	    : "memory");
}
// KIND 2: kind::tf32 (fp32 rows read as tf32, K = 8 per instruction = the same 32 bytes; the fp32 base rows are the operand).
// KIND 0: kind::f16 (bf16 x bf16 -> fp32). KIND 1: kind::i8 (u8 x u8 -> s32, K = 32 per instruction: the same 32 bytes
// of every operand row as a bf16 step, so descriptors and k-block layout are unchanged).
template <int KIND>
__device__ __forceinline__ void tc_mma_2sm(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
	if (KIND == 0)
		asm volatile(
		    "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
		    "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
		    "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
		    : "memory");
	else if (KIND == 1)
		asm volatile(
		    "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
		    "tcgen05.mma.cta_group::2.kind::i8 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
		    "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
		    : "memory");
	else
		asm volatile(
		    "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
		    "tcgen05.mma.cta_group::2.kind::tf32 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
		    "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
		    : "memory");
}
// commit of the leader's MMAs: one arrival on the barrier at this shared-memory offset in BOTH CTAs of the pair
__device__ __forceinline__ void tc_commit_2sm(uint64_t* bar) {
	asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(smem_u32(bar)),
	             "h"((uint16_t)3)
	             : "memory");
}
template <int KIND>
__device__ __forceinline__ uint32_t tc3_instr_desc() { // K-major operands, M = 256 (pair), N = 256
	// KIND 0: D = fp32 (bits 4-5 = 1), A = B = bf16 (format 1 at bits 7-9 / 10-12); KIND 1: D = s32 (2), A = B = unsigned 8 bit (0)
	// KIND 2: D = fp32, A = B = tf32 (format 2)
	const uint32_t fmt = KIND == 0 ? (1u << 4) | (1u << 7) | (1u << 10) : KIND == 1 ? (2u << 4) : (1u << 4) | (2u << 7) | (2u << 10);
	return fmt | ((uint32_t)(TC3_TN >> 3) << 17) | ((uint32_t)(256 >> 4) << 24);
}

struct Tc3Cfg {
	static constexpr int EPI_WARPS = 8;
	static constexpr int THREADS = 32 * (EPI_WARPS + 2);
	// astream: the query operand is not resident; its k-blocks travel through the ring beside the row tiles
	static size_t smem_bytes(uint32_t kd, int nstage, bool astream = false) {
		return 1024 + (astream ? 0 : (size_t)(kd / TC_BK) * TC_A_KB_BYTES) + 4096 +
		       (size_t)nstage * (TC2_STAGE_BYTES + (astream ? TC_A_KB_BYTES : 0)) + (size_t)EPI_WARPS * 4096 +
		       (size_t)(1 + 2 * nstage + 4) * 8 + 16;
	}
};

// p.rsplit = buffers per query = 2 * gridDim.y; p.ntiles / p.tiles_per_cta count 256-row tiles.
// ASTREAM = false: each CTA's 128 queries (all k-blocks) stay resident in shared memory (d <= 128 real-valued, <= 384
// bf16-exact). ASTREAM = true: any d -- a ring slot carries [A k-block 16 KB | B k-block 16 KB | norm terms 4 KB], the
// classic GEMM main loop (A comes from L2 again for every row tile: 64 B/clk per SM of L2 -> shared memory traffic).
// KIND 1 epilogue: int32 dot products of one 32-column chunk -> a'' = cnb[row] - 2 dot (inner product: bias - dot), kept as
// the integer's bit pattern. The biased code norms are the same for every lane: broadcast 16-byte loads through L1.
__device__ __forceinline__ void tc3_code_chunk(uint32_t (&d)[32], const int4* __restrict__ c4, bool ipm, int32_t ibias) {
#pragma unroll
	for (int i = 0; i < 8; ++i) {
		const int4 c = ipm ? make_int4(ibias, ibias, ibias, ibias) : __ldg(c4 + i);
		const int mul = ipm ? 1 : 2;
		d[4 * i] = (uint32_t)(c.x - mul * (int)d[4 * i]);
		d[4 * i + 1] = (uint32_t)(c.y - mul * (int)d[4 * i + 1]);
		d[4 * i + 2] = (uint32_t)(c.z - mul * (int)d[4 * i + 2]);
		d[4 * i + 3] = (uint32_t)(c.w - mul * (int)d[4 * i + 3]);
	}
}

// KIND = 1 (the compressed scan of the quantised engine on the tensor cores): operands are 8-bit codes (query codes x base
// codes, tcgen05.mma.kind::i8, exact int32 dot products in TMEM; p.kd counts 2-byte units, i.e. code bytes / 2), there is
// no norm step, and the epilogue forms  a'' = cnb[row] - 2 dot  (p.cnb = sum c^2 + bias; inner product: bias - dot) as an
// integer in (0, 2^31 - 2^23): such integers order like the positive floats with the same bits, so bound, lists, keys and
// the select kernels of the float path are used unchanged (slack 0).
template <int NSTAGE, bool ASTREAM = false, int KIND = 0>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(Tc3Cfg::THREADS, 1)
tc_filter_pair_kernel(const __grid_constant__ CUtensorMap tmap_q, const __grid_constant__ CUtensorMap tmap_x,
                      const __grid_constant__ CUtensorMap tmap_xx, const Tc2Params p) {
	extern __shared__ uint8_t tc_smem_raw[];
	uint8_t* smem = (uint8_t*)(((uintptr_t)tc_smem_raw + 1023) & ~(uintptr_t)1023);
	const uint32_t nkb = p.kd / TC_BK;
	constexpr uint32_t STG = TC2_STAGE_BYTES + (ASTREAM ? TC_A_KB_BYTES : 0); // ring slot
	constexpr int KELEM = KIND == 0 ? TC_BK : KIND == 1 ? 2 * TC_BK : TC_BK / 2; // tensor-map elements per 128-byte k-block: 64 bf16 | 128 codes | 32 floats
	constexpr uint32_t BOFF = ASTREAM ? TC_A_KB_BYTES : 0;                    // B inside the slot; norm terms follow B
	uint8_t* smA = smem;                                                    // [nkb][128 q][64 bf16] (resident form)
	uint8_t* smOnes = smA + (ASTREAM ? 0 : (size_t)nkb * TC_A_KB_BYTES);    // [128 q][16 bf16]
	uint8_t* smB = smOnes + 4096;                                           // [NSTAGE] slots
	float* sc_all = reinterpret_cast<float*>(smB + (size_t)NSTAGE * STG);
	uint64_t* bars = reinterpret_cast<uint64_t*>(sc_all + Tc3Cfg::EPI_WARPS * 1024);
	uint64_t* a_full = bars;
	uint64_t* full = bars + 1;
	uint64_t* empty = full + NSTAGE;
	uint64_t* tm_full = empty + NSTAGE;
	uint64_t* tm_empty = tm_full + 2;
	uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tm_empty + 2);

	const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
	const uint32_t rank = cluster_ctarank();
	const bool leader = rank == 0;
	const uint32_t q0 = (blockIdx.x >> 1) * 256 + rank * TC_BM;
	const uint32_t split = blockIdx.y;
	const uint32_t t_begin = split * p.tiles_per_cta;
	uint32_t t_end = t_begin + p.tiles_per_cta;
	if (t_end > p.ntiles) t_end = p.ntiles;
	const uint32_t my_tiles = t_begin < t_end ? t_end - t_begin : 0;

	if (tid == 0) {
		mbar_init(a_full, 1);
		for (int s = 0; s < NSTAGE; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
		for (int s = 0; s < 2; ++s) { mbar_init(&tm_full[s], 1); mbar_init(&tm_empty[s], 2 * Tc3Cfg::EPI_WARPS); }
		mbar_fence_init();
	}
	if (KIND == 2) { // constant A rows of the norm step: 32-byte rows of 8 floats [1 1 1 0 | 1 1 1 0]
		for (int i = tid; i < 128 * 8; i += Tc3Cfg::THREADS) reinterpret_cast<float*>(smOnes)[i] = (i & 3) < 3 ? 1.f : 0.f;
	} else {
		for (int i = tid; i < 128 * 16; i += Tc3Cfg::THREADS)
			reinterpret_cast<__nv_bfloat16*>(smOnes)[i] = __float2bfloat16_rn((i & 7) < 3 ? 1.f : 0.f);
	}
	fence_proxy_async();
	if (warp == Tc3Cfg::EPI_WARPS + 1) { // the same logical warp of both CTAs
		asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(tmem_slot)) : "memory");
		asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
	}
	tc_fence_before();
	cluster_sync_all(); // barrier inits, ones tiles and the TMEM address of both CTAs are in place
	tc_fence_after();
	const uint32_t tmem_base = *tmem_slot;

	if (warp == Tc3Cfg::EPI_WARPS) {
		// ===================== TMA producer (both CTAs; bytes complete on the leader's barriers) =====================
		if (my_tiles) {
			const uint32_t a_full_l = mapa_u32(a_full, 0);
			if (elect_one()) {
				tma_prefetch_desc(&tmap_q);
				tma_prefetch_desc(&tmap_x);
				tma_prefetch_desc(&tmap_xx);
				if (!ASTREAM) {
					if (leader) mbar_expect_tx(a_full, 2 * nkb * TC_A_KB_BYTES);
					for (uint32_t kb = 0; kb < nkb; ++kb)
						tma_load_2d_2sm(smA + (size_t)kb * TC_A_KB_BYTES, &tmap_q, (int)(kb * KELEM), (int)q0, a_full_l);
				}
			}
			__syncwarp();
			uint32_t it = 0;
			for (uint32_t lt = 0; lt < my_tiles; ++lt) {
				const uint32_t row = (p.tile_first + t_begin + lt) * TC3_TN + rank * TC2_TN; // this CTA's half of the tile
				for (uint32_t kb = 0; kb < nkb; ++kb, ++it) {
					const uint32_t s = it % NSTAGE, use = it / NSTAGE;
					if (use > 0) mbar_wait(&empty[s], (use - 1) & 1);
					if (elect_one()) {
						const bool last = KIND != 1 && kb + 1 == nkb; // the norm terms of the tile travel with its last k-block
						const uint32_t full_l = mapa_u32(&full[s], 0);
						if (leader) mbar_expect_tx(&full[s], 2 * (BOFF + TC2_B_STAGE_BYTES + (last ? TC2_X_BYTES : 0)));
						if (ASTREAM) tma_load_2d_2sm(smB + (size_t)s * STG, &tmap_q, (int)(kb * KELEM), (int)q0, full_l);
						tma_load_2d_2sm(smB + (size_t)s * STG + BOFF, &tmap_x, (int)(kb * KELEM), (int)row, full_l);
						if (last) tma_load_2d_2sm(smB + (size_t)s * STG + BOFF + TC2_B_STAGE_BYTES, &tmap_xx, 0, (int)row, full_l);
					}
					__syncwarp();
				}
			}
		}
	} else if (warp == Tc3Cfg::EPI_WARPS + 1) {
		// ===================== MMA issuer (leader CTA only) =====================
		if (my_tiles && leader) {
			const uint32_t idesc = tc3_instr_desc<KIND>();
			const uint64_t a_desc0 = tc_smem_desc(smem_u32(smA)), b_desc0 = tc_smem_desc(smem_u32(smB) + BOFF);
			const uint64_t as_desc0 = tc_smem_desc(smem_u32(smB)); // streamed A: first 16 KB of a slot
			const uint64_t ones_desc = tc_smem_desc_sw32(smem_u32(smOnes)), x_desc0 = tc_smem_desc_sw32(smem_u32(smB) + BOFF + TC2_B_STAGE_BYTES);
			if (!ASTREAM) {
				mbar_wait(a_full, 0);
				tc_fence_after();
			}
			if (KIND == 0 && !ASTREAM && nkb == 2 && NSTAGE == 6) {
				// d <= 128, bf16-exact mode: a tile is 2 ring slots = 9 MMAs of 128 clk. One elected block per TILE with
				// incrementally stepped slot numbers: the per-slot barrier / elect / descriptor overhead (~450 clk per slot
				// in the generic loop) is otherwise not covered by so few queued MMAs.
				uint32_t s0 = 0, ph = 0;
				for (uint32_t lt = 0; lt < my_tiles; ++lt) {
					const uint32_t acc = lt & 1, use = lt >> 1;
					if (use > 0) { mbar_wait(&tm_empty[acc], (use - 1) & 1); }
					mbar_wait(&full[s0], ph);
					mbar_wait(&full[s0 + 1], ph);
					tc_fence_after();
					if (elect_one()) {
						const uint32_t d_tmem = tmem_base + acc * TC3_TN;
						const uint64_t bd0 = b_desc0 + (uint64_t)((s0 * TC2_STAGE_BYTES) >> 4), bd1 = bd0 + (uint64_t)(TC2_STAGE_BYTES >> 4);
						const uint64_t ad1 = a_desc0 + (uint64_t)(TC_A_KB_BYTES >> 4);
#pragma unroll
						for (int k = 0; k < 4; ++k) tc_mma_2sm<0>(d_tmem, a_desc0 + 2 * k, bd0 + 2 * k, idesc, k != 0 ? 1u : 0u);
						tc_commit_2sm(&empty[s0]);
#pragma unroll
						for (int k = 0; k < 4; ++k) tc_mma_2sm<0>(d_tmem, ad1 + 2 * k, bd1 + 2 * k, idesc, 1u);
						tc_mma_2sm<0>(d_tmem, ones_desc, x_desc0 + (uint64_t)(((s0 + 1) * TC2_STAGE_BYTES) >> 4), idesc, 1u);
						tc_commit_2sm(&empty[s0 + 1]);
						tc_commit_2sm(&tm_full[acc]);
					}
					__syncwarp();
					s0 += 2;
					if (s0 == 6) { s0 = 0; ph ^= 1; }
				}
			} else {
			uint32_t it = 0;
			for (uint32_t lt = 0; lt < my_tiles; ++lt) {
				const uint32_t acc = lt & 1, use = lt >> 1;
				if (use > 0) { mbar_wait(&tm_empty[acc], (use - 1) & 1); tc_fence_after(); }
				const uint32_t d_tmem = tmem_base + acc * TC3_TN;
				for (uint32_t kb = 0; kb < nkb; ++kb, ++it) {
					const uint32_t s = it % NSTAGE;
					mbar_wait(&full[s], (it / NSTAGE) & 1);
					tc_fence_after();
					if (elect_one()) {
						const uint64_t bd0 = b_desc0 + (uint64_t)((s * STG) >> 4);
						const uint64_t ad0 = ASTREAM ? as_desc0 + (uint64_t)((s * STG) >> 4) : a_desc0 + (uint64_t)((kb * TC_A_KB_BYTES) >> 4);
#pragma unroll
						for (int k = 0; k < TC_BK / TC_UMMA_K; ++k)
							tc_mma_2sm<KIND>(d_tmem, ad0 + 2 * k, bd0 + 2 * k, idesc, (kb | (uint32_t)k) != 0 ? 1u : 0u);
						if (KIND != 1 && kb + 1 == nkb) tc_mma_2sm<KIND>(d_tmem, ones_desc, x_desc0 + (uint64_t)((s * STG) >> 4), idesc, 1u);
						tc_commit_2sm(&empty[s]);
						if (kb + 1 == nkb) tc_commit_2sm(&tm_full[acc]);
					}
					__syncwarp();
				}
			}
			}
		}
	} else {
		// ===================== epilogue: one thread = (one query, one half of the tile's rows) =====================
		const int quad = warp & 3, half = warp >> 2;
		const uint32_t qg = q0 + quad * 32 + lane;
		const bool qvalid = qg < p.m;
		const float INF = __int_as_float(0x7f800000);
		Tc2Thread t;
		t.slack = 0.f;
		if (qvalid && (p.eps_rel > 0.f || p.eps_abs > 0.f)) t.slack = 2.f * (p.eps_rel * sqrtf(p.qn[qg]) * p.xnorm_max + p.eps_abs);
#pragma unroll
		for (int j = 0; j < TC2_KK; ++j) t.v[j] = j < TC2_KK - (int)p.k ? -INF : INF;
		t.thr = qvalid ? INF : -INF;
		if (p.seed_in && qvalid) {
			// the k smallest a' of the first launch's rows (two sorted lists, one per half tile): all of them sit in that
			// launch's candidate buffers, so the k-th of their union with this stream's values bounds the final k-th
			const float* s0 = p.seed_in + (size_t)qg * 32;
#pragma unroll
			for (int j = 0; j < TC2_KK; ++j) t.v[j] = s0[j];
			for (int j = TC2_KK - (int)p.k; j < TC2_KK; ++j) {
				const float aj = s0[16 + j];
#pragma unroll
				for (int i = TC2_KK - 1; i >= 1; --i) t.v[i] = fminf(t.v[i], fmaxf(t.v[i - 1], aj));
				t.v[0] = fminf(t.v[0], aj);
			}
			t.thr = __fadd_ru(t.v[TC2_KK - 1], t.slack);
		}
		float pub = t.thr;
		t.cnt = 0;
		t.cap = p.cap;
		const uint32_t bsplit = p.buf_first + split * 2 + half;
		t.cbuf = p.cand + ((size_t)(qvalid ? qg : 0) * p.rsplit + bsplit) * p.cap;
		float* sc = sc_all + warp * 1024 + lane; // hit-path scratch: value j of this lane at sc[j * 32]
		const uint32_t tm_empty_l0 = mapa_u32(&tm_empty[0], 0), tm_empty_l1 = mapa_u32(&tm_empty[1], 0);
		uint32_t gnext = f32_orderable(INF);
		// more than 16 results per query (thr_out): a stream's bound is ITS 16th smallest; adopting the smallest such bound of
		// all streams would cut into the k'-th smallest overall, so the streams keep to themselves
		const bool share = p.thr_out == nullptr;
		for (uint32_t lt = 0; lt < my_tiles; ++lt) {
			const uint32_t acc = lt & 1;
			const uint32_t row0 = (p.tile_first + t_begin + lt) * TC3_TN + half * TC2_TN;
			if (qvalid && share) {
				t.thr = fminf(t.thr, f32_from_orderable(gnext));
				gnext = __ldcg(p.gthr + qg);
			}
			mbar_wait(&tm_full[acc], (lt >> 1) & 1);
			tc_fence_after();
			const uint32_t tbase = tmem_base + ((uint32_t)(quad * 32) << 16) + acc * TC3_TN + half * TC2_TN;
			if (KIND != 1) {
				// the thread's 128 accumulator columns go to registers in one go and the accumulator is handed back to the MMA
				// warp BEFORE they are looked at: a warp that has hits to work through no longer holds the other 15 warps of
				// the pair (and the next MMA) up, it only has to be back in time for its next tile
				uint32_t dA[32], dB[32], dC[32], dD[32];
				tc_ld32_nowait(tbase, dA);
				tc_ld32_nowait(tbase + 32, dB);
				tc_ld32_nowait(tbase + 64, dC);
				tc_ld32_nowait(tbase + 96, dD);
				tc_wait_ld(dA);
				tc_wait_ld(dB);
				tc_wait_ld(dC);
				tc_wait_ld(dD);
				tc_fence_before();
				__syncwarp();
				if (lane == 0) mbar_arrive_cluster(acc ? tm_empty_l1 : tm_empty_l0); // all 16 epilogue warps of the pair -> leader
					const float mnA = tc2_chunk_min(dA);
				if (__any_sync(FULL, mnA <= t.thr)) tc2_chunk_hits(t, dA, sc, row0, p.n);
				const float mnB = tc2_chunk_min(dB);
				if (__any_sync(FULL, mnB <= t.thr)) tc2_chunk_hits(t, dB, sc, row0 + 32, p.n);
				const float mnC = tc2_chunk_min(dC);
				if (__any_sync(FULL, mnC <= t.thr)) tc2_chunk_hits(t, dC, sc, row0 + 64, p.n);
				const float mnD = tc2_chunk_min(dD);
				if (__any_sync(FULL, mnD <= t.thr)) tc2_chunk_hits(t, dD, sc, row0 + 96, p.n);
			} else {
				// 8-bit codes: two chunks of 32 columns at a time (the code norms need registers too), the accumulator is handed
				// back after the last TMEM read — a tile is 32 MMA steps at d = 960, there is time
				const int4* c4 = reinterpret_cast<const int4*>(p.cnb + row0);
				const bool ipm = p.scale > -1.5f;
				uint32_t dA[32], dB[32];
				tc_ld32_nowait(tbase, dA);
				tc_ld32_nowait(tbase + 32, dB);
				tc_wait_ld(dA);
				tc_wait_ld(dB);
				tc3_code_chunk(dA, c4, ipm, p.ibias);
				const float mnA = tc2_chunk_min(dA);
				if (__any_sync(FULL, mnA <= t.thr)) tc2_chunk_hits(t, dA, sc, row0, p.n);
				tc_ld32_nowait(tbase + 64, dA);
				tc3_code_chunk(dB, c4 + 8, ipm, p.ibias);
				const float mnB = tc2_chunk_min(dB);
				if (__any_sync(FULL, mnB <= t.thr)) tc2_chunk_hits(t, dB, sc, row0 + 32, p.n);
				tc_wait_ld(dA);
				tc_ld32_nowait(tbase + 96, dB);
				tc3_code_chunk(dA, c4 + 16, ipm, p.ibias);
				const float mnC = tc2_chunk_min(dA);
				if (__any_sync(FULL, mnC <= t.thr)) tc2_chunk_hits(t, dA, sc, row0 + 64, p.n);
				tc_wait_ld(dB);
				tc_fence_before();
				__syncwarp();
				if (lane == 0) mbar_arrive_cluster(acc ? tm_empty_l1 : tm_empty_l0);
				tc3_code_chunk(dB, c4 + 24, ipm, p.ibias);
				const float mnD = tc2_chunk_min(dB);
				if (__any_sync(FULL, mnD <= t.thr)) tc2_chunk_hits(t, dB, sc, row0 + 96, p.n);
			}
			if (share && t.thr < pub) { atomicMin(p.gthr + qg, f32_orderable(t.thr)); pub = t.thr; }
		}
		if (qvalid) {
			p.cnt[(size_t)qg * p.rsplit + bsplit] = t.cnt;
			if (p.thr_out) p.thr_out[(size_t)qg * p.rsplit + bsplit] = f32_orderable(t.thr); // the bound this stream ended with
			if (p.seed_out) {
#pragma unroll
				for (int j = 0; j < TC2_KK; ++j) p.seed_out[((size_t)qg * p.seed_lists + split * 2 + half) * 16 + j] = t.v[j];
			}
		}
	}
	tc_fence_before();
	cluster_sync_all(); // nobody leaves (or frees TMEM) while the peer may still read its shared memory / arrive on its barriers
	if (warp == Tc3Cfg::EPI_WARPS + 1) {
		tc_fence_after();
		asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, 512;" ::"r"(tmem_base) : "memory");
	}
}

} // namespace expann
