// K3 (second form) — single-launch tensor-core candidate filter for k <= 16 (sm_100a: tcgen05.mma + TMEM + TMA).
//
// Same quantity as tc_filter.cuh:  a'(q,x) = ||x||^2 + scale * q.x  on bf16 operands with fp32 accumulation in TMEM
// (a' = dist2(q,x) - ||q||^2, src/vec.h:139-142; scale = -2, or -1 with ||x||^2 := 0 for the inner product), but
//   * the WHOLE of a' comes out of the tensor core: the query operand is pre-scaled (exact, a power of two) and
//     ||x||^2 enters through one extra K=16 MMA step per tile, [1 1 1 0..] x [xn_hi xn_mid xn_lo 0..] (three bf16 terms
//     = 24 bits; operand rows of 32 bytes, 32B-swizzle descriptors, see tc_convert_kernel) — the epilogue is
//     tcgen05.ld + a min tree, no FMA and no norm loads;
//   * the filter threshold is maintained INSIDE the kernel instead of by a host loop over row chunks:
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
#include "tc_prep.cuh"

namespace expann {

constexpr int TC2_TN = 128;                         // rows per tile (UMMA N)
constexpr int TC2_B_STAGE_BYTES = TC2_TN * TC_BK * 2; // 16 KB k-block of B
constexpr int TC2_X_BYTES = TC2_TN * 32;              // 4 KB: [128 rows][16 bf16] norm terms of the tile
constexpr int TC2_STAGE_BYTES = TC2_B_STAGE_BYTES + TC2_X_BYTES; // ring slot: k-block + (last k-block only) norm terms
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
	// pair form only: a short first launch over the leading tiles leaves every query's k smallest a' in seed_out; the main
	// launch starts its lists from them (seed_in), so its streams do not each warm a bound up from +inf
	uint32_t tile_first, buf_first; // this launch covers tiles [tile_first, tile_first + ntiles), buffers from buf_first
	const float* seed_in;           // [m][2][16] or null
	float* seed_out;                // [m][seed_lists][16] or null: list 2 * (row split of this launch) + half tile
	uint32_t seed_lists;            // lists per query in seed_out (2 x the seeding launch's row splits)
	// more than 16 results per query (the code scan's k' survivors): every stream still keeps ITS 16 smallest (k = 16) and
	// reports the bound it ended with; the select kernel takes the k'-th smallest key T over all buffers and verifies
	// T <= every stream's final bound -- then no stream has dropped a row at or below T and the buffers hold the exact top
	// k'; otherwise (row order that piles more than 16 of the top k' into one stream) the batch is flagged and redone
	uint32_t* thr_out;              // [m][rsplit] orderable final bound per stream, or null
	// pair form, KIND 1 (8-bit codes, tcgen05.mma.kind::i8): biased code norms and the bias (see tc_filter3.cuh)
	const int32_t* cnb;             // [ceil(n/256)*256 + 256] sum c^2 + ibias, rows >= n: 0x7F000000
	int32_t ibias;
};

__device__ __forceinline__ uint32_t tc2_instr_desc() { // kind::f16, bf16 x bf16 -> fp32, K-major A and B, M = 128, N = 128
	return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(TC2_TN >> 3) << 17) | ((uint32_t)(TC_BM >> 4) << 24);
}
// K-major operand tile with 32-byte rows (K = 16 bf16 = one UMMA_K step), 32B swizzle: 8-row groups are 256 B apart
__device__ __forceinline__ uint64_t tc_smem_desc_sw32(uint32_t saddr) {
	uint64_t d = (uint64_t)((saddr >> 4) & 0x3FFF);
	d |= (uint64_t)1 << 16;
	d |= (uint64_t)(256 >> 4) << 32;
	d |= (uint64_t)1 << 46;
	d |= (uint64_t)6 << 61;
	return d;
}
__device__ __forceinline__ bool elect_one() {
	uint32_t pred;
	asm volatile("{\n\t.reg .pred P;\n\telect.sync _|P, 0xffffffff;\n\tselp.b32 %0, 1, 0, P;\n\t}" : "=r"(pred));
	return pred != 0;
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


// per-thread (= per-query) filter state of the epilogue
struct Tc2Thread {
	float v[TC2_KK]; // ascending; the first 16-k slots hold -inf for good, so v[15] is the k-th smallest a' seen
	float thr, slack;
	uint32_t cnt, cap;
	uint64_t* cbuf;
};
__device__ __forceinline__ void tc2_event(Tc2Thread& t, float aj, uint32_t row) {
	if (t.cnt < t.cap) t.cbuf[t.cnt] = make_key(aj, row);
	++t.cnt;
#pragma unroll
	for (int i = TC2_KK - 1; i >= 1; --i) t.v[i] = fminf(t.v[i], fmaxf(t.v[i - 1], aj));
	t.v[0] = fminf(t.v[0], aj);
	t.thr = fminf(t.thr, __fadd_ru(t.v[TC2_KK - 1], t.slack));
}
__device__ __forceinline__ uint64_t pack_b64(uint32_t lo, uint32_t hi) {
	uint64_t r;
	asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "r"(lo), "r"(hi));
	return r;
}
__device__ __forceinline__ void unpack_f32x2(uint64_t v, float& lo, float& hi) {
	asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ uint64_t fma_f32x2(uint64_t a, uint64_t b, uint64_t c) {
	uint64_t r;
	asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
	return r;
}
// pass 1 of a chunk: only the minimum of its 32 accumulator columns (= a' values of one query). The minima of the
// eleven triples are left in mn[] for pass 2.
__device__ __forceinline__ float tc2_chunk_min(const uint32_t (&d)[32], float (&mn)[11]) {
#pragma unroll
	for (int j = 0; j < 10; ++j) mn[j] = min3f(__uint_as_float(d[3 * j]), __uint_as_float(d[3 * j + 1]), __uint_as_float(d[3 * j + 2]));
	mn[10] = fminf(__uint_as_float(d[30]), __uint_as_float(d[31]));
	const float m0 = min3f(mn[0], mn[1], mn[2]), m1 = min3f(mn[3], mn[4], mn[5]), m2 = min3f(mn[6], mn[7], mn[8]);
	return min3f(min3f(m0, m1, m2), mn[9], mn[10]);
}
__device__ __forceinline__ float tc2_chunk_min(const uint32_t (&d)[32]) {
	float mn[11];
	return tc2_chunk_min(d, mn);
}
// pass 2 (rare per lane, frequent per warp on short row ranges): a chunk known to hold some a' <= thr. The values are
// parked in this lane's scratch column and the hit mask is built without branches; only the set bits (usually one) are
// visited. thr tightens while we go, hence the second compare. Rows >= n (zero-filled by TMA: a' = 0) are masked here.
__device__ __forceinline__ void tc2_chunk_hits(Tc2Thread& t, const uint32_t (&d)[32], float* sc, uint32_t row_c, uint32_t n) {
	const uint32_t sca = smem_u32(sc);
	uint32_t mask = 0;
#pragma unroll
	for (int j = 0; j < 32; ++j) {
		asm volatile("st.shared.b32 [%0], %1;" ::"r"(sca + j * 128), "r"(d[j]) : "memory");
		mask |= __uint_as_float(d[j]) <= t.thr ? 1u << j : 0u;
	}
	while (mask) {
		const int j = __ffs(mask) - 1;
		mask &= mask - 1;
		float aj;
		asm volatile("ld.shared.f32 %0, [%1];" : "=f"(aj) : "r"(sca + j * 128) : "memory");
		if (aj <= t.thr && row_c + j < n) tc2_event(t, aj, row_c + j);
	}
}
// Four other forms of this pass were built and measured on B200 (filter kernel, 10 000 queries x 125k / 1M rows of C2;
// this form: 0.418-0.434 / 1.78-1.84 ms; profiles/r2_notes.md):
//   * branching down the min tree to the hit triple and parking only that triple: 0.555 / 2.27 ms;
//   * the whole warp working on one hit lane at a time (16-byte parking, one vote per hit lane for its mask, values
//     handed over by shuffles): 0.856 / 2.20 ms -- all lanes that have hits run the loop below SIDE BY SIDE, serialising
//     over hit lanes multiplies the event cost by their number;
//   * 16-byte parking into a lane-private row + set/lop3 mask: the compiler turns each pair into FSETP + SEL + LOP3
//     (no gain); + an 11-bit mask of the hit triples instead of the 32-bit one: 0.475 / 1.96 ms.
__device__ __forceinline__ void tc_ld32_nowait(uint32_t taddr, uint32_t (&v)[32]) {
	asm volatile(
	    "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
	    "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
	    : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]),
	      "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]), "=r"(v[17]), "=r"(v[18]),
	      "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]),
	      "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
	    : "r"(taddr)
	    : "memory");
}
// the wait names the loaded registers as in/out operands so that no use of them can be scheduled above it
__device__ __forceinline__ void tc_wait_ld(uint32_t (&v)[32]) {
	asm volatile("tcgen05.wait::ld.sync.aligned;"
	             : "+r"(v[0]), "+r"(v[1]), "+r"(v[2]), "+r"(v[3]), "+r"(v[4]), "+r"(v[5]), "+r"(v[6]), "+r"(v[7]), "+r"(v[8]), "+r"(v[9]),
	               "+r"(v[10]), "+r"(v[11]), "+r"(v[12]), "+r"(v[13]), "+r"(v[14]), "+r"(v[15]), "+r"(v[16]), "+r"(v[17]), "+r"(v[18]),
	               "+r"(v[19]), "+r"(v[20]), "+r"(v[21]), "+r"(v[22]), "+r"(v[23]), "+r"(v[24]), "+r"(v[25]), "+r"(v[26]), "+r"(v[27]),
	               "+r"(v[28]), "+r"(v[29]), "+r"(v[30]), "+r"(v[31])
	             :
	             : "memory");
}

template <int NQB> struct Tc2Cfg {
	static constexpr int EPI_WARPS = 4 * NQB;
	static constexpr int THREADS = 32 * (EPI_WARPS + 2);
	static constexpr int TMEM_COLS = 2 * NQB * TC2_TN; // 256 or 512
	static size_t smem_bytes(uint32_t kd, int nstage) {
		return 1024 + (size_t)NQB * (kd / TC_BK) * TC_A_KB_BYTES + 4096 + (size_t)nstage * TC2_STAGE_BYTES + (size_t)EPI_WARPS * 4096 +
		       (size_t)(1 + 2 * nstage + 4) * 8 + 16;
	}
};

template <int NQB, int NSTAGE>
__global__ void __launch_bounds__(32 * (4 * NQB + 2), 1)
tc_filter2_kernel(const __grid_constant__ CUtensorMap tmap_q, const __grid_constant__ CUtensorMap tmap_x,
                  const __grid_constant__ CUtensorMap tmap_xx, const Tc2Params p) {
	using Cfg = Tc2Cfg<NQB>;
	extern __shared__ uint8_t tc_smem_raw[];
	uint8_t* smem = (uint8_t*)(((uintptr_t)tc_smem_raw + 1023) & ~(uintptr_t)1023);
	const uint32_t nkb = p.kd / TC_BK;
	uint8_t* smA = smem;                                            // [NQB][nkb][128 q][64 bf16]
	uint8_t* smOnes = smA + (size_t)NQB * nkb * TC_A_KB_BYTES;      // [128 q][16 bf16] = [1 1 1 0 0 0 0 0 | 1 1 1 0 0 0 0 0]
	uint8_t* smB = smOnes + 4096;                                   // [NSTAGE]{[128 rows][64 bf16] k-block, [128 rows][16 bf16] norm terms}
	float* sc_all = reinterpret_cast<float*>(smB + (size_t)NSTAGE * TC2_STAGE_BYTES);  // [EPI_WARPS][32][32]
	uint64_t* bars = reinterpret_cast<uint64_t*>(sc_all + Cfg::EPI_WARPS * 1024);
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
	for (int i = tid; i < 128 * 16; i += Cfg::THREADS) // the constant A rows of the norm step
		reinterpret_cast<__nv_bfloat16*>(smOnes)[i] = __float2bfloat16_rn((i & 7) < 3 ? 1.f : 0.f);
	fence_proxy_async(); // generic-proxy writes -> visible to the tensor core (async proxy)
	if (warp == Cfg::EPI_WARPS + 1) {
		asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "n"(Cfg::TMEM_COLS) : "memory");
		asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
	}
	tc_fence_before();
	__syncthreads();
	tc_fence_after();
	const uint32_t tmem_base = *tmem_slot;

	if (warp == Cfg::EPI_WARPS) {
		// ===================== TMA producer =====================
		// The whole warp walks the loop converged and one elected lane issues: ptxas then emits the uniform-datapath
		// instructions (UTMALDG, UTCHMMA, UTCBAR) directly instead of an ELECT / BRA.U.ANY loop around each of them.
		if (my_tiles) {
			if (elect_one()) {
				tma_prefetch_desc(&tmap_q);
				tma_prefetch_desc(&tmap_x);
				tma_prefetch_desc(&tmap_xx);
				mbar_expect_tx(a_full, NQB * nkb * TC_A_KB_BYTES);
				for (int qb = 0; qb < NQB; ++qb)
					for (uint32_t kb = 0; kb < nkb; ++kb)
						tma_load_2d(smA + ((size_t)qb * nkb + kb) * TC_A_KB_BYTES, &tmap_q, (int)(kb * TC_BK), (int)(q0 + qb * TC_BM), a_full);
			}
			__syncwarp();
			uint32_t it = 0;
			for (uint32_t lt = 0; lt < my_tiles; ++lt) {
				const uint32_t row = (t_begin + lt) * TC2_TN;
				for (uint32_t kb = 0; kb < nkb; ++kb, ++it) {
					const uint32_t s = it % NSTAGE, use = it / NSTAGE;
					if (use > 0) mbar_wait(&empty[s], (use - 1) & 1);
					if (elect_one()) {
						const bool last = kb + 1 == nkb; // the tile's norm terms travel with its last k-block
						mbar_expect_tx(&full[s], TC2_B_STAGE_BYTES + (last ? TC2_X_BYTES : 0));
						tma_load_2d(smB + (size_t)s * TC2_STAGE_BYTES, &tmap_x, (int)(kb * TC_BK), (int)row, &full[s]);
						if (last) tma_load_2d(smB + (size_t)s * TC2_STAGE_BYTES + TC2_B_STAGE_BYTES, &tmap_xx, 0, (int)row, &full[s]);
					}
					__syncwarp();
				}
			}
		}
	} else if (warp == Cfg::EPI_WARPS + 1) {
		// ===================== MMA issuer (converged warp, one elected lane issues) =====================
		// (A second issuer warp taking every other tile was tried: +3 %, and it needs 2*nkb <= NSTAGE or the parity
		// waits alias — not kept.)
		if (my_tiles) {
			const uint32_t idesc = tc2_instr_desc();
			const uint64_t a_desc0 = tc_smem_desc(smem_u32(smA)), b_desc0 = tc_smem_desc(smem_u32(smB));
			const uint64_t ones_desc = tc_smem_desc_sw32(smem_u32(smOnes)), x_desc0 = tc_smem_desc_sw32(smem_u32(smB) + TC2_B_STAGE_BYTES);
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
					if (elect_one()) {
						// descriptors: the start-address field counts 16-byte units, so a k step of 16 bf16 (32 B) adds 2
						const uint64_t bd0 = b_desc0 + (uint64_t)((s * TC2_STAGE_BYTES) >> 4);
						const bool last = kb + 1 == nkb;
#pragma unroll
						for (int qb = 0; qb < NQB; ++qb) {
							const uint32_t d_tmem = tmem_base + (acc * NQB + qb) * TC2_TN;
							const uint64_t ad0 = a_desc0 + (uint64_t)(((qb * nkb + kb) * TC_A_KB_BYTES) >> 4);
#pragma unroll
							for (int k = 0; k < TC_BK / TC_UMMA_K; ++k)
								tc_mma(d_tmem, ad0 + 2 * k, bd0 + 2 * k, idesc, (kb | (uint32_t)k) != 0 ? 1u : 0u);
							if (last) // + ||x||^2: [1 1 1 0..] x [xn_hi xn_mid xn_lo 0..], 32-byte operand rows
								tc_mma(d_tmem, ones_desc, x_desc0 + (uint64_t)((s * TC2_STAGE_BYTES) >> 4), idesc, 1u);
						}
						tc_commit(&empty[s]);                        // stage reusable once these MMAs have read it
						if (kb + 1 == nkb) tc_commit(&tm_full[acc]); // both query blocks' accumulators of this tile complete
					}
					__syncwarp();
				}
			}
		}
	} else {
		// ===================== epilogue: one thread = one query =====================
		const int qb = warp >> 2, quad = warp & 3;
		const uint32_t qg = q0 + qb * TC_BM + quad * 32 + lane;
		const bool qvalid = qg < p.m;
		const float INF = __int_as_float(0x7f800000);
		Tc2Thread t;
		t.slack = 0.f;
		if (qvalid && (p.eps_rel > 0.f || p.eps_abs > 0.f)) t.slack = 2.f * (p.eps_rel * sqrtf(p.qn[qg]) * p.xnorm_max + p.eps_abs);
#pragma unroll
		for (int j = 0; j < TC2_KK; ++j) t.v[j] = j < TC2_KK - (int)p.k ? -INF : INF;
		t.thr = qvalid ? INF : -INF; // invalid lanes never append and never publish
		float pub = t.thr;
		t.cnt = 0;
		t.cap = p.cap;
		t.cbuf = p.cand + ((size_t)(qvalid ? qg : 0) * p.rsplit + split) * p.cap;
		float* sc = sc_all + warp * 1024 + lane; // slow-path scratch: value j of this lane at sc[j * 32]
		uint32_t gnext = f32_orderable(INF); // bound of the other splits, fetched one tile ahead (any stale value is valid)
		for (uint32_t lt = 0; lt < my_tiles; ++lt) {
			const uint32_t acc = lt & 1;
			const uint32_t row0 = (t_begin + lt) * TC2_TN;
			if (qvalid) {
				t.thr = fminf(t.thr, f32_from_orderable(gnext));
				gnext = __ldcg(p.gthr + qg);
			}
			mbar_wait(&tm_full[acc], (lt >> 1) & 1);
			tc_fence_after();
			const uint32_t tbase = tmem_base + ((uint32_t)(quad * 32) << 16) + (acc * NQB + qb) * TC2_TN;
			// pass 1, branch-free: the minimum of each of the 4 chunks of 32 columns (two TMEM loads in flight)
			uint32_t dA[32], dB[32];
			tc_ld32_nowait(tbase, dA);
			tc_ld32_nowait(tbase + 32, dB);
			tc_wait_ld(dA);
			tc_wait_ld(dB);
			const float mn0 = tc2_chunk_min(dA);
			tc_ld32_nowait(tbase + 64, dA);
			const float mn1 = tc2_chunk_min(dB);
			tc_ld32_nowait(tbase + 96, dB);
			tc_wait_ld(dA);
			tc_wait_ld(dB);
			const float mn2 = tc2_chunk_min(dA);
			const float mn3 = tc2_chunk_min(dB);
			const uint32_t hm = (mn0 <= t.thr ? 1u : 0u) | (mn1 <= t.thr ? 2u : 0u) | (mn2 <= t.thr ? 4u : 0u) | (mn3 <= t.thr ? 8u : 0u);
			// pass 2, rare: chunks in which some lane has a candidate are read again from TMEM (warp-collective load)
			if (__any_sync(FULL, hm != 0)) {
#pragma unroll 1
				for (int c = 0; c < 4; ++c) {
					if (!__any_sync(FULL, (hm >> c) & 1u)) continue;
					tc_ld32_nowait(tbase + c * 32, dA);
					tc_wait_ld(dA);
					if ((hm >> c) & 1u) tc2_chunk_hits(t, dA, sc, row0 + c * 32, p.n);
				}
			}
			tc_fence_before();
			__syncwarp();
			if (lane == 0) mbar_arrive(&tm_empty[acc]);
			if (t.thr < pub) { atomicMin(p.gthr + qg, f32_orderable(t.thr)); pub = t.thr; }
		}
		if (qvalid) p.cnt[(size_t)qg * p.rsplit + split] = t.cnt;
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
// One warp per query. The per-split counts become an exclusive prefix in shared memory, and both passes walk the
// FLAT index space of all buffered keys (binary search of the prefix per key, 4 independent loads in flight per lane):
// with up to 296 splits a split-by-split walk is a chain of dependent L2 round trips (180 us measured at 148 splits).
constexpr int TC2_MAX_SPLIT = 296;

// Seeding launch with several row splits: every (split, half tile) stream leaves its sorted list of 16; the main launch wants
// ONE list per query (its seed_in: [m][2][16], second list unused = +inf). One warp per query: lane l holds lists l and
// l + 32 (lists <= 64), k rounds of "smallest value left anywhere" by shuffles. Slots below 16 - k stay -inf.
static __global__ void tc_seed_merge_kernel(const float* __restrict__ lists, uint32_t nlists, uint32_t m, uint32_t k, float* __restrict__ seed) {
	const uint32_t q = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
	const int lane = threadIdx.x & 31;
	if (q >= m) return;
	const float INF = __int_as_float(0x7f800000);
	float v[32];
#pragma unroll
	for (int h = 0; h < 2; ++h) {
		const uint32_t l = (uint32_t)lane + 32u * h;
		const float4* src = reinterpret_cast<const float4*>(lists + ((size_t)q * nlists + l) * 16);
#pragma unroll
		for (int j = 0; j < 4; ++j) {
			const float4 x = l < nlists ? __ldg(src + j) : make_float4(INF, INF, INF, INF);
			v[h * 16 + 4 * j] = x.x; v[h * 16 + 4 * j + 1] = x.y; v[h * 16 + 4 * j + 2] = x.z; v[h * 16 + 4 * j + 3] = x.w;
		}
	}
	// the fixed -inf slots of a list (below 16 - k) are not values
#pragma unroll
	for (int j = 0; j < 32; ++j)
		if ((j & 15) < TC2_KK - (int)k) v[j] = INF;
	float* out = seed + (size_t)q * 32;
	if (lane < 16) { out[16 + lane] = INF; if (lane < TC2_KK - (int)k) out[lane] = -INF; }
	for (uint32_t r = 0; r < k; ++r) {
		float mn = v[0];
#pragma unroll
		for (int j = 1; j < 32; ++j) mn = fminf(mn, v[j]);
		float wm = mn;
#pragma unroll
		for (int of = 16; of >= 1; of >>= 1) wm = fminf(wm, __shfl_xor_sync(FULL, wm, of));
		const uint32_t has = __ballot_sync(FULL, mn == wm);
		if (lane == __ffs(has) - 1) { // one instance leaves
			bool done = false;
#pragma unroll
			for (int j = 0; j < 32; ++j)
				if (!done && v[j] == wm) { v[j] = INF; done = true; }
			out[TC2_KK - k + r] = wm;
		}
	}
}

__device__ __forceinline__ uint64_t tc2_flat_key(const uint64_t* __restrict__ qbuf, const uint32_t* pre, uint32_t rsplit, uint32_t cap,
                                                 uint32_t i, uint32_t total) {
	if (i >= total) return KEY_EMPTY;
	uint32_t lo = 0, hi = rsplit; // largest s with pre[s] <= i
	while (hi - lo > 1) {
		const uint32_t mid = (lo + hi) >> 1;
		if (pre[mid] <= i) lo = mid; else hi = mid;
	}
	return __ldg(qbuf + (size_t)lo * cap + (i - pre[lo]));
}

// k > 16 (Tc2Params::thr_out): the selection bound t must not exceed the bound any stream of the query ended with
__device__ __forceinline__ bool tc2_bound_verified(const uint32_t* __restrict__ thr_fin, uint32_t q, uint32_t rsplit, float t, int lane) {
	if (thr_fin == nullptr) return true;
	const uint32_t to = f32_orderable(t);
	bool ok = true;
	for (uint32_t s = lane; s < rsplit; s += 32) ok = ok && to <= thr_fin[(size_t)q * rsplit + s];
	return __all_sync(FULL, ok);
}

template <int LPL>
__global__ void __launch_bounds__(128) tc_select2_kernel(const uint64_t* __restrict__ cand, const uint32_t* __restrict__ cnt, uint32_t cap,
                                                          uint32_t rsplit, uint32_t m, uint32_t k, const float* __restrict__ qnorm2,
                                                          float eps_rel, float eps_abs, float xnorm_max, uint64_t* __restrict__ out,
                                                          uint32_t pad, uint32_t* __restrict__ flags,
                                                          const uint32_t* __restrict__ thr_fin = nullptr) {
	__shared__ uint32_t pre_s[4][TC2_MAX_SPLIT + 8];
	const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
	const uint32_t q = blockIdx.x * (blockDim.x >> 5) + wib;
	if (q >= m) return;
	uint32_t* pre = pre_s[wib];
	const uint64_t* qbuf = cand + (size_t)q * rsplit * cap;
	const uint32_t* qcnt = cnt + (size_t)q * rsplit;
	uint32_t total = 0;
	bool over = false;
	for (uint32_t s0 = 0; s0 < rsplit; s0 += 32) {
		const uint32_t s = s0 + lane;
		uint32_t c = s < rsplit ? qcnt[s] : 0u;
		if (c > cap) { over = true; c = cap; }
		uint32_t incl = c;
#pragma unroll
		for (int o = 1; o < 32; o <<= 1) {
			const uint32_t up = __shfl_up_sync(FULL, incl, o);
			if (lane >= o) incl += up;
		}
		if (s < rsplit) pre[s] = total + incl - c;
		total += __shfl_sync(FULL, incl, 31);
	}
	over = __any_sync(FULL, over);
	__syncwarp();
	uint64_t* o = out + (size_t)q * pad;
	if (total <= 512) {
		// the usual case (a few hundred keys per query): all keys in registers, 16 independent loads per lane in flight;
		// the k-th smallest a' by a bitwise search over the orderable encoding (32 rounds of 16 compares + one warp sum)
		// instead of ~200 dependent list insertions; survivors are compacted from the same registers
		uint64_t kk[16];
#pragma unroll
		for (int u = 0; u < 16; ++u) kk[u] = tc2_flat_key(qbuf, pre, rsplit, cap, u * 32 + lane, total);
		float t = __int_as_float(0x7f800000);
		if (total >= k) {
			uint32_t v = 0;
#pragma unroll 1
			for (int bit = 31; bit >= 0; --bit) {
				const uint32_t trial = v | (1u << bit);
				uint32_t c = 0;
#pragma unroll
				for (int u = 0; u < 16; ++u) c += (uint32_t)(kk[u] >> 32) < trial ? 1u : 0u; // KEY_EMPTY has hi = 0xffffffff: never below
				c = __reduce_add_sync(FULL, c);
				if (c < k) v = trial;
			}
			const float eps = eps_rel * sqrtf(qnorm2[q]) * xnorm_max + eps_abs;
			t = __fadd_ru(f32_from_orderable(v), 2.f * eps);
		}
		uint32_t w = 0;
#pragma unroll
		for (int u = 0; u < 16; ++u) {
			const bool keep = kk[u] != KEY_EMPTY && key_dist(kk[u]) <= t;
			const unsigned mk = __ballot_sync(FULL, keep);
			const uint32_t pos = w + __popc(mk & ((1u << lane) - 1));
			if (keep && pos < pad) o[pos] = kk[u];
			w += __popc(mk);
		}
		for (uint32_t i = w + lane; i < pad; i += 32) o[i] = KEY_EMPTY;
		const bool ver = tc2_bound_verified(thr_fin, q, rsplit, t, lane);
		if ((over || w > pad || !ver) && lane == 0) atomicAdd(flags, 1u);
		return;
	}
	WarpList<LPL> best;
	best.init();
	uint64_t tk = KEY_EMPTY;
	for (uint32_t i0 = 0; i0 < total; i0 += 128) {
		uint64_t kk[4];
#pragma unroll
		for (int u = 0; u < 4; ++u) kk[u] = tc2_flat_key(qbuf, pre, rsplit, cap, i0 + u * 32 + lane, total);
#pragma unroll
		for (int u = 0; u < 4; ++u) {
			unsigned mask = __ballot_sync(FULL, kk[u] < tk);
			while (mask) {
				const int src = __ffs(mask) - 1;
				mask &= mask - 1;
				const uint64_t c = shfl_u64(kk[u], src);
				if (c < tk) { best.insert(c, lane); tk = best.at(k - 1); }
			}
		}
	}
	float t = __int_as_float(0x7f800000);
	if (tk != KEY_EMPTY) {
		const float eps = eps_rel * sqrtf(qnorm2[q]) * xnorm_max + eps_abs;
		t = __fadd_ru(key_dist(tk), 2.f * eps);
	}
	uint32_t w = 0;
	for (uint32_t i0 = 0; i0 < total; i0 += 128) {
		uint64_t kk[4];
#pragma unroll
		for (int u = 0; u < 4; ++u) kk[u] = tc2_flat_key(qbuf, pre, rsplit, cap, i0 + u * 32 + lane, total);
#pragma unroll
		for (int u = 0; u < 4; ++u) {
			const bool keep = kk[u] != KEY_EMPTY && key_dist(kk[u]) <= t;
			const unsigned mk = __ballot_sync(FULL, keep);
			const uint32_t pos = w + __popc(mk & ((1u << lane) - 1));
			if (keep && pos < pad) o[pos] = kk[u];
			w += __popc(mk);
		}
	}
	for (uint32_t i = w + lane; i < pad; i += 32) o[i] = KEY_EMPTY;
	const bool ver = tc2_bound_verified(thr_fin, q, rsplit, t, lane);
	if ((over || w > pad || !ver) && lane == 0) atomicAdd(flags, 1u);
}


// Block-per-query form of the select for launches with many row splits (small batches: every split warms up cold in
// the same wave, ~k*ln(rows/k) keys each, 148 splits -> ~10^4 keys per query). 8 warps take contiguous eighths of the
// flat key space (the split index advances incrementally), their lists meet in shared memory, warp 0 derives t, and
// the second pass appends the survivors through a shared counter.
template <int LPL>
__global__ void __launch_bounds__(256) tc_select2_block_kernel(const uint64_t* __restrict__ cand, const uint32_t* __restrict__ cnt, uint32_t cap,
                                                                uint32_t rsplit, uint32_t m, uint32_t k, const float* __restrict__ qnorm2,
                                                                float eps_rel, float eps_abs, float xnorm_max, uint64_t* __restrict__ out,
                                                                uint32_t pad, uint32_t* __restrict__ flags, const uint32_t* __restrict__ gthr,
                                                                const uint32_t* __restrict__ thr_fin = nullptr) {
	__shared__ uint32_t pre[TC2_MAX_SPLIT + 8];
	__shared__ uint64_t lists[8][32 * LPL];
	__shared__ uint64_t fk[2048];
	__shared__ uint32_t fcnt, red[2][8];
	__shared__ float t_s;
	__shared__ uint32_t w_s, over_s;
	const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	const uint32_t q = blockIdx.x;
	const uint64_t* qbuf = cand + (size_t)q * rsplit * cap;
	const uint32_t* qcnt = cnt + (size_t)q * rsplit;
	if (warp == 0) {
		uint32_t total = 0;
		bool over = false;
		for (uint32_t s0 = 0; s0 < rsplit; s0 += 32) {
			const uint32_t s = s0 + lane;
			uint32_t c = s < rsplit ? qcnt[s] : 0u;
			if (c > cap) { over = true; c = cap; }
			uint32_t incl = c;
#pragma unroll
			for (int o = 1; o < 32; o <<= 1) {
				const uint32_t up = __shfl_up_sync(FULL, incl, o);
				if (lane >= o) incl += up;
			}
			if (s < rsplit) pre[s] = total + incl - c;
			total += __shfl_sync(FULL, incl, 31);
		}
		over = __any_sync(FULL, over);
		if (lane == 0) { pre[rsplit] = total; w_s = 0; over_s = over ? 1u : 0u; fcnt = 0; }
	}
	__syncthreads();
	const uint32_t total = pre[rsplit];
	if (gthr != nullptr) {
		// Fast path. The filter's final shared bound gthr[q] is >= A_k + 2e (it is the smallest local k-th + 2e), so only keys
		// at or below it can matter: usually ~10^3 of the ~10^4 buffered keys of a 148-split launch. They are parked in
		// shared memory (8 independent loads in flight per thread), then the exact k-th smallest a' comes from a bitwise
		// search over the orderable encoding with the keys in registers -- no list insertions.
		const uint32_t gb = gthr[q];
		const uint32_t tid = threadIdx.x;
		// warp w walks splits w, w + 8, ...: four splits (3 x 32 keys each) in flight per round, lanes read neighbouring
		// keys; the counts come from the prefix in shared memory, so no load depends on another
		auto park = [&](uint64_t key) {
			if (key != KEY_EMPTY && (uint32_t)(key >> 32) <= gb) {
				const uint32_t pos = atomicAdd(&fcnt, 1u);
				if (pos < 2048) fk[pos] = key;
			}
		};
		for (uint32_t s0 = warp; s0 < rsplit; s0 += 32) {
			uint64_t kk[4][3];
			uint32_t c[4];
#pragma unroll
			for (int u = 0; u < 4; ++u) {
				const uint32_t sp = s0 + 8 * u;
				c[u] = sp < rsplit ? pre[sp + 1] - pre[sp] : 0u;
#pragma unroll
				for (int j = 0; j < 3; ++j)
					kk[u][j] = (uint32_t)(lane + 32 * j) < c[u] ? __ldg(qbuf + (size_t)sp * cap + lane + 32 * j) : KEY_EMPTY;
			}
#pragma unroll
			for (int u = 0; u < 4; ++u) {
#pragma unroll
				for (int j = 0; j < 3; ++j) park(kk[u][j]);
				for (uint32_t i = 96 + lane; i < c[u]; i += 32) park(__ldg(qbuf + (size_t)(s0 + 8 * u) * cap + i)); // long buffers: rare
			}
		}
		__syncthreads();
		const uint32_t nf = fcnt;
		if (nf <= 2048) {
			uint64_t r[8];
#pragma unroll
			for (int u = 0; u < 8; ++u) r[u] = tid + u * 256 < nf ? fk[tid + u * 256] : KEY_EMPTY;
			float t = __int_as_float(0x7f800000);
			if (nf >= k) {
				uint32_t v = 0;
#pragma unroll 1
				for (int bit = 31; bit >= 0; --bit) {
					const uint32_t trial = v | (1u << bit);
					uint32_t c = 0;
#pragma unroll
					for (int u = 0; u < 8; ++u) c += (uint32_t)(r[u] >> 32) < trial ? 1u : 0u;
					c = __reduce_add_sync(FULL, c);
					if (lane == 0) red[bit & 1][warp] = c;
					__syncthreads(); // one barrier per round: the two halves of red[] alternate
					uint32_t tot = 0;
#pragma unroll
					for (int w = 0; w < 8; ++w) tot += red[bit & 1][w];
					if (tot < k) v = trial;
				}
				const float eps = eps_rel * sqrtf(qnorm2[q]) * xnorm_max + eps_abs;
				t = __fadd_ru(f32_from_orderable(v), 2.f * eps);
			}
			uint64_t* o = out + (size_t)q * pad;
#pragma unroll
			for (int u = 0; u < 8; ++u)
				if (r[u] != KEY_EMPTY && key_dist(r[u]) <= t) {
					const uint32_t pos = atomicAdd(&w_s, 1u);
					if (pos < pad) o[pos] = r[u];
				}
			__syncthreads();
			const uint32_t w = w_s;
			for (uint32_t i = w + tid; i < pad; i += 256) o[i] = KEY_EMPTY;
			if ((over_s || w > pad) && tid == 0) atomicAdd(flags, 1u);
			return;
		}
	}
	const uint32_t per = (total + 7) / 8, i_begin = min(total, warp * per), i_end = min(total, i_begin + per);
	// lane-private cursor into the prefix: flat indices of a lane only grow
	auto fetch = [&](uint32_t i, uint32_t& sp) -> uint64_t {
		if (i >= i_end) return KEY_EMPTY;
		while (pre[sp + 1] <= i) ++sp;
		return __ldg(qbuf + (size_t)sp * cap + (i - pre[sp]));
	};
	uint32_t sp0 = 0;
	{ // split of this warp's first index
		uint32_t lo = 0, hi = rsplit;
		while (hi - lo > 1) { const uint32_t mid = (lo + hi) >> 1; if (pre[mid] <= i_begin) lo = mid; else hi = mid; }
		sp0 = lo;
	}
	WarpList<LPL> best;
	best.init();
	uint64_t tk = KEY_EMPTY;
	uint32_t sp = sp0;
	for (uint32_t i0 = i_begin; i0 < i_end; i0 += 128) {
		uint64_t kk[4];
#pragma unroll
		for (int u = 0; u < 4; ++u) kk[u] = fetch(i0 + u * 32 + lane, sp);
#pragma unroll
		for (int u = 0; u < 4; ++u) {
			unsigned mask = __ballot_sync(FULL, kk[u] < tk);
			while (mask) {
				const int src = __ffs(mask) - 1;
				mask &= mask - 1;
				const uint64_t c = shfl_u64(kk[u], src);
				if (c < tk) { best.insert(c, lane); tk = best.at(k - 1); }
			}
		}
	}
#pragma unroll
	for (int s = 0; s < LPL; ++s) lists[warp][s * 32 + lane] = best.e[s];
	__syncthreads();
	if (warp == 0) {
		for (int w = 1; w < 8; ++w) {
#pragma unroll
			for (int s = 0; s < LPL; ++s) {
				const uint64_t key = lists[w][s * 32 + lane];
				unsigned mask = __ballot_sync(FULL, key < tk);
				while (mask) {
					const int src = __ffs(mask) - 1;
					mask &= mask - 1;
					const uint64_t c = shfl_u64(key, src);
					if (c < tk) { best.insert(c, lane); tk = best.at(k - 1); }
				}
			}
		}
		if (lane == 0) {
			float t = __int_as_float(0x7f800000);
			if (tk != KEY_EMPTY) t = __fadd_ru(key_dist(tk), 2.f * (eps_rel * sqrtf(qnorm2[q]) * xnorm_max + eps_abs));
			t_s = t;
		}
	}
	__syncthreads();
	const float t = t_s;
	uint64_t* o = out + (size_t)q * pad;
	sp = sp0;
	for (uint32_t i0 = i_begin; i0 < i_end; i0 += 128) {
		uint64_t kk[4];
#pragma unroll
		for (int u = 0; u < 4; ++u) kk[u] = fetch(i0 + u * 32 + lane, sp);
#pragma unroll
		for (int u = 0; u < 4; ++u) {
			const bool keep = kk[u] != KEY_EMPTY && key_dist(kk[u]) <= t;
			const unsigned mk = __ballot_sync(FULL, keep);
			if (mk) {
				uint32_t base = 0;
				if (lane == 0) base = atomicAdd(&w_s, (uint32_t)__popc(mk));
				base = __shfl_sync(FULL, base, 0);
				const uint32_t pos = base + __popc(mk & ((1u << lane) - 1));
				if (keep && pos < pad) o[pos] = kk[u];
			}
		}
	}
	__syncthreads();
	const uint32_t w = w_s;
	for (uint32_t i = w + threadIdx.x; i < pad; i += 256) o[i] = KEY_EMPTY;
	bool ver = true;
	if (warp == 0) ver = tc2_bound_verified(thr_fin, q, rsplit, t, lane);
	if ((over_s || w > pad || !ver) && threadIdx.x == 0) atomicAdd(flags, 1u);
}

// exactness of the bf16 single-GEMM mode for THIS query batch, decided on the device (no host round trip): the queries
// must be bf16-exact integers too, no partial sum of the accumulator may leave the 2^24 integer range, and the fp32
// diff-form key of the SIMT scan / K4 must be exact as well (d * (hi - lo)^2 < 2^24, the predicate of the first form,
// tc.cu) so that both paths stay bit-identical; otherwise flags[0] is raised and the caller redoes the batch with the
// SIMT scan. The accumulator runs through sum (-2 q_i) x_i and takes ||x||^2 LAST (the norm step is the tile's final
// MMA): for non-negative data the partial sums lie in [-2 d hi^2, 0] and the total in [-2 d hi^2, ||x||^2_max]; with
// signed values the dot product can have either sign, so the total reaches 2 d amax^2 + ||x||^2_max.
static __global__ void tc_check_exact_kernel(const TcStats* __restrict__ qs, float xmin, float xmax, float d, float xnorm2_max,
                                      uint32_t* __restrict__ flags) {
	const float lo = fminf(xmin, f32_from_orderable(qs->vmin)), hi = fmaxf(xmax, f32_from_orderable(qs->vmax));
	const float amax = fmaxf(fabsf(lo), fabsf(hi));
	const double lim = 16777216.0, dd = (double)d;
	const double acc_max = lo >= 0.f ? fmax(2.0 * dd * amax * amax, (double)xnorm2_max) : 2.0 * dd * amax * amax + (double)xnorm2_max;
	const double span = (double)hi - (double)lo;
	const bool exact = !qs->inexact && amax <= 256.f && acc_max < lim && dd * span * span < lim;
	if (!exact) atomicAdd(flags, 1u);
}
// start of a batch in one launch: query statistics reset, the batch's flags cleared, shared bounds set to +inf
static __global__ void tc_batch_init_kernel(TcStats* s, uint32_t* flags, uint32_t* gthr, uint32_t m) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i < m) gthr[i] = f32_orderable(__int_as_float(0x7f800000));
	if (i < 4) flags[i] = 0u;
	if (i == 0) {
		s->vmin = f32_orderable(3.4e38f);
		s->vmax = f32_orderable(-3.4e38f);
		s->inexact = 0u;
		s->nmax = f32_orderable(0.f);
	}
}

static __global__ void tc_stats_init_kernel(TcStats* s) {
	s->vmin = f32_orderable(3.4e38f);
	s->vmax = f32_orderable(-3.4e38f);
	s->inexact = 0u;
	s->nmax = f32_orderable(0.f);
}

} // namespace expann
