// Operand format 4 of K3: tcgen05.mma.kind::tf32 on the fp32 rows themselves (pair kernel, tc_filter3.cuh KIND = 2).
// The format between one bf16 term per value (error 2^-6 |q| max|x|) and the two-term bf16 split (three times the
// flops): the tensor core reads the top 19 bits of every fp32 value (10 mantissa bits; truncation, relative error
// <= 2^-10), the queries are rounded to tf32 beforehand (2^-11), so a' is off by at most 2 (2^-10 + 2^-11) |q| max|x| —
// a slack 5x tighter than format 3 — at twice its MMA time (K = 8 per instruction instead of 16) and with NO operand copy:
// the TMA loads the engine's own fp32 base rows ([128 rows][32 floats] boxes, 128-byte swizzle). This file holds the
// kernel instantiations; the host logic is tc.cu's (tc_batch2).
#include "engine.h"
#include "tc_filter3.cuh"

namespace expann {

template <int NSTAGE, bool ASTREAM>
static int launch_pair_tf32(expann_gpu* h, const CUtensorMap& tq, const Tc2Params& p, dim3 grid, cudaStream_t st) {
	const size_t smem = Tc3Cfg::smem_bytes(p.kd, NSTAGE, ASTREAM);
	if (int rc = ensure_dyn_smem(h, (const void*)tc_filter_pair_kernel<NSTAGE, ASTREAM, 2>, smem)) return rc;
	tc_filter_pair_kernel<NSTAGE, ASTREAM, 2><<<grid, Tc3Cfg::THREADS, smem, st>>>(tq, h->tmap_x32, h->tmap_xx32, p); // __cluster_dims__(2,1,1)
	EXPANN_CUDA(h, cudaGetLastError());
	return EXPANN_OK;
}

int launch_filter_pair_tf32(expann_gpu* h, const CUtensorMap& tq, const Tc2Params& p, dim3 grid, cudaStream_t st) {
	if (Tc3Cfg::smem_bytes(p.kd, 4) > 227 * 1024) return launch_pair_tf32<5, true>(h, tq, p, grid, st); // query operand through the ring
	if (Tc3Cfg::smem_bytes(p.kd, 6) <= 227 * 1024) return launch_pair_tf32<6, false>(h, tq, p, grid, st);
	return launch_pair_tf32<4, false>(h, tq, p, grid, st);
}

} // namespace expann
