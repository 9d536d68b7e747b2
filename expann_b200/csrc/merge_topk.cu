// Host side of K7 (merge) and the key decoder.
#include "engine.h"
#include "merge_topk.cuh"

namespace expann {

int launch_merge(expann_gpu* h, const uint64_t* in, size_t m, size_t nlists, size_t klist, size_t stride_q,
                 size_t stride_s, size_t k, uint32_t id_add, uint64_t* out, cudaStream_t st, const uint32_t* gate) {
	if (m == 0) return EXPANN_OK;
	if (k == 0 || k > 128) return fail(h, EXPANN_ERR_UNSUPPORTED, "merge: k must be in 1..128");
	MergeParams p{};
	p.in = in;
	p.stride_q = stride_q;
	p.stride_s = stride_s;
	p.nlists = (uint32_t)nlists;
	p.klist = (uint32_t)klist;
	p.m = (uint32_t)m;
	p.k = (uint32_t)k;
	p.id_add = id_add;
	p.out_keys = out;
	p.gate = gate;
	const int wpb = 4;
	dim3 grid((unsigned)((m + wpb - 1) / wpb));
	if (nlists * klist >= 512 && m <= 4096) { // few queries, many lists: one block per query
		if (k <= 32) merge_topk_block_kernel<1><<<(unsigned)m, 256, 0, st>>>(p);
		else merge_topk_block_kernel<4><<<(unsigned)m, 256, 0, st>>>(p);
	} else if (k <= 32) merge_topk_kernel<1><<<grid, wpb * 32, 0, st>>>(p);
	else merge_topk_kernel<4><<<grid, wpb * 32, 0, st>>>(p);
	EXPANN_CUDA(h, cudaGetLastError());
	h->stats.merge_launches++;
	return EXPANN_OK;
}

int launch_decode(expann_gpu* h, const uint64_t* keys, size_t m, size_t k, int kind, uint64_t* ids, float* dists,
                  uint64_t* counts, cudaStream_t st) {
	if (m == 0) return EXPANN_OK;
	decode_keys_kernel<<<(unsigned)((m + 127) / 128), 128, 0, st>>>(keys, (uint32_t)m, (uint32_t)k, 0ull, kind, 0, ids,
	                                                                  dists, counts);
	EXPANN_CUDA(h, cudaGetLastError());
	h->stats.aux_launches++;
	return EXPANN_OK;
}

} // namespace expann
