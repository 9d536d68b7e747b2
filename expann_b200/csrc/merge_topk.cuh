// K7 — merge sorted/unsorted lists of 64-bit (key,id) entries into the k smallest per query, and decode keys
// into the reference's output form (ids as size_t nearest-first, src/brute_force_engine.h:39-45).
// Used (a) to merge the row-split partials of one GPU, (b) to merge the per-GPU lists gathered over NCCL —
// because keys carry (distance, id) the result is bit-identical for any partitioning.
#pragma once
#include "common.cuh"

namespace expann {

struct MergeParams {
	const uint64_t* in; // list s of query q starts at in + q*stride_q + s*stride_s, klist keys each
	uint64_t stride_q, stride_s;
	uint32_t nlists, klist;
	uint32_t m, k;
	uint32_t id_add;     // added to the 32-bit id of every non-empty key written to out_keys (shard offset)
	uint64_t* out_keys;  // [m][k] or null
	const uint32_t* gate; // device-side gate (a redo queued behind a flag): when set, the launch does nothing unless *gate != 0
};

template <int LPL> __global__ void __launch_bounds__(128) merge_topk_kernel(const MergeParams p) {
	const int lane = threadIdx.x & 31;
	const uint32_t q = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
	if (q >= p.m) return;
	if (p.gate != nullptr && *reinterpret_cast<const volatile uint32_t*>(p.gate) == 0u) return;
	WarpList<LPL> out;
	out.init();
	uint64_t tk = KEY_EMPTY;
	const uint64_t* base = p.in + (size_t)q * p.stride_q;
	const uint32_t total = p.nlists * p.klist;
	for (uint32_t i0 = 0; i0 < total; i0 += 32) {
		const uint32_t i = i0 + lane;
		uint64_t key = KEY_EMPTY;
		if (i < total) {
			const uint32_t s = i / p.klist, j = i - s * p.klist;
			key = base[(size_t)s * p.stride_s + j];
		}
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
#pragma unroll
	for (int s = 0; s < LPL; ++s) {
		const uint32_t j = s * 32 + lane;
		if (j < p.k) {
			uint64_t key = out.e[s];
			if (key != KEY_EMPTY) key += p.id_add;
			p.out_keys[(size_t)q * p.k + j] = key;
		}
	}
}


// Same merge with one 256-thread block per query: 8 warps each take every 8th 32-key slice, their lists meet in shared
// memory and warp 0 merges those. For a few queries with many row splits (small-batch scans: 8 queries x 148 lists) the
// one-warp kernel above is a chain of dependent L2 round trips (51 us measured for 218 lists).
template <int LPL> __global__ void __launch_bounds__(256) merge_topk_block_kernel(const MergeParams p) {
	__shared__ uint64_t sm[8][32 * LPL];
	const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	const uint32_t q = blockIdx.x;
	if (p.gate != nullptr && *reinterpret_cast<const volatile uint32_t*>(p.gate) == 0u) return; // uniform over the block
	WarpList<LPL> out;
	out.init();
	uint64_t tk = KEY_EMPTY;
	const uint64_t* base = p.in + (size_t)q * p.stride_q;
	const uint32_t total = p.nlists * p.klist;
	for (uint32_t i0 = warp * 32; i0 < total; i0 += 256) {
		const uint32_t i = i0 + lane;
		uint64_t key = KEY_EMPTY;
		if (i < total) {
			const uint32_t s = i / p.klist, j = i - s * p.klist;
			key = base[(size_t)s * p.stride_s + j];
		}
		unsigned mask = __ballot_sync(FULL, key < tk);
		while (mask) {
			const int src = __ffs(mask) - 1;
			mask &= mask - 1;
			const uint64_t kk = shfl_u64(key, src);
			if (kk < tk) { out.insert(kk, lane); tk = out.at(p.k - 1); }
		}
	}
#pragma unroll
	for (int s = 0; s < LPL; ++s) sm[warp][s * 32 + lane] = out.e[s];
	__syncthreads();
	if (warp != 0) return;
	for (int w = 1; w < 8; ++w) {
#pragma unroll
		for (int s = 0; s < LPL; ++s) {
			const uint64_t key = sm[w][s * 32 + lane];
			unsigned mask = __ballot_sync(FULL, key < tk);
			while (mask) {
				const int src = __ffs(mask) - 1;
				mask &= mask - 1;
				const uint64_t kk = shfl_u64(key, src);
				if (kk < tk) { out.insert(kk, lane); tk = out.at(p.k - 1); }
			}
		}
	}
#pragma unroll
	for (int s = 0; s < LPL; ++s) {
		const uint32_t j = s * 32 + lane;
		if (j < p.k) {
			uint64_t key = out.e[s];
			if (key != KEY_EMPTY) key += p.id_add;
			p.out_keys[(size_t)q * p.k + j] = key;
		}
	}
}

// keys -> (ids, dists, counts). kind 0: fp32 keys (dist = decoded float), 1: int32 scores (dist = (float)score)
__global__ void decode_keys_kernel(const uint64_t* keys, uint32_t m, uint32_t k, uint64_t id_offset, int kind,
                                   int negate, uint64_t* ids, float* dists, uint64_t* counts) {
	const uint32_t q = blockIdx.x * blockDim.x + threadIdx.x;
	if (q >= m) return;
	uint32_t c = 0;
	for (uint32_t j = 0; j < k; ++j) {
		const uint64_t key = keys[(size_t)q * k + j];
		const bool ok = key != KEY_EMPTY;
		c += ok;
		ids[(size_t)q * k + j] = ok ? (uint64_t)key_id(key) + id_offset : 0xFFFFFFFFFFFFFFFFull;
		if (dists) {
			float d = __int_as_float(0x7f800000);
			if (ok) d = kind == 0 ? key_dist(key) : (float)key_score_i32(key);
			if (ok && negate) d = -d;
			dists[(size_t)q * k + j] = d;
		}
	}
	if (counts) counts[q] = c;
}

} // namespace expann
