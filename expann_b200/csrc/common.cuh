// Shared device helpers for the expANN B200 query path: ordered 64-bit (key,id) encoding, the warp-resident
// top-k list, mbarrier / TMA wrappers. sm_100a only.
//
// Ordering contract (reference: src/brute_force_engine.h:30-38 and src/topk_t.h:24-35): results are the k
// smallest elements under the lexicographic order of std::pair<T,size_t>, i.e. (distance asc, id asc).
// A single unsigned 64-bit compare on  (orderable(dist) << 32) | id  implements exactly that order, on every
// lane, warp, CTA and GPU, so merges are associative and the result does not depend on the partitioning.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace expann {

constexpr uint64_t KEY_EMPTY = 0xFFFFFFFFFFFFFFFFull; // sorts after every real key (incl. +inf distances)
constexpr unsigned FULL = 0xFFFFFFFFu;

// monotone map float -> uint32 (total order, -0 < +0, NaN with sign bit clear sorts above +inf)
__host__ __device__ __forceinline__ uint32_t f32_orderable(float f) {
#ifdef __CUDA_ARCH__
	uint32_t u = __float_as_uint(f);
#else
	union { float f; uint32_t u; } c; c.f = f; uint32_t u = c.u;
#endif
	return u ^ ((u >> 31) ? 0xFFFFFFFFu : 0x80000000u);
}
__host__ __device__ __forceinline__ float f32_from_orderable(uint32_t o) {
	uint32_t u = o ^ ((o >> 31) ? 0x80000000u : 0xFFFFFFFFu);
#ifdef __CUDA_ARCH__
	return __uint_as_float(u);
#else
	union { float f; uint32_t u; } c; c.u = u; return c.f;
#endif
}
__host__ __device__ __forceinline__ uint64_t make_key(float d, uint32_t id) {
	return ((uint64_t)f32_orderable(d) << 32) | id;
}
// integer scores (compressed scans): order-preserving bias so that int32 compares as unsigned
__host__ __device__ __forceinline__ uint64_t make_key_i32(int32_t s, uint32_t id) {
	return ((uint64_t)((uint32_t)s ^ 0x80000000u) << 32) | id;
}
__host__ __device__ __forceinline__ int32_t key_score_i32(uint64_t k) {
	return (int32_t)((uint32_t)(k >> 32) ^ 0x80000000u);
}
__host__ __device__ __forceinline__ float key_dist(uint64_t k) { return f32_from_orderable((uint32_t)(k >> 32)); }
__host__ __device__ __forceinline__ uint32_t key_id(uint64_t k) { return (uint32_t)k; }

#ifdef __CUDACC__
__device__ __forceinline__ uint64_t shfl_u64(uint64_t v, int src) {
	uint32_t lo = __shfl_sync(FULL, (uint32_t)v, src), hi = __shfl_sync(FULL, (uint32_t)(v >> 32), src);
	return ((uint64_t)hi << 32) | lo;
}
__device__ __forceinline__ uint64_t shfl_up_u64(uint64_t v, int delta) {
	uint32_t lo = __shfl_up_sync(FULL, (uint32_t)v, delta), hi = __shfl_up_sync(FULL, (uint32_t)(v >> 32), delta);
	return ((uint64_t)hi << 32) | lo;
}

// ---------------------------------------------------------------------------------------------------------
// Warp-resident sorted list of the K = 32*LPL smallest keys seen so far (ascending). Element i lives in
// lane i%32, slot i/32. All 32 lanes call every member with warp-uniform arguments.
// This is the device form of topk_t<T> (src/topk_t.h:9-67): `consider` == insert(), `worst_val` == the
// threshold at position k-1, `to_combined_vector` == the list itself (already ascending). The dedupe set of
// topk_t is structural here: a scan visits every id once, the graph search keeps its own visited set.
// ---------------------------------------------------------------------------------------------------------
template <int LPL> struct WarpList {
	uint64_t e[LPL];
	__device__ __forceinline__ void init() {
#pragma unroll
		for (int s = 0; s < LPL; ++s) e[s] = KEY_EMPTY;
	}
	// key at sorted position p (warp-uniform p)
	__device__ __forceinline__ uint64_t at(int p) const {
		uint64_t r = KEY_EMPTY;
#pragma unroll
		for (int s = 0; s < LPL; ++s) {
			uint64_t v = shfl_u64(e[s], p & 31);
			if ((p >> 5) == s) r = v;
		}
		return r;
	}
	// insert key (warp-uniform) keeping the list sorted; the largest element falls off the end.
	__device__ __forceinline__ void insert(uint64_t key, int lane) {
		int pos = 0;
#pragma unroll
		for (int s = 0; s < LPL; ++s) pos += __popc(__ballot_sync(FULL, e[s] < key));
		if (pos >= 32 * LPL) return;
#pragma unroll
		for (int s = LPL - 1; s >= 0; --s) {
			uint64_t up = shfl_up_u64(e[s], 1);                       // element i-1 for lanes >= 1
			uint64_t carry = s > 0 ? shfl_u64(e[s - 1], 31) : KEY_EMPTY; // element i-1 for lane 0
			if (lane == 0) up = carry;
			int i = s * 32 + lane;
			if (i == pos) e[s] = key;
			else if (i > pos) e[s] = up;
		}
	}
};
template <int LPL> __device__ __forceinline__ uint64_t warplist_threshold(const WarpList<LPL>& l, int k) {
	return l.at(k - 1);
}

// ---------------------------------------------------------------------------------------------------------
// mbarrier + TMA (cp.async.bulk.tensor) wrappers — the PTX the CUTLASS helpers would emit
// ---------------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
	asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_fence_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
	asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
	asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
	asm volatile(
	    "{\n\t.reg .pred p;\n\t"
	    "WAIT_%=:\n\t"
	    "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
	    "@p bra DONE_%=;\n\t"
	    "bra WAIT_%=;\n\t"
	    "DONE_%=:\n\t}" ::"r"(smem_u32(bar)),
	    "r"(parity)
	    : "memory");
}
// 2D tiled TMA load: box at (c0 = innermost coordinate, c1 = row coordinate) -> smem, completes on `bar`
__device__ __forceinline__ void tma_load_2d(void* smem_dst, const void* tmap, int c0, int c1, uint64_t* bar) {
	asm volatile(
	    "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::
	        "r"(smem_u32(smem_dst)),
	    "l"(tmap), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
	    : "memory");
}
__device__ __forceinline__ void tma_prefetch_desc(const void* tmap) {
	asm volatile("prefetch.tensormap [%0];" ::"l"(tmap) : "memory");
}
#endif // __CUDACC__

} // namespace expann
