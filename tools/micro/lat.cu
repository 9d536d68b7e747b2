// pointer-chase latency (cycles at the running SM clock) for a working set in L2 / in DRAM, and a 32-lane 16-byte gather
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <vector>
#include <algorithm>
#include <random>
__global__ void chase(const uint32_t* __restrict__ next, int steps, uint32_t* out, long long* cyc) {
	uint32_t i = 0;
	for (int s = 0; s < 64; ++s) i = next[i]; // warm a little
	long long t0 = clock64();
	for (int s = 0; s < steps; ++s) i = next[i];
	long long t1 = clock64();
	*out = i; *cyc = t1 - t0;
}
__global__ void chase_ldg(const uint32_t* __restrict__ next, int steps, uint32_t* out, long long* cyc) {
	uint32_t i = 0;
	long long t0 = clock64();
	for (int s = 0; s < steps; ++s) i = __ldg(next + i);
	long long t1 = clock64();
	*out = i; *cyc = t1 - t0;
}
// a warp reads 4 x 16 B per lane from a random 16 KB-aligned block, dependent on the previous block's first word
__global__ void chase_block(const uint4* __restrict__ blocks, int steps, uint32_t nblk, uint32_t* out, long long* cyc) {
	uint32_t b = 0; float acc = 0.f;
	const int lane = threadIdx.x;
	long long t0 = clock64();
	for (int s = 0; s < steps; ++s) {
		const uint4* p = blocks + (size_t)b * 1024 + lane;
		uint4 x0 = __ldg(p), x1 = __ldg(p + 32), x2 = __ldg(p + 64), x3 = __ldg(p + 96);
		acc += __uint_as_float(x1.y) + __uint_as_float(x2.z) + __uint_as_float(x3.w);
		b = __shfl_sync(0xffffffffu, x0.x, 0) % nblk;
	}
	long long t1 = clock64();
	if (lane == 0) { *out = b + (acc == 1.234f); *cyc = t1 - t0; }
}
int main() {
	uint32_t* d_out; long long* d_c; cudaMalloc(&d_out, 4); cudaMalloc(&d_c, 8);
	for (size_t mb : {16, 64, 512, 4096}) {
		const size_t n = mb * 1024 * 1024 / 4, stride = 64; // one word per 256 B
		const size_t m = n / stride;
		std::vector<uint32_t> perm(m); for (size_t i = 0; i < m; ++i) perm[i] = (uint32_t)i;
		std::mt19937 g(1); std::shuffle(perm.begin() + 1, perm.end(), g);
		std::vector<uint32_t> h(n, 0);
		for (size_t i = 0; i + 1 < m; ++i) h[(size_t)perm[i] * stride] = perm[i + 1] * (uint32_t)stride;
		h[(size_t)perm[m - 1] * stride] = 0;
		uint32_t* d; cudaMalloc(&d, n * 4); cudaMemcpy(d, h.data(), n * 4, cudaMemcpyHostToDevice);
		const int steps = 20000; long long c;
		for (int rep = 0; rep < 2; ++rep) { chase<<<1, 1>>>(d, steps, d_out, d_c); cudaMemcpy(&c, d_c, 8, cudaMemcpyDeviceToHost); }
		printf("chase   %5zu MB: %.1f cycles per dependent load\n", mb, (double)c / steps);
		chase_ldg<<<1, 1>>>(d, steps, d_out, d_c); cudaMemcpy(&c, d_c, 8, cudaMemcpyDeviceToHost);
		printf("ldg     %5zu MB: %.1f cycles per dependent load\n", mb, (double)c / steps);
		// blocks: fill first word of every 16 KB block with a random block number
		const uint32_t nblk = (uint32_t)(mb * 1024 / 16);
		std::vector<uint32_t> hb(n); for (size_t i = 0; i < n; ++i) hb[i] = (uint32_t)g();
		cudaMemcpy(d, hb.data(), n * 4, cudaMemcpyHostToDevice);
		for (int rep = 0; rep < 2; ++rep) { chase_block<<<1, 32>>>((const uint4*)d, 5000, nblk, d_out, d_c); cudaMemcpy(&c, d_c, 8, cudaMemcpyDeviceToHost); }
		printf("block   %5zu MB: %.1f cycles per dependent 2 KB warp gather\n", mb, (double)c / 5000);
		cudaFree(d);
	}
	int clk; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0); printf("clock rate attr %d kHz\n", clk);
	return 0;
}
