// Host side of K5: graph hand-off (the reference's index file, or CSR arrays), HBM layout, launch.
#include "engine.h"
#include "graph_search.cuh"
#include "graph_inline.cuh"
#include <algorithm>
#include <cstdio>
#include <cstring>
#include <stdexcept>

namespace expann {

__global__ void quantize_u8_kernel(const float* __restrict__ base, uint8_t* __restrict__ codes, size_t total) {
	// quantizer_simple<uint8_t>::build (src/quantizer.h:132-141): code = uint8_t(x), i.e. truncation on [0,256)
	size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	const size_t stride = (size_t)gridDim.x * blockDim.x;
	for (; i < total; i += stride) codes[i] = (uint8_t)(uint32_t)base[i];
}

// up_vec[c] = base[meta[c].z]: one warp per row
__global__ void gather_upper_rows_kernel(const float* __restrict__ base, const uint4* __restrict__ meta, float* __restrict__ out,
                                         uint32_t nup, uint32_t d_pad) {
	const uint32_t w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
	if (w >= nup) return;
	const float4* src = reinterpret_cast<const float4*>(base + (size_t)meta[w].z * d_pad);
	float4* dst = reinterpret_cast<float4*>(out + (size_t)w * d_pad);
	for (uint32_t i = lane; i < d_pad / 4; i += 32) dst[i] = src[i];
}

// CSR over (vertex, layer) pairs in vertex-major order -> fixed-stride device layout
int graph_set_csr(expann_gpu* h, size_t n, const uint32_t* nlayers, const uint64_t* offsets, const uint32_t* edges,
                  uint64_t start, uint64_t max_layer) {
	HostGraph& g = h->graph;
	if (n == 0 || start >= n) return fail(h, EXPANN_ERR_ARG, "graph: empty graph or bad starting vertex");
	size_t maxb = 0, maxu = 0, urows = 0, li = 0;
	for (size_t v = 0; v < n; ++v) {
		if (nlayers[v] == 0) return fail(h, EXPANN_ERR_ARG, "graph: vertex without a bottom layer");
		for (uint32_t l = 0; l < nlayers[v]; ++l, ++li) {
			const size_t deg = (size_t)(offsets[li + 1] - offsets[li]);
			if (l == 0) maxb = std::max(maxb, deg); else maxu = std::max(maxu, deg);
		}
		urows += nlayers[v] - 1;
	}
	if (maxb > GS_MAXNB || maxu > GS_MAXNB) return fail(h, EXPANN_ERR_UNSUPPORTED, "graph: vertex degree above 256");
	g.n = n;
	g.m0s = std::max<size_t>(4, (maxb + 3) / 4 * 4);
	g.ms = std::max<size_t>(4, (maxu + 3) / 4 * 4);
	g.start = (uint32_t)start;
	g.max_layer = (uint32_t)max_layer;
	g.bottom.assign(n * g.m0s, NO_VERTEX);
	g.upper.assign(std::max<size_t>(1, urows) * g.ms, NO_VERTEX);
	g.upper_idx.assign(n, NO_VERTEX);
	g.upper_nl.assign(n, 1);
	size_t urow = 0;
	li = 0;
	for (size_t v = 0; v < n; ++v) {
		g.upper_nl[v] = nlayers[v];
		if (nlayers[v] > 1) g.upper_idx[v] = (uint32_t)urow;
		for (uint32_t l = 0; l < nlayers[v]; ++l, ++li) {
			const uint32_t* e = edges + offsets[li];
			const size_t deg = (size_t)(offsets[li + 1] - offsets[li]);
			uint32_t* dst = l == 0 ? &g.bottom[v * g.m0s] : &g.upper[(urow + l - 1) * g.ms];
			for (size_t j = 0; j < deg; ++j) {
				if (e[j] >= n) return fail(h, EXPANN_ERR_ARG, "graph: edge to a vertex out of range");
				dst[j] = e[j];
			}
		}
		urow += nlayers[v] - 1;
	}
	g.present = true;
	h->built = false;
	return EXPANN_OK;
}

// reference index file, src/antitopo_engine.h:931-991 (writer) / :994-1074 (reader): raw little-endian dump.
// Nothing in the file is trusted: every count is checked against the bytes that are left before anything is allocated,
// ids are range-checked in 64 bits, and no exception crosses the C ABI (a truncated or corrupt file is EXPANN_ERR_IO).
static int graph_load_index_impl(expann_gpu* h, const char* path, std::vector<float>& vectors, size_t& n_out, size_t& d_out) {
	FILE* f = std::fopen(path, "rb");
	if (!f) return fail(h, EXPANN_ERR_IO, std::string("cannot open index file ") + path);
	struct Closer { FILE* f; ~Closer() { std::fclose(f); } } closer{ f };
	const auto bad = [&](const char* why) { return fail(h, EXPANN_ERR_IO, std::string("malformed index file ") + path + ": " + why); };
	uint64_t fsize = 0, pos = 0;
	if (std::fseek(f, 0, SEEK_END) != 0) return bad("not seekable");
	{ const long long e = (long long)ftello(f); if (e < 0) return bad("not seekable"); fsize = (uint64_t)e; }
	std::rewind(f);
	auto rd = [&](void* p, uint64_t sz) {
		if (sz > fsize - pos || std::fread(p, 1, (size_t)sz, f) != sz) return false;
		pos += sz;
		return true;
	};
	uint64_t start, M, M0, u64, max_layer, n = 0;
	uint8_t b;
	float fl;
	bool ok = rd(&start, 8) && rd(&M, 8) && rd(&M0, 8) && rd(&u64, 8) /* ef_search_mult */ && rd(&b, 1);
	if (ok && b) ok = rd(&u64, 8);                       // optional ef_search
	ok = ok && rd(&u64, 8) && rd(&u64, 8)                // ef_construction, ortho_count
	     && rd(&fl, 4) && rd(&fl, 4) && rd(&u64, 8)      // ortho_factor, ortho_bias, prune_overflow
	     && rd(&b, 1) && rd(&b, 1)                       // use_compression, use_largest_direction_filtering
	     && rd(&max_layer, 8) && rd(&n, 8);
	if (!ok) return bad("truncated header");
	if (n == 0) return bad("no vectors");
	// every vector record is 8 + 4 d bytes, every vertex record at least 8: n is bounded by what is left in the file
	if (n > (fsize - pos) / 8) return bad("vector count exceeds the file size");
	if (n >= 0xFFFFFFFFull) return fail(h, EXPANN_ERR_UNSUPPORTED, "index file: vertex ids must fit 32 bits on device");
	if (start >= n) return bad("starting vertex out of range");
	if (max_layer > 64) return bad("implausible max_layer");
	uint64_t d = 0;
	for (uint64_t i = 0; i < n; ++i) {
		uint64_t sz;
		if (!rd(&sz, 8)) return bad("truncated vector section");
		if (i == 0) {
			d = sz;
			if (d == 0 || d > (1u << 20)) return bad("implausible dimension");
			if (n > (fsize - pos + 8) / (8 + 4 * d)) return bad("vector section exceeds the file size");
			vectors.resize((size_t)(n * d)); // n * (8 + 4 d) <= file size: no overflow
		}
		if (sz != d) return bad("vectors of different dimensions");
		if (!rd(vectors.data() + i * d, 4 * d)) return bad("truncated vector section");
	}
	uint64_t nv = 0;
	if (!rd(&nv, 8) || nv != n) return bad("vertex count differs from vector count");
	if (n > (fsize - pos) / 8) return bad("vertex section exceeds the file size");
	std::vector<uint32_t> nlayers((size_t)n), edges;
	std::vector<uint64_t> offsets(1, 0);
	std::vector<uint8_t> rec;
	edges.reserve((size_t)std::min<uint64_t>((fsize - pos) / 12, (uint64_t)1 << 32));
	for (uint64_t v = 0; v < n; ++v) {
		uint64_t nl;
		if (!rd(&nl, 8)) return bad("truncated vertex section");
		if (nl == 0 || nl > max_layer + 1 || nl > 64) return bad("vertex with an implausible layer count");
		nlayers[v] = (uint32_t)nl;
		for (uint64_t l = 0; l < nl; ++l) {
			uint64_t deg;
			if (!rd(&deg, 8)) return bad("truncated adjacency list");
			if (deg > (fsize - pos) / 12) return bad("adjacency list exceeds the file size");
			rec.resize((size_t)deg * 12); // {float len; size_t id} packed as written field by field
			if (!rd(rec.data(), deg * 12)) return bad("truncated adjacency list");
			for (uint64_t e = 0; e < deg; ++e) {
				uint64_t id;
				std::memcpy(&id, rec.data() + e * 12 + 4, 8);
				if (id >= n) return bad("edge to a vertex out of range");
				edges.push_back((uint32_t)id);
			}
			offsets.push_back(edges.size());
		}
	}
	n_out = (size_t)n;
	d_out = (size_t)d;
	h->cfg.M = M;
	h->cfg.M0 = M0;
	return graph_set_csr(h, (size_t)n, nlayers.data(), offsets.data(), edges.data(), start, max_layer);
}

int graph_load_index(expann_gpu* h, const char* path, std::vector<float>& vectors, size_t& n_out, size_t& d_out) {
	try {
		return graph_load_index_impl(h, path, vectors, n_out, d_out);
	} catch (const std::exception& e) { // bad_alloc / length_error must not cross the extern "C" boundary
		return fail(h, EXPANN_ERR_IO, std::string("index file ") + path + ": " + e.what());
	}
}

// shared memory of graph_search_inline_kernel next to its visited bitmap
static size_t inline_fixed_smem(size_t d_pad) { return d_pad * 5 + 2 * GS_MAXNB * 8 + 256 * 8 + 32 * 8 + 6 * 16 + 6 * 8 + 128; }

int graph_finalize(expann_gpu* h) {
	HostGraph& g = h->graph;
	if (!g.present) return fail(h, EXPANN_ERR_STATE, "graph engine: no graph loaded (expann_gpu_load_index / expann_gpu_set_graph)");
	if (g.n != h->n) return fail(h, EXPANN_ERR_STATE, "graph engine: graph and stored vectors disagree on n");
	cudaStream_t st = h->stream;
	EXPANN_CUDA(h, h->g_bottom.ensure(g.bottom.size() * 4));
	EXPANN_CUDA(h, h->g_upper.ensure(g.upper.size() * 4));
	EXPANN_CUDA(h, h->g_upper_idx.ensure(g.n * 4));
	EXPANN_CUDA(h, h->g_upper_nl.ensure(g.n * 4));
	EXPANN_CUDA(h, cudaMemcpyAsync(h->g_bottom.p, g.bottom.data(), g.bottom.size() * 4, cudaMemcpyHostToDevice, st));
	EXPANN_CUDA(h, cudaMemcpyAsync(h->g_upper.p, g.upper.data(), g.upper.size() * 4, cudaMemcpyHostToDevice, st));
	EXPANN_CUDA(h, cudaMemcpyAsync(h->g_upper_idx.p, g.upper_idx.data(), g.n * 4, cudaMemcpyHostToDevice, st));
	EXPANN_CUDA(h, cudaMemcpyAsync(h->g_upper_nl.p, g.upper_nl.data(), g.n * 4, cudaMemcpyHostToDevice, st));
	// resident warp slots: every SM holds GS_CTAS_PER_SM CTAs of GS_WARPS warps; one visited byte-array per slot
	int sms = 148;
	cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, h->device);
	h->g_ctas = (uint32_t)sms * (uint32_t)h->opt_graph_occ; // resident CTAs (GS_LAUNCH picks the matching register cap)
	const size_t slots = (size_t)h->g_ctas * GS_WARPS;
	h->g_vis_stride = (g.n + 15) / 16 * 16;
	EXPANN_CUDA(h, h->g_visited.ensure(slots * h->g_vis_stride));
	EXPANN_CUDA(h, cudaMemsetAsync(h->g_visited.p, 0, slots * h->g_vis_stride, st));
	EXPANN_CUDA(h, h->g_epoch.ensure(slots * 4 + 8));
	EXPANN_CUDA(h, cudaMemsetAsync(h->g_epoch.p, 0, slots * 4 + 8, st));
	// compact form of the upper layers for the latency form of the search (GraphParams::up_meta)
	h->g_nup = 0;
	if (h->opt_graph_upper_compact && g.max_layer > 1) {
		std::vector<uint32_t> cidx(g.n, NO_VERTEX);
		std::vector<uint4> meta;
		for (size_t v = 0; v < g.n; ++v)
			if (g.upper_idx[v] != NO_VERTEX) {
				cidx[v] = (uint32_t)meta.size();
				meta.push_back(make_uint4(g.upper_idx[v], g.upper_nl[v], (uint32_t)v, 0));
			}
		std::vector<uint32_t> uc(g.upper.size());
		bool ok = !meta.empty() && meta.size() * h->d_pad * 4 <= ((size_t)2 << 30);
		for (size_t i = 0; ok && i < uc.size(); ++i) {
			uc[i] = g.upper[i] == NO_VERTEX ? NO_VERTEX : cidx[g.upper[i]];
			if (g.upper[i] != NO_VERTEX && uc[i] == NO_VERTEX) ok = false; // a neighbour without upper layers: keep the plain form
		}
		if (ok) {
			EXPANN_CUDA(h, h->g_up_meta.ensure(meta.size() * 16));
			EXPANN_CUDA(h, h->g_upper_c.ensure(uc.size() * 4));
			EXPANN_CUDA(h, h->g_up_vec.ensure(meta.size() * h->d_pad * 4));
			EXPANN_CUDA(h, cudaMemcpyAsync(h->g_up_meta.p, meta.data(), meta.size() * 16, cudaMemcpyHostToDevice, st));
			EXPANN_CUDA(h, cudaMemcpyAsync(h->g_upper_c.p, uc.data(), uc.size() * 4, cudaMemcpyHostToDevice, st));
			gather_upper_rows_kernel<<<(unsigned)((meta.size() * 32 + 255) / 256), 256, 0, st>>>(
			    h->base, h->g_up_meta.as<uint4>(), h->g_up_vec.as<float>(), (uint32_t)meta.size(), (uint32_t)h->d_pad);
			EXPANN_CUDA(h, cudaGetLastError());
			EXPANN_CUDA(h, cudaStreamSynchronize(st)); // meta / uc are host temporaries
			h->stats.aux_launches++;
			h->g_nup = meta.size();
		}
	}
	if (h->cfg.use_compression) {
		const size_t total = h->n * h->d_pad;
		EXPANN_CUDA(h, h->g_codes.ensure(total));
		quantize_u8_kernel<<<(unsigned)std::min<size_t>(65535, (total + 255) / 256), 256, 0, st>>>(h->base, h->g_codes.as<uint8_t>(), total);
		EXPANN_CUDA(h, cudaGetLastError());
		h->stats.aux_launches++;
	}
	// neighbourhood blocks for the latency form (graph_inline.cuh): HBM capacity traded for one memory trip per hop
	h->g_inline = false;
	if (h->opt_graph_inline != 0 && (h->g_nup || g.max_layer <= 1)) {
		const bool comp = h->cfg.use_compression != 0; // compressed graphs: the bottom blocks hold the 8-bit codes
		const size_t bb_idb = (g.m0s * 4 + 127) / 128 * 128, bb = comp ? (bb_idb + g.m0s * h->d_pad + 127) / 128 * 128 : bb_idb + g.m0s * h->d_pad * 4;
		const size_t ub_idb = (g.ms * 4 + 127) / 128 * 128, ub = (ub_idb + g.ms * 16 + g.ms * h->d_pad * 4 + 127) / 128 * 128;
		const size_t urows = h->g_nup ? g.upper.size() / g.ms : 0;
		const size_t total = g.n * bb + urows * ub;
		size_t freeb = 0, totalb = 0;
		EXPANN_CUDA(h, cudaMemGetInfo(&freeb, &totalb));
		freeb += h->g_blk_bottom.cap + h->g_blk_upper.cap; // a rebuild reuses what the handle already holds
		const bool fits_smem = inline_fixed_smem(h->d_pad) <= 200 * 1024;
		const bool fits_hbm = h->opt_graph_inline >= 1 ? total + ((size_t)1 << 30) <= freeb : total <= freeb / 3;
		if (fits_smem && fits_hbm && bb < ((size_t)1 << 31) && ub < ((size_t)1 << 31)) {
			bool got = true;
			for (auto pr : { std::make_pair(&h->g_blk_bottom, g.n * bb), std::make_pair(&h->g_blk_upper, std::max<size_t>(256, urows * ub)) })
				if (got && pr.first->cap < pr.second) { // exact size: these are the largest allocations of the handle
					pr.first->release();
					const cudaError_t e = cudaMalloc(&pr.first->p, pr.second);
					if (e != cudaSuccess) { // (several engines building on one GPU at the same time may all have seen the same free memory)
						(void)cudaGetLastError();
						pr.first->p = nullptr;
						got = false;
						if (h->opt_graph_inline >= 1) return cuda_fail(h, e, "cudaMalloc of the neighbourhood blocks (graph_inline)");
					} else
						pr.first->cap = pr.second;
				}
			if (got) {
			h->g_bb_bytes = (uint32_t)bb; h->g_bb_idb = (uint32_t)bb_idb; h->g_ub_bytes = (uint32_t)ub; h->g_ub_idb = (uint32_t)ub_idb;
			const size_t wb = g.n * g.m0s;
			if (comp)
				build_bottom_code_blocks_kernel<<<(unsigned)((wb * 32 + 255) / 256), 256, 0, st>>>(h->g_codes.as<uint8_t>(), h->g_bottom.as<uint32_t>(), h->g_blk_bottom.as<uint8_t>(),
				                                                                                (uint32_t)g.n, (uint32_t)g.m0s, (uint32_t)h->d_pad, (uint32_t)bb, (uint32_t)bb_idb);
			else
				build_bottom_blocks_kernel<<<(unsigned)((wb * 32 + 255) / 256), 256, 0, st>>>(h->base, h->g_bottom.as<uint32_t>(), h->g_blk_bottom.as<uint8_t>(),
				                                                                           (uint32_t)g.n, (uint32_t)g.m0s, (uint32_t)h->d_pad, (uint32_t)bb, (uint32_t)bb_idb);
			EXPANN_CUDA(h, cudaGetLastError());
			if (urows) {
				const size_t wu = urows * g.ms;
				build_upper_blocks_kernel<<<(unsigned)((wu * 32 + 255) / 256), 256, 0, st>>>(h->g_up_vec.as<float>(), h->g_up_meta.as<uint4>(), h->g_upper_c.as<uint32_t>(),
				                                                                          h->g_blk_upper.as<uint8_t>(), (uint32_t)urows, (uint32_t)g.ms, (uint32_t)h->d_pad, (uint32_t)ub, (uint32_t)ub_idb);
				EXPANN_CUDA(h, cudaGetLastError());
			}
			h->stats.aux_launches += 2;
			h->g_inline = true;
			} // (automatic mode without the memory: the row forms serve every call)
		}
	}
	if (!h->g_inline) { h->g_blk_bottom.release(); h->g_blk_upper.release(); }
	EXPANN_CUDA(h, cudaStreamSynchronize(st));
	return EXPANN_OK;
}

int launch_graph_search(expann_gpu* h, const float* d_q, size_t m, size_t k, size_t ef, uint64_t* d_keys_out, cudaStream_t st) {
	if (ef == 0) ef = k * std::max<uint64_t>(1, h->cfg.ef_search_mult); // src/antitopo_engine.h:858-859
	const size_t cap = std::max(k, ef);
	if (cap > 256) return fail(h, EXPANN_ERR_UNSUPPORTED, "graph search: max(k, ef) above 256");
	int sms = 148;
	cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, h->device);
	GraphParams p{};
	p.base = h->base;
	p.codes = h->cfg.use_compression ? h->g_codes.as<uint8_t>() : nullptr;
	p.bottom = h->g_bottom.as<uint32_t>();
	p.upper_idx = h->g_upper_idx.as<uint32_t>();
	p.upper_nl = h->g_upper_nl.as<uint32_t>();
	p.upper = h->g_upper.as<uint32_t>();
	p.queries = d_q;
	p.out_keys = d_keys_out;
	p.visited = h->g_visited.as<uint8_t>();
	p.epoch = h->g_epoch.as<uint32_t>();
	p.distcomps = reinterpret_cast<unsigned long long*>(h->g_epoch.as<uint8_t>() + (size_t)h->g_ctas * GS_WARPS * 4);
	p.n = (uint32_t)h->n;
	p.vis_stride = (uint32_t)h->g_vis_stride;
	p.d_pad = (uint32_t)h->d_pad;
	p.m0s = (uint32_t)h->graph.m0s;
	p.ms = (uint32_t)h->graph.ms;
	p.m = (uint32_t)m;
	p.k = (uint32_t)k;
	p.ef = (uint32_t)ef;
	p.start = h->graph.start;
	p.max_layer = h->graph.max_layer;
	p.id_add = (uint32_t)h->cfg.id_offset;
	p.compressed = h->cfg.use_compression ? 1 : 0;
	// automatic: up to two waves of one CTA per SM (measured on 1M x 128, ef 40: 128 queries 188 vs 394 us, 256 queries
	// 320 vs 410 us, 512 queries 573 vs 419 us per call)
	// (over the neighbourhood blocks the one-CTA-per-query form stays ahead for longer: 296 queries 183 vs 398 us, 592 queries
	// 310 vs 408 us, 1024 queries 484 vs 426 us per call: automatic up to five waves)
	const bool blocks_ok = h->g_inline && h->opt_graph_inline != 0 && !(h->opt_graph_cta_width == 2 && ef >= 2);
	if (m <= (size_t)(h->opt_graph_cta_batch < 0 ? (blocks_ok ? 5 : 2) * sms : h->opt_graph_cta_batch)) {
		// latency form: one CTA per query (few queries per call; nothing to hide the hops' round trips behind)
		const size_t smem1 = (size_t)h->d_pad * 5 + GS_MAXNB * 8 + 16 + 256 * 8 + GS_MAXNB * 4; // query | u8 query | nb | nbd | ctl | beam copy
		if (smem1 > 200 * 1024) return fail(h, EXPANN_ERR_UNSUPPORTED, "graph search: d too large");
		const unsigned grid1 = (unsigned)std::min<size_t>(m, (size_t)h->g_ctas * GS_WARPS);
		if (h->g_nup && h->opt_graph_upper_compact) {
			p.up_meta = h->g_up_meta.as<uint4>();
			p.upper_c = h->g_upper_c.as<uint32_t>();
			p.up_vec = h->g_up_vec.as<float>();
			p.up_start_ui = h->graph.upper_idx[h->graph.start];
			p.up_start_nl = h->graph.upper_nl[h->graph.start];
		}
		p.width = h->opt_graph_cta_width == 2 && 2 * h->graph.m0s <= (size_t)GS_MAXNB && ef >= 2 ? 2 : 1;
		if (h->fuse_ids) { // small host call: the search decodes its own results (api.cu)
			p.dec_ids = h->fuse_ids;
			p.dec_dists = h->fuse_dists;
			p.dec_counts = h->fuse_counts;
			h->fused = true;
		}
		cudaEvent_t t0 = timing_begin(h, st);
		if (h->g_inline && h->opt_graph_inline != 0 && p.width == 1) {
			// neighbourhood blocks (graph_inline.cuh): one bulk copy per hop
			InlineParams ip{};
			ip.g = p;
			ip.blk_bottom = h->g_blk_bottom.as<uint8_t>();
			ip.blk_upper = h->g_blk_upper.as<uint8_t>();
			ip.bb_bytes = h->g_bb_bytes; ip.bb_idb = h->g_bb_idb; ip.ub_bytes = h->g_ub_bytes; ip.ub_idb = h->g_ub_idb;
			const size_t fixed = inline_fixed_smem(h->d_pad), lim = 227 * 1024;
			const size_t bitmap = ((h->n + 127) / 128) * 16; // bytes, whole uint4s
			ip.bitmap_words = fixed + bitmap <= lim && h->opt_graph_inline != 2 ? (uint32_t)(bitmap / 4) : 0;
			const size_t smemi = fixed + (size_t)ip.bitmap_words * 4;
#define GI_LAUNCH_N(L, N, R, C)                                                                                   \
	do {                                                                                                          \
		if (int rc_ = ensure_dyn_smem(h, (const void*)graph_search_inline_kernel<L, N, R, C>, smemi)) return rc_;       \
		graph_search_inline_kernel<L, N, R, C><<<grid1, GI_THREADS, smemi, st>>>(ip);                            \
	} while (0)
#define GI_LAUNCH(L)                                                                                              \
	do {                                                                                                          \
		const bool wide = std::max(h->graph.m0s, h->graph.ms) > 32 && std::max(h->graph.m0s, h->graph.ms) <= 128; \
		if (p.compressed && h->d_pad == 128) GI_LAUNCH_N(L, 4, 1, true);                                         \
		else if (p.compressed) GI_LAUNCH_N(L, 0, 1, true);                                                       \
		else if (h->d_pad == 128 && wide) GI_LAUNCH_N(L, 4, 4, false);                                           \
		else if (h->d_pad == 128) GI_LAUNCH_N(L, 4, 1, false);                                                   \
		else if (h->d_pad == 96 && wide) GI_LAUNCH_N(L, 3, 4, false);                                            \
		else if (h->d_pad == 96) GI_LAUNCH_N(L, 3, 1, false);                                                    \
		else GI_LAUNCH_N(L, 0, 1, false);                                                                        \
	} while (0)
			if (cap <= 32) GI_LAUNCH(1);
			else if (cap <= 64) GI_LAUNCH(2);
			else if (cap <= 128) GI_LAUNCH(4);
			else GI_LAUNCH(8);
#undef GI_LAUNCH_N
#undef GI_LAUNCH
			timing_end(h, t0, st);
			EXPANN_CUDA(h, cudaGetLastError());
			h->stats.graph_launches++;
			return EXPANN_OK;
		}
#define GC_LAUNCH(L)                                                                                              \
	do {                                                                                                          \
		if (int rc_ = ensure_dyn_smem(h, (const void*)graph_search_cta_kernel<L>, smem1)) return rc_;                    \
		graph_search_cta_kernel<L><<<grid1, 32 * GC_WARPS, smem1, st>>>(p);                                      \
	} while (0)
		if (cap <= 32) GC_LAUNCH(1);
		else if (cap <= 64) GC_LAUNCH(2);
		else if (cap <= 128) GC_LAUNCH(4);
		else GC_LAUNCH(8);
#undef GC_LAUNCH
		timing_end(h, t0, st);
		EXPANN_CUDA(h, cudaGetLastError());
		h->stats.graph_launches++;
		return EXPANN_OK;
	}
	size_t smem = ((size_t)h->d_pad * 5 + GS_MAXNB * 8) * GS_WARPS;
	if (smem > 200 * 1024) return fail(h, EXPANN_ERR_UNSUPPORTED, "graph search: d too large");
	{
		// per-warp visited hash in shared memory (graph_search.cuh): 2048 ids per warp when that leaves the resident CTAs
		// per SM the kernel is launched for (d = 128: 43.5 KB per CTA, five CTAs per SM); else the HBM bytes only
		const size_t occ = std::max<size_t>(1, h->g_ctas / (size_t)std::max(1, sms));
		const size_t with_hash = smem + (size_t)GS_HASH * 4 * GS_WARPS;
		p.hash = h->opt_graph_hash && (with_hash + 1024) * occ <= 227 * 1024 ? GS_HASH : 0;
		if (p.hash) smem = with_hash;
	}
	const unsigned grid = (unsigned)std::min<size_t>(h->g_ctas, (m + GS_WARPS - 1) / GS_WARPS);
	cudaEvent_t t0 = timing_begin(h, st);
	if (h->g_inline && h->opt_graph_inline != 0 && h->opt_graph_inline_batch && !p.compressed && !p.hash) {
		// the one-warp-per-query form over the neighbourhood blocks (graph_inline.cuh)
		InlineParams ip{};
		ip.g = p;
		ip.blk_bottom = h->g_blk_bottom.as<uint8_t>();
		ip.blk_upper = h->g_blk_upper.as<uint8_t>();
		ip.bb_bytes = h->g_bb_bytes; ip.bb_idb = h->g_bb_idb; ip.ub_bytes = h->g_ub_bytes; ip.ub_idb = h->g_ub_idb;
		const size_t smemb = ((size_t)h->d_pad * 4 + GS_MAXNB * 8) * GS_WARPS;
#define GB_LAUNCH_B(L, B)                                                                                         \
	do {                                                                                                          \
		if (int rc_ = ensure_dyn_smem(h, (const void*)graph_search_blocks_kernel<L, B>, smemb)) return rc_;              \
		graph_search_blocks_kernel<L, B><<<grid, 32 * GS_WARPS, smemb, st>>>(ip);                                \
	} while (0)
#define GB_LAUNCH(L)                                                                                              \
	do {                                                                                                          \
		const uint32_t occ = h->g_ctas / (uint32_t)std::max(1, sms);                                             \
		if (occ >= 8) GB_LAUNCH_B(L, 8);                                                                         \
		else if (occ >= 5) GB_LAUNCH_B(L, 5);                                                                    \
		else if (occ >= 4) GB_LAUNCH_B(L, 4);                                                                    \
		else GB_LAUNCH_B(L, 2);                                                                                  \
	} while (0)
		if (cap <= 32) GB_LAUNCH(1);
		else if (cap <= 64) GB_LAUNCH(2);
		else if (cap <= 128) GB_LAUNCH(4);
		else GB_LAUNCH(8);
#undef GB_LAUNCH
#undef GB_LAUNCH_B
		timing_end(h, t0, st);
		EXPANN_CUDA(h, cudaGetLastError());
		h->stats.graph_launches++;
		return EXPANN_OK;
	}
#define GS_LAUNCH_B(L, B)                                                                                         \
	do {                                                                                                          \
		if (int rc_ = ensure_dyn_smem(h, (const void*)graph_search_kernel<L, B>, smem)) return rc_;                      \
		graph_search_kernel<L, B><<<grid, 32 * GS_WARPS, smem, st>>>(p);                                         \
	} while (0)
#define GS_LAUNCH(L)                                                                                              \
	do {                                                                                                          \
		const uint32_t occ = h->g_ctas / (uint32_t)std::max(1, sms);                                             \
		if (occ >= 8) GS_LAUNCH_B(L, 8);                                                                         \
		else if (occ >= 5) GS_LAUNCH_B(L, 5);                                                                    \
		else if (occ >= 4) GS_LAUNCH_B(L, 4);                                                                    \
		else GS_LAUNCH_B(L, 2);                                                                                  \
	} while (0)
	if (cap <= 32) GS_LAUNCH(1);
	else if (cap <= 64) GS_LAUNCH(2);
	else if (cap <= 128) GS_LAUNCH(4);
	else GS_LAUNCH(8);
#undef GS_LAUNCH
#undef GS_LAUNCH_B
	timing_end(h, t0, st);
	EXPANN_CUDA(h, cudaGetLastError());
	h->stats.graph_launches++;
	return EXPANN_OK;
}

int graph_read_distcomps(expann_gpu* h) {
	if (!h->g_epoch.p) return EXPANN_OK;
	unsigned long long v = 0;
	EXPANN_CUDA(h, cudaMemcpy(&v, h->g_epoch.as<uint8_t>() + (size_t)h->g_ctas * GS_WARPS * 4, 8, cudaMemcpyDeviceToHost));
	h->stats.distcomps = v;
	return EXPANN_OK;
}

} // namespace expann
