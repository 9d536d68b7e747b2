// Host side of the GPU graph construction + the index-file writer.
#include "engine.h"
#include "graph_build.cuh"
#include <cub/cub.cuh>
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <random>

namespace expann {

namespace {

struct LayerOut { // host result of one layer over `ns` points (local ids)
	std::vector<uint32_t> ids;
	std::vector<float> len;
	uint32_t stride = 0;
};

struct Scratch {
	DevBuf keys, cand, fwd_ids, fwd_len, fwd_keys, tkey, tval, tkey2, tval2, hist, offs, cub_tmp, fin_ids, fin_len, rowmap, sub;
	void release() {
		for (DevBuf* b : { &keys, &cand, &fwd_ids, &fwd_len, &fwd_keys, &tkey, &tval, &tkey2, &tval2, &hist, &offs, &cub_tmp,
		                   &fin_ids, &fin_len, &rowmap, &sub })
			b->release();
	}
};

// one layer: exact kNN over the engine's CURRENT base (h->base / h->n / tensor map), prune, reverse, prune
int build_layer(expann_gpu* h, Scratch& s, const uint32_t* d_rowmap_for_prune, const float* prune_base, uint32_t K,
                uint32_t mcap, float ortho_factor, float ortho_bias, uint32_t leniency, LayerOut& out) {
	cudaStream_t st = h->stream;
	const uint32_t ns = (uint32_t)h->n;
	out.stride = mcap;
	out.ids.assign((size_t)ns * mcap, NO_VERTEX);
	out.len.assign((size_t)ns * mcap, 0.f);
	if (ns < 2) return EXPANN_OK;
	K = std::min(K, ns - 1);
	const uint32_t kk = K + 1;
	// ---- candidates: exact K nearest neighbours (K1 scan, queries = the base rows themselves)
	const uint32_t batch = 32768;
	EXPANN_CUDA(h, s.keys.ensure((size_t)batch * kk * 8));
	EXPANN_CUDA(h, s.cand.ensure((size_t)ns * K * 8));
	const int saved_metric = h->cfg.metric;
	const uint64_t saved_off = h->cfg.id_offset;
	h->cfg.metric = EXPANN_METRIC_L2;
	h->cfg.id_offset = 0;
	int rc = EXPANN_OK;
	for (uint32_t r0 = 0; r0 < ns && !rc; r0 += batch) {
		const uint32_t rows = std::min(batch, ns - r0);
		rc = launch_scan_f32(h, h->base + (size_t)r0 * h->d_pad, rows, kk, s.keys.as<uint64_t>(), st);
		if (rc) break;
		knn_strip_self_kernel<<<(rows + 3) / 4, 128, 0, st>>>(s.keys.as<uint64_t>(), kk, r0, rows, s.cand.as<uint64_t>(), K);
		h->stats.aux_launches++;
	}
	h->cfg.metric = saved_metric;
	h->cfg.id_offset = saved_off;
	if (rc) return rc;
	EXPANN_CUDA(h, cudaGetLastError());
	// ---- forward prune
	const size_t tot = (size_t)ns * mcap;
	EXPANN_CUDA(h, s.fwd_ids.ensure(tot * 4));
	EXPANN_CUDA(h, s.fwd_len.ensure(tot * 4));
	PruneParams p{};
	p.base = prune_base;
	p.rowmap = d_rowmap_for_prune;
	p.candA = s.cand.as<uint64_t>();
	p.candB = nullptr;
	p.offB = nullptr;
	p.out_ids = s.fwd_ids.as<uint32_t>();
	p.out_len = s.fwd_len.as<float>();
	p.n = ns;
	p.ka = K;
	p.capB = 0;
	p.d_pad = (uint32_t)h->d_pad;
	p.stride = mcap;
	p.mcap = mcap;
	p.leniency = leniency;
	p.ortho_factor = ortho_factor;
	p.ortho_bias = ortho_bias;
	p.lazy = 0;
	const size_t smem = (size_t)GB_MAXC * 8 + (size_t)GB_MAXC * (h->d_pad + 1) * 4;
	if (smem > 220 * 1024) return fail(h, EXPANN_ERR_UNSUPPORTED, "graph build: d too large for the prune kernel");
	if (int rc = ensure_dyn_smem(h, (const void*)prune_select_kernel, smem)) return rc;
	prune_select_kernel<<<ns, GB_THREADS, smem, st>>>(p);
	EXPANN_CUDA(h, cudaGetLastError());
	h->stats.aux_launches++;
	// ---- reverse edges: tuples sorted by (target, dist); segment offsets from a histogram
	EXPANN_CUDA(h, s.tkey.ensure(tot * 8));
	EXPANN_CUDA(h, s.tkey2.ensure(tot * 8));
	EXPANN_CUDA(h, s.tval.ensure(tot * 4));
	EXPANN_CUDA(h, s.tval2.ensure(tot * 4));
	EXPANN_CUDA(h, s.hist.ensure(((size_t)ns + 1) * 4));
	EXPANN_CUDA(h, s.offs.ensure(((size_t)ns + 1) * 4));
	EXPANN_CUDA(h, cudaMemsetAsync(s.hist.p, 0, ((size_t)ns + 1) * 4, st));
	reverse_tuples_kernel<<<(unsigned)((tot + 255) / 256), 256, 0, st>>>(s.fwd_ids.as<uint32_t>(), s.fwd_len.as<float>(), ns, mcap,
	                                                                      s.tkey.as<uint64_t>(), s.tval.as<uint32_t>(),
	                                                                      s.hist.as<uint32_t>());
	size_t tmp_bytes = 0, tmp2 = 0;
	cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, s.tkey.as<uint64_t>(), s.tkey2.as<uint64_t>(), s.tval.as<uint32_t>(),
	                                s.tval2.as<uint32_t>(), (int)tot, 0, 64, st);
	cub::DeviceScan::ExclusiveSum(nullptr, tmp2, s.hist.as<uint32_t>(), s.offs.as<uint32_t>(), (int)ns + 1, st);
	EXPANN_CUDA(h, s.cub_tmp.ensure(std::max(tmp_bytes, tmp2)));
	size_t tb = s.cub_tmp.cap;
	EXPANN_CUDA(h, cub::DeviceRadixSort::SortPairs(s.cub_tmp.p, tb, s.tkey.as<uint64_t>(), s.tkey2.as<uint64_t>(),
	                                                s.tval.as<uint32_t>(), s.tval2.as<uint32_t>(), (int)tot, 0, 64, st));
	tb = s.cub_tmp.cap;
	EXPANN_CUDA(h, cub::DeviceScan::ExclusiveSum(s.cub_tmp.p, tb, s.hist.as<uint32_t>(), s.offs.as<uint32_t>(), (int)ns + 1, st));
	reverse_keys_kernel<<<(unsigned)((tot + 255) / 256), 256, 0, st>>>(s.tkey2.as<uint64_t>(), s.tval2.as<uint32_t>(), tot,
	                                                                    s.tkey.as<uint64_t>());
	EXPANN_CUDA(h, s.fwd_keys.ensure(tot * 8));
	lists_to_keys_kernel<<<(unsigned)((tot + 255) / 256), 256, 0, st>>>(s.fwd_ids.as<uint32_t>(), s.fwd_len.as<float>(), tot,
	                                                                     s.fwd_keys.as<uint64_t>());
	h->stats.aux_launches += 5;
	// ---- lazy prune of forward ∪ reverse (src/antitopo_engine.h:442-455)
	EXPANN_CUDA(h, s.fin_ids.ensure(tot * 4));
	EXPANN_CUDA(h, s.fin_len.ensure(tot * 4));
	p.candA = s.fwd_keys.as<uint64_t>();
	p.ka = mcap;
	p.candB = s.tkey.as<uint64_t>();
	p.offB = s.offs.as<uint32_t>();
	p.capB = GB_MAXC - mcap;
	p.out_ids = s.fin_ids.as<uint32_t>();
	p.out_len = s.fin_len.as<float>();
	p.lazy = 1;
	prune_select_kernel<<<ns, GB_THREADS, smem, st>>>(p);
	EXPANN_CUDA(h, cudaGetLastError());
	h->stats.aux_launches++;
	EXPANN_CUDA(h, cudaMemcpyAsync(out.ids.data(), s.fin_ids.p, tot * 4, cudaMemcpyDeviceToHost, st));
	EXPANN_CUDA(h, cudaMemcpyAsync(out.len.data(), s.fin_len.p, tot * 4, cudaMemcpyDeviceToHost, st));
	EXPANN_CUDA(h, cudaStreamSynchronize(st));
	return EXPANN_OK;
}

} // namespace

int graph_build_gpu(expann_gpu* h, uint64_t M, uint64_t M0, uint64_t K, float ortho_factor, float ortho_bias,
                    uint64_t prune_overflow, uint64_t seed) {
	if (h->n == 0) return fail(h, EXPANN_ERR_STATE, "build_graph: no vectors stored");
	if (M < 2 || M0 < 2 || M0 > 96 || M > 96) return fail(h, EXPANN_ERR_ARG, "build_graph: need 2 <= M, M0 <= 96");
	if (K == 0) K = std::min<uint64_t>(127, 2 * M0);
	if (K > 127) return fail(h, EXPANN_ERR_ARG, "build_graph: K must be <= 127");
	const size_t n = h->n;
	// levels: floor(-ln(U)/ln(M)) with mt19937(seed) — the reference's generator (src/antitopo_engine.h:159,323)
	std::mt19937 gen((uint32_t)seed);
	std::uniform_real_distribution<> dist(0, 1);
	std::vector<uint32_t> level(n);
	uint32_t top = 0, start = 0;
	for (size_t v = 0; v < n; ++v) {
		level[v] = (uint32_t)std::floor(-std::log(dist(gen)) * 1 / std::log(double(M)));
		if (level[v] + 1 > top) { top = level[v] + 1; start = (uint32_t)v; } // :459-462
	}
	if (int rc = make_base_tensor_map(h)) return rc;
	Scratch s;
	HostGraph& g = h->graph;
	g = HostGraph();
	g.n = n;
	g.m0s = (M0 + 3) / 4 * 4;
	g.ms = (M + 3) / 4 * 4;
	g.start = start;
	g.max_layer = top;
	const uint32_t leniency = (uint32_t)prune_overflow + 1;
	// ---- bottom layer over all points. Its exact kNN (n queries against n rows, k = K + 1) is a large-batch brute-force
	// query: give it the tensor-core path (K3 first form, k <= 128; results are bit-identical to K1's, so the graph is
	// the same) by preparing the bf16 operand for the duration of this layer -- 38.6 vs 166.6 ms per 16384 rows at 1M x 128.
	if (h->opt_tc && n >= 65536) {
		const int saved_engine = h->cfg.engine, saved_metric = h->cfg.metric;
		h->cfg.engine = EXPANN_ENGINE_BRUTE_FORCE;
		h->cfg.metric = EXPANN_METRIC_L2;
		const int trc = tc_finalize(h);
		h->cfg.engine = saved_engine;
		h->cfg.metric = saved_metric;
		if (trc) h->tc_mode = 0; // not fatal: the SIMT scan does the same job
	}
	LayerOut lo;
	int rc = build_layer(h, s, nullptr, h->base, (uint32_t)K, (uint32_t)M0, ortho_factor, ortho_bias, leniency, lo);
	if (h->tc_mode) { // the upper layers scan row subsets, and the search does not use the operand: give the memory back
		h->tc_mode = 0;
		cudaStreamSynchronize(h->stream);
		for (DevBuf* b : { &h->tc_xb, &h->tc_xn, &h->tc_xx, &h->tc_cand, &h->tc_qb, &h->tc_sel }) b->release();
	}
	if (rc) { s.release(); return rc; }
	g.bottom.assign(n * g.m0s, NO_VERTEX);
	g.bottom_len.assign(n * g.m0s, 0.f);
	for (size_t v = 0; v < n; ++v)
		for (uint32_t j = 0; j < lo.stride; ++j) {
			g.bottom[v * g.m0s + j] = lo.ids[v * lo.stride + j];
			g.bottom_len[v * g.m0s + j] = lo.len[v * lo.stride + j];
		}
	// ---- upper layers: each over the subset of points whose level reaches it
	g.upper_idx.assign(n, NO_VERTEX);
	g.upper_nl.assign(n, 1);
	size_t urows = 0;
	for (size_t v = 0; v < n; ++v) {
		g.upper_nl[v] = level[v] + 1;
		if (level[v] > 0) { g.upper_idx[v] = (uint32_t)urows; urows += level[v]; }
	}
	g.upper.assign(std::max<size_t>(1, urows) * g.ms, NO_VERTEX);
	g.upper_len.assign(std::max<size_t>(1, urows) * g.ms, 0.f);
	float* full_base = h->base;
	const size_t full_n = h->n;
	const CUtensorMap full_map = h->tmap_base;
	for (uint32_t layer = 1; layer < top && !rc; ++layer) {
		std::vector<uint32_t> members;
		for (size_t v = 0; v < n; ++v) if (level[v] >= layer) members.push_back((uint32_t)v);
		const uint32_t ns = (uint32_t)members.size();
		if (ns < 2) continue;
		cudaError_t e = s.rowmap.ensure((size_t)ns * 4);
		if (e == cudaSuccess) e = s.sub.ensure((size_t)ns * h->d_pad * 4);
		if (e == cudaSuccess) e = cudaMemcpyAsync(s.rowmap.p, members.data(), (size_t)ns * 4, cudaMemcpyHostToDevice, h->stream);
		if (e != cudaSuccess) { rc = cuda_fail(h, e, "build_graph upper layer"); break; }
		gather_rows_kernel<<<(unsigned)(((size_t)ns * h->d_pad + 255) / 256), 256, 0, h->stream>>>(full_base, s.rowmap.as<uint32_t>(), ns,
		                                                                                          (uint32_t)h->d_pad, s.sub.as<float>());
		// the scan works on "the engine's base": point it at the subset for this layer
		h->base = s.sub.as<float>();
		h->n = ns;
		rc = make_base_tensor_map(h);
		LayerOut up;
		if (!rc) rc = build_layer(h, s, nullptr, s.sub.as<float>(), (uint32_t)K, (uint32_t)M, ortho_factor, ortho_bias, leniency, up);
		h->base = full_base;
		h->n = full_n;
		h->tmap_base = full_map;
		if (rc) break;
		for (uint32_t i = 0; i < ns; ++i) {
			const uint32_t v = members[i];
			const size_t row = (size_t)g.upper_idx[v] + (layer - 1);
			for (uint32_t j = 0; j < up.stride; ++j) {
				const uint32_t loc = up.ids[(size_t)i * up.stride + j];
				g.upper[row * g.ms + j] = loc == NO_VERTEX ? NO_VERTEX : members[loc];
				g.upper_len[row * g.ms + j] = up.len[(size_t)i * up.stride + j];
			}
		}
	}
	s.release();
	if (rc) return rc;
	h->cfg.M = M;
	h->cfg.M0 = M0;
	h->build_ortho_factor = ortho_factor;
	h->build_ortho_bias = ortho_bias;
	h->build_prune_overflow = prune_overflow;
	h->build_K = K;
	g.present = true;
	h->built = false;
	return EXPANN_OK;
}

// index file writer — the reference's serialize format (src/antitopo_engine.h:931-991), readable by its
// deserialize (:994-1074)
int graph_save_index(expann_gpu* h, const char* path) {
	const HostGraph& g = h->graph;
	if (!g.present) return fail(h, EXPANN_ERR_STATE, "save_index: no graph");
	std::vector<float> rows(h->n * h->d);
	cudaError_t e = cudaMemcpy2D(rows.data(), h->d * 4, h->base, h->d_pad * 4, h->d * 4, h->n, cudaMemcpyDeviceToHost);
	if (e != cudaSuccess) return cuda_fail(h, e, "save_index");
	FILE* f = std::fopen(path, "wb");
	if (!f) return fail(h, EXPANN_ERR_IO, std::string("cannot create ") + path);
	auto wr = [&](const void* p, size_t sz) { return std::fwrite(p, 1, sz, f) == sz; };
	const uint64_t start = g.start, M = h->cfg.M, M0 = h->cfg.M0, efm = std::max<uint64_t>(1, h->cfg.ef_search_mult);
	const uint8_t has_ef = 0, comp = h->cfg.use_compression ? 1 : 0, ldf = 0;
	const uint64_t efc = h->build_K, oc = 1, po = h->build_prune_overflow, ml = g.max_layer, n = h->n, d = h->d;
	bool ok = wr(&start, 8) && wr(&M, 8) && wr(&M0, 8) && wr(&efm, 8) && wr(&has_ef, 1) && wr(&efc, 8) && wr(&oc, 8) &&
	          wr(&h->build_ortho_factor, 4) && wr(&h->build_ortho_bias, 4) && wr(&po, 8) && wr(&comp, 1) && wr(&ldf, 1) &&
	          wr(&ml, 8) && wr(&n, 8);
	for (uint64_t i = 0; ok && i < n; ++i) ok = wr(&d, 8) && wr(rows.data() + i * d, 4 * d);
	ok = ok && wr(&n, 8);
	std::vector<uint8_t> rec;
	for (uint64_t v = 0; ok && v < n; ++v) {
		const uint64_t nl = g.upper_nl[v];
		ok = wr(&nl, 8);
		for (uint64_t l = 0; ok && l < nl; ++l) {
			const uint32_t* ids = l == 0 ? &g.bottom[v * g.m0s] : &g.upper[((size_t)g.upper_idx[v] + l - 1) * g.ms];
			const float* len = nullptr;
			if (l == 0 && !g.bottom_len.empty()) len = &g.bottom_len[v * g.m0s];
			if (l > 0 && !g.upper_len.empty()) len = &g.upper_len[((size_t)g.upper_idx[v] + l - 1) * g.ms];
			const size_t stride = l == 0 ? g.m0s : g.ms;
			uint64_t deg = 0;
			while (deg < stride && ids[deg] != NO_VERTEX) ++deg;
			rec.resize(deg * 12);
			for (uint64_t j = 0; j < deg; ++j) {
				const float ln = len ? len[j] : 0.f;
				const uint64_t id = ids[j];
				std::memcpy(rec.data() + j * 12, &ln, 4);
				std::memcpy(rec.data() + j * 12 + 4, &id, 8);
			}
			ok = wr(&deg, 8) && (deg == 0 || wr(rec.data(), rec.size()));
		}
	}
	ok = std::fclose(f) == 0 && ok;
	if (!ok) return fail(h, EXPANN_ERR_IO, std::string("short write to ") + path);
	return EXPANN_OK;
}

} // namespace expann
