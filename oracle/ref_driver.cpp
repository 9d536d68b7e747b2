// TEST INFRASTRUCTURE — not part of the product path.
//
// C-ABI driver that compiles the UNMODIFIED reference engines from
// /root/reference/src (never copied into this repo) into oracle/_ref/*.so so
// that tests / golden generation / the bench's cpu_baseline leg can run the
// reference's own code on identical inputs.
//
// Reference entry points exercised:
//   brute_force_engine<float>::_store_vector/_build/_query_k  (src/brute_force_engine.h:19-46)
//   antitopo_engine<float>  ctor(config) / _store_vector / _build / _query_k /
//     set_ef_search / serialize                              (src/antitopo_engine.h:103-1074)
//   dist2 / dot (Eigen)                                      (src/vec.h:109-111,139-142)
//   distance_compare_avx512f_* / dot_avx512f_f16             (src/distance.h:14-190)
//   quantizer_simple<uint8_t>::build                         (src/quantizer.h:125-150)
//
// Built by oracle/Makefile with -DDIM=<d> (fixed Eigen vectors, what the shipped
// binary uses: CMakeLists.txt:87) or without DIM (dynamic Eigen vectors).
#define RECORD_STATS 1
#include <atomic>
#include <cstdint>
#include <cstring>
#include <fstream>
#include <thread>
#include <vector>

#include "antitopo_engine.h"
#include "brute_force_engine.h"
#include "vec.h"

namespace {
vec<float> make_vec(const float* p, size_t d) {
	std::vector<float> tmp(p, p + d);
	return vec<float>(tmp);
}
struct bf_handle {
	brute_force_engine<float> eng;
	size_t d = 0;
};
struct graph_handle {
	antitopo_engine<float>* eng = nullptr;
	size_t d = 0;
	~graph_handle() { delete eng; }
};
} // namespace

extern "C" {

int ref_compiled_dim() {
#ifdef DIM
	return DIM;
#else
	return 0;
#endif
}

// ---- primitives -----------------------------------------------------------
float ref_dist2(const float* a, const float* b, size_t d) {
	return dist2(make_vec(a, d), make_vec(b, d)); // src/vec.h:139-142
}
float ref_dot(const float* a, const float* b, size_t d) {
	return dot(make_vec(a, d), make_vec(b, d)); // src/vec.h:156
}
float ref_dist2_batch128(const float* a, const float* b, size_t d) {
	return distance_compare_avx512f_f16_batch128(a, b, d); // src/distance.h:86-111
}
float ref_dist2_f16(const float* a, const float* b, size_t d) {
	return distance_compare_avx512f_f16(a, b, d); // src/distance.h:136-148
}
float ref_dot_f16(const float* a, const float* b, size_t d) {
	return dot_avx512f_f16(a, b, d); // src/distance.h:181-190
}
int32_t ref_dist2_i16(const int16_t* a, const int16_t* b, size_t d) {
	return distance_compare_avx512f_i32(a, b, d); // src/distance.h:14-27
}
int32_t ref_dist2_i8(const int8_t* a, const int8_t* b, size_t d) {
	return distance_compare_avx512f_i64(a, b, d); // src/distance.h:29-53
}
// u8 codes vs float query through the exact code path of the compressed graph
// search: swizzle (src/antitopo_engine.h:726-737) + dist2_compressed (:38-61).
int32_t ref_dist2_compressed(const float* q, const uint8_t* code, size_t d) {
	Eigen::VectorX<uint32_t> q_swizzled(d);
	for (size_t i = 0; i < d; i += 64)
		for (size_t j = 0; j < 4; ++j)
			for (size_t k = 0; k < 16; ++k)
				q_swizzled[i + j * 16 + k] = uint32_t(q[i + k * 4 + j]);
	Eigen::VectorX<uint8_t> b(d);
	std::memcpy(b.data(), code, d);
	return dist2_compressed(q_swizzled, b);
}

// ---- brute force engine ---------------------------------------------------
void* ref_bf_create() { return new bf_handle(); }
void ref_bf_destroy(void* h) { delete static_cast<bf_handle*>(h); }
void ref_bf_store(void* h, const float* rows, size_t n, size_t d) {
	auto* b = static_cast<bf_handle*>(h);
	b->d = d;
	b->eng.all_entries.reserve(b->eng.all_entries.size() + n);
	for (size_t i = 0; i < n; ++i)
		b->eng.store_vector(make_vec(rows + i * d, d));
}
void ref_bf_build(void* h) { static_cast<bf_handle*>(h)->eng.build(); }
// m queries, `threads` host threads over one shared read-only engine (safe:
// _query_k only reads all_entries). ids_out is m*k, padded with UINT64_MAX;
// dist_out (optional) receives dist2(query, base[id]) recomputed by the reference.
void ref_bf_query(void* h, const float* q, size_t m, size_t k, uint64_t* ids_out,
									float* dist_out, uint64_t* counts_out, int threads) {
	auto* b = static_cast<bf_handle*>(h);
	const size_t d = b->d;
	if (threads < 1)
		threads = 1;
	std::atomic<size_t> next(0);
	auto work = [&]() {
		for (;;) {
			size_t i = next++;
			if (i >= m)
				break;
			vec<float> qv = make_vec(q + i * d, d);
			std::vector<size_t> ans = b->eng._query_k(qv, k);
			for (size_t j = 0; j < k; ++j) {
				ids_out[i * k + j] = j < ans.size() ? uint64_t(ans[j]) : UINT64_MAX;
				if (dist_out)
					dist_out[i * k + j] = j < ans.size()
																		? dist2(qv, b->eng.all_entries[ans[j]])
																		: std::numeric_limits<float>::infinity();
			}
			if (counts_out)
				counts_out[i] = ans.size();
		}
	};
	std::vector<std::thread> pool;
	for (int t = 1; t < threads; ++t)
		pool.emplace_back(work);
	work();
	for (auto& t : pool)
		t.join();
}

// ---- antitopo (graph) engine ----------------------------------------------
void* ref_graph_create(size_t M, size_t M0, size_t ef_search_mult,
											 size_t ef_construction, size_t ortho_count,
											 float ortho_factor, float ortho_bias,
											 size_t prune_overflow, int use_compression) {
	auto* g = new graph_handle();
	antitopo_engine_config conf(M, M0, ef_search_mult, ef_construction,
															ortho_count, ortho_factor, ortho_bias,
															prune_overflow, use_compression != 0);
	g->eng = new antitopo_engine<float>(conf);
	return g;
}
void ref_graph_destroy(void* h) { delete static_cast<graph_handle*>(h); }
void ref_graph_store(void* h, const float* rows, size_t n, size_t d) {
	auto* g = static_cast<graph_handle*>(h);
	g->d = d;
	for (size_t i = 0; i < n; ++i)
		g->eng->_store_vector(make_vec(rows + i * d, d), /*silent=*/true);
}
void ref_graph_build(void* h) { static_cast<graph_handle*>(h)->eng->build(); }
void ref_graph_set_ef(void* h, size_t ef) {
	static_cast<graph_handle*>(h)->eng->ef_search = ef;
}
void ref_graph_set_compression(void* h, int on) {
	auto* g = static_cast<graph_handle*>(h);
	g->eng->use_compression = on != 0;
	if (on)
		g->eng->quant->build(g->eng->all_entries, g->eng->hadj_bottom);
}
int ref_graph_save(void* h, const char* path) {
	std::ofstream out(path, std::ios::binary);
	if (!out)
		return -1;
	static_cast<graph_handle*>(h)->eng->serialize(out);
	return out.good() ? 0 : -1;
}
int ref_graph_load(void* h, const char* path, size_t d) {
	auto* g = static_cast<graph_handle*>(h);
	std::ifstream in(path, std::ios::binary);
	if (!in)
		return -1;
	g->eng->deserialize(in);
	g->d = d;
	return 0;
}
uint64_t ref_graph_distcomps(void* h) {
	return static_cast<graph_handle*>(h)->eng->num_distcomps;
}
void ref_graph_reset_stats(void* h) {
	auto* e = static_cast<graph_handle*>(h)->eng;
	e->num_distcomps = 0;
	e->num_distcomps_compressed = 0;
}
// sequential (engine is not re-entrant: src/antitopo_engine.h:208-209)
void ref_graph_query(void* h, const float* q, size_t m, size_t k,
										 uint64_t* ids_out, uint64_t* counts_out) {
	auto* g = static_cast<graph_handle*>(h);
	const size_t d = g->d;
	for (size_t i = 0; i < m; ++i) {
		std::vector<size_t> ans = g->eng->_query_k(make_vec(q + i * d, d), k);
		for (size_t j = 0; j < k; ++j)
			ids_out[i * k + j] = j < ans.size() ? uint64_t(ans[j]) : UINT64_MAX;
		if (counts_out)
			counts_out[i] = ans.size();
	}
}
uint64_t ref_graph_info(void* h, int what) {
	auto* e = static_cast<graph_handle*>(h)->eng;
	switch (what) {
	case 0:
		return e->starting_vertex;
	case 1:
		return e->max_layer;
	case 2:
		return e->all_entries.size();
	case 3:
		return e->M;
	case 4:
		return e->M0;
	}
	return 0;
}
} // extern "C"
