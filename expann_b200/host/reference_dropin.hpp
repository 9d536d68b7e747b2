// The GPU engines AS REFERENCE ENGINES: this header is meant to be compiled INSIDE the reference tree (it includes the
// reference's own "ann_engine.h" / "vec.h"), next to brute_force_engine.h and antitopo_engine.h. The classes derive
// from the real  template <typename T, class Derived> struct ann_engine  (src/ann_engine.h:16-29), take the real vec<T>
// (src/vec.h:17-23: contiguous Eigen storage, internal.data()), and carry the nested `config` + `Engine(config)`
// constructor that job<Engine, Engine::config> needs (src/bench_runner.h:15-33,116), so they can be listed in a JobTuple
// and driven by the unmodified basic_bench::get_benchmark_data (src/basic_bench.h:58-149). Everything forwards to the C
// ABI in include/expann_gpu.h; nothing is computed on the host.
//
// The repo's test infrastructure compiles exactly this against /root/reference (dropin_harness.cpp, make target
// `dropin`) and tests/test_gpu_dropin.py runs the reference's own harness over it.
#pragma once
#include <stdexcept>
#include <string>
#include <vector>

#include "ann_engine.h" // the reference's (src/ann_engine.h): ann_engine, param_list_t, add_param; pulls in vec.h
#include "expann_gpu.h"

namespace expann_dropin {
inline void check(expann_gpu_t* h, int rc) {
	if (rc != EXPANN_OK) throw std::runtime_error(expann_gpu_last_error(h));
}
} // namespace expann_dropin

// ---- brute_force_engine<T> (src/brute_force_engine.h:10-46) on the GPU -------------------------------------------
struct gpu_brute_force_engine_config {
	int gpus = 1; // GPUs 0 .. gpus-1 behind the one engine
	gpu_brute_force_engine_config(int _gpus = 1) : gpus(_gpus) {}
};
template <typename T> struct gpu_brute_force_engine : public ann_engine<T, gpu_brute_force_engine<T>> {
	static_assert(sizeof(T) == 4, "the GPU path computes in fp32 like the reference's float engines");
	using config = gpu_brute_force_engine_config;
	config conf;
	expann_gpu_t* h = nullptr;
	explicit gpu_brute_force_engine(config c = config()) : conf(c) {
		expann_gpu_cfg g{};
		g.engine = EXPANN_ENGINE_BRUTE_FORCE;
		g.metric = EXPANN_METRIC_L2;
		g.n_devices = c.gpus;
		if (expann_gpu_create(&g, &h) != EXPANN_OK) throw std::runtime_error(expann_gpu_last_error(nullptr));
	}
	gpu_brute_force_engine(const gpu_brute_force_engine&) = delete;
	~gpu_brute_force_engine() { expann_gpu_destroy(h); }
	void _store_vector(const vec<T>& v) { expann_dropin::check(h, expann_gpu_add(h, v.internal.data(), 1, v.size())); }
	void _build() { expann_dropin::check(h, expann_gpu_build(h)); }
	std::vector<size_t> _query_k(const vec<T>& v, size_t k) {
		std::vector<uint64_t> ids(k);
		uint64_t cnt = 0;
		expann_dropin::check(h, expann_gpu_query(h, v.internal.data(), 1, k, 0, ids.data(), nullptr, &cnt));
		return std::vector<size_t>(ids.begin(), ids.begin() + cnt);
	}
	// all m queries of a harness in one call (rows of q contiguous, m x dim)
	std::vector<std::vector<size_t>> query_k_batch(const std::vector<vec<T>>& qs, size_t k) {
		const size_t m = qs.size(), d = m ? qs[0].size() : 0;
		std::vector<T> flat(m * d);
		for (size_t i = 0; i < m; ++i) std::copy(qs[i].internal.data(), qs[i].internal.data() + d, flat.begin() + i * d);
		std::vector<uint64_t> ids(m * k), cnt(m);
		expann_dropin::check(h, expann_gpu_query(h, flat.data(), m, k, 0, ids.data(), nullptr, cnt.data()));
		std::vector<std::vector<size_t>> out(m);
		for (size_t i = 0; i < m; ++i) out[i].assign(ids.begin() + i * k, ids.begin() + i * k + cnt[i]);
		return out;
	}
	const std::string _name() { return "Brute-Force Engine (B200)"; }
	const param_list_t _param_list() {
		param_list_t pl;
		add_param(pl, conf.gpus);
		return pl;
	}
};

// ---- antitopo_engine<T> (src/antitopo_engine.h:103-1074): its config, its index file, its query semantics --------
// The constructor arguments are antitopo_engine_config's (src/antitopo_engine.h:73-101) in the same order, so a job
// list written for antitopo_engine<float> works unchanged. read_index = true searches the graph the reference built and
// serialised (src/antitopo_engine.h:931-991); otherwise the graph is built on the GPU (ef_construction = candidates per
// vertex, at most 127) and write_index writes it in the reference's format.
struct gpu_graph_engine_config {
	size_t M, M0, ef_search_mult, ef_construction, ortho_count;
	float ortho_factor, ortho_bias;
	size_t prune_overflow;
	bool use_compression, use_largest_direction_filtering;
	std::string index_filename;
	bool read_index, write_index;
	int gpus = 1;
	gpu_graph_engine_config(size_t _M, size_t _M0, size_t _ef_search_mult, size_t _ef_construction, size_t _ortho_count,
	                        float _ortho_factor, float _ortho_bias, size_t _prune_overflow, bool _use_compression = false,
	                        bool _use_largest_direction_filtering = false, std::string _index_filename = "",
	                        bool _read_index = false, bool _write_index = false, int _gpus = 1)
	    : M(_M), M0(_M0), ef_search_mult(_ef_search_mult), ef_construction(_ef_construction), ortho_count(_ortho_count),
	      ortho_factor(_ortho_factor), ortho_bias(_ortho_bias), prune_overflow(_prune_overflow), use_compression(_use_compression),
	      use_largest_direction_filtering(_use_largest_direction_filtering), index_filename(_index_filename),
	      read_index(_read_index), write_index(_write_index), gpus(_gpus) {}
};
template <typename T> struct gpu_graph_engine : public ann_engine<T, gpu_graph_engine<T>> {
	static_assert(sizeof(T) == 4, "the GPU path computes in fp32 like the reference's float engines");
	using config = gpu_graph_engine_config;
	config conf;
	expann_gpu_t* h = nullptr;
	size_t ef_search = 0;
	explicit gpu_graph_engine(config c) : conf(c) {
		// the GPU build takes exact nearest neighbours as candidates: there are no repeated candidate searches to count
		if (!c.read_index && c.ortho_count != 1) throw std::invalid_argument("gpu_graph_engine: ortho_count must be 1 when the graph is built on the GPU");
		expann_gpu_cfg g{};
		g.engine = EXPANN_ENGINE_GRAPH;
		g.metric = EXPANN_METRIC_L2;
		g.M = c.M;
		g.M0 = c.M0;
		g.ef_search_mult = c.ef_search_mult;
		g.use_compression = c.use_compression;
		g.n_devices = c.gpus;
		if (expann_gpu_create(&g, &h) != EXPANN_OK) throw std::runtime_error(expann_gpu_last_error(nullptr));
	}
	gpu_graph_engine(const gpu_graph_engine&) = delete;
	~gpu_graph_engine() { expann_gpu_destroy(h); }
	void set_ef_search(size_t ef) { ef_search = ef; } // src/antitopo_engine.h:189-195
	void _store_vector(const vec<T>& v) {
		if (conf.read_index) return; // src/antitopo_engine.h:312-313
		expann_dropin::check(h, expann_gpu_add(h, v.internal.data(), 1, v.size()));
	}
	void _build() { // src/antitopo_engine.h:467-493
		if (conf.read_index) expann_dropin::check(h, expann_gpu_load_index(h, conf.index_filename.c_str()));
		else {
			expann_dropin::check(h, expann_gpu_build_graph(h, conf.M, conf.M0, conf.ef_construction < 127 ? conf.ef_construction : 127, conf.ortho_factor,
			                                               conf.ortho_bias, conf.prune_overflow, 0));
			if (conf.write_index) expann_dropin::check(h, expann_gpu_save_index(h, conf.index_filename.c_str()));
		}
		expann_dropin::check(h, expann_gpu_build(h));
	}
	std::vector<size_t> _query_k(const vec<T>& v, size_t k) {
		std::vector<uint64_t> ids(k);
		uint64_t cnt = 0;
		expann_dropin::check(h, expann_gpu_query(h, v.internal.data(), 1, k, ef_search, ids.data(), nullptr, &cnt)); // ef 0 = k * ef_search_mult (:858-859)
		return std::vector<size_t>(ids.begin(), ids.begin() + cnt);
	}
	const std::string _name() { return "Anti-Topo Engine (B200 query path)"; }
	const param_list_t _param_list() { // the names of src/antitopo_engine.h:211-238
		param_list_t pl;
		add_param(pl, conf.M);
		add_param(pl, conf.M0);
		add_param(pl, conf.ef_search_mult);
		add_param(pl, conf.ef_construction);
		add_param(pl, conf.ortho_count);
		add_param(pl, conf.ortho_factor);
		add_param(pl, conf.ortho_bias);
		add_param(pl, conf.prune_overflow);
		add_param(pl, conf.use_compression);
		add_param(pl, conf.read_index);
		add_param(pl, conf.gpus);
		return pl;
	}
};
