// expann_gpu — the reference's `expann --config f.json` workflow (src/main.cpp:52-109) with the GPU engines:
//   config JSON / --flag overrides (keys dataset, ds_name, num_threads, n, m, d, k; src/main.cpp:32-67)
//   -> dataset (seeded form of vec_generator, src/randomgeometry.h:80-96; ground truth by exhaustive search,
//      src/dataset_loader.h:27-38; or .fvecs/.ivecs files, src/dataset_loader.h:96-168)
//   -> job grid (src/bench_runner.h:133-172) -> per job: store all, build, timed queries, recall
//      (src/basic_bench.h:58-149)
//   -> data/<ds_name>/data/{latest,all}.json in the bench_data schema (src/bench_data.h:8-43,
//      src/bench_data_manager.h:64-73) that src/pyplotter.py reads.
// Host code only; every distance is computed by libexpann_gpu.so on the GPU.
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <iostream>
#include <random>
#include <set>
#include <sstream>
#include <sys/stat.h>
#include "gpu_engines.hpp"

using namespace expann_host;

struct bench_data { // src/bench_data.h:8-18
	double time_per_query_ns = 0, time_to_build_ns = 0, average_distance = 0, average_squared_distance = 0, recall = 0;
	std::string engine_name;
	param_list_t param_list;
};

static std::string json_escape(const std::string& s) {
	std::string o;
	for (char c : s) { if (c == '"' || c == '\\') o += '\\'; o += c; }
	return o;
}
static std::string to_json(const bench_data& b, int indent) {
	std::ostringstream o;
	std::string p(indent, ' '), p2(indent + 4, ' '), p3(indent + 8, ' ');
	o.precision(17);
	o << p << "{\n" << p2 << "\"average_distance\": " << b.average_distance << ",\n"
	  << p2 << "\"average_squared_distance\": " << b.average_squared_distance << ",\n"
	  << p2 << "\"engine_name\": \"" << json_escape(b.engine_name) << "\",\n" << p2 << "\"param_list\": {";
	bool first = true;
	for (auto& kv : b.param_list) {
		o << (first ? "\n" : ",\n") << p3 << "\"" << json_escape(kv.first) << "\": \"" << json_escape(kv.second) << "\"";
		first = false;
	}
	o << (first ? "" : "\n" + p2) << "},\n" << p2 << "\"recall\": " << b.recall << ",\n"
	  << p2 << "\"time_per_query_ns\": " << b.time_per_query_ns << ",\n"
	  << p2 << "\"time_to_build_ns\": " << b.time_to_build_ns << "\n" << p << "}";
	return o.str();
}

// flat JSON object {"key": "string" | number, ...} — all the reference's config files are of this form
static std::map<std::string, std::string> parse_flat_json(const std::string& text) {
	std::map<std::string, std::string> kv;
	size_t i = 0;
	auto skip = [&] { while (i < text.size() && (isspace((unsigned char)text[i]) || text[i] == ',' || text[i] == '{' || text[i] == '}')) ++i; };
	while (true) {
		skip();
		if (i >= text.size() || text[i] != '"') break;
		size_t e = text.find('"', i + 1);
		std::string key = text.substr(i + 1, e - i - 1);
		i = text.find(':', e) + 1;
		while (i < text.size() && isspace((unsigned char)text[i])) ++i;
		std::string val;
		if (text[i] == '"') { e = text.find('"', i + 1); val = text.substr(i + 1, e - i - 1); i = e + 1; }
		else { e = i; while (e < text.size() && text[e] != ',' && text[e] != '}' && !isspace((unsigned char)text[e])) ++e; val = text.substr(i, e - i); i = e; }
		kv[key] = val;
	}
	return kv;
}

static void mkdirs(const std::string& path) {
	for (size_t i = 1; i <= path.size(); ++i)
		if (i == path.size() || path[i] == '/') mkdir(path.substr(0, i).c_str(), 0755);
}

struct dataset { // test_dataset, src/dataset.h:9-31
	std::string name;
	size_t n = 0, m = 0, k = 0, dim = 0;
	std::vector<float> vecs, queries;
	std::vector<std::vector<size_t>> answers;
};

// .fvecs / .ivecs records: (int32 d, d x 4 bytes) — Tvecs_read, src/dataset_loader.h:96-125
template <class T> static bool tvecs_read(const std::string& path, std::vector<T>& out, size_t& rows, size_t& dim, size_t limit) {
	std::ifstream f(path, std::ios::binary);
	if (!f) return false;
	rows = 0;
	int32_t d;
	while (rows < limit && f.read(reinterpret_cast<char*>(&d), 4)) {
		dim = (size_t)d;
		size_t off = out.size();
		out.resize(off + dim);
		if (!f.read(reinterpret_cast<char*>(out.data() + off), 4 * dim)) { out.resize(off); break; }
		++rows;
	}
	return rows > 0;
}

// Binary dataset cache: the reference keeps a dataset (vectors, queries, ground truth) as one text-JSON file
// ./data/<name>.dataset (save_imtd / load_imtd / exists_imtd, src/dataset_loader.h:42-95) so that the brute-force ground
// truth is computed once; the same idea as raw little-endian arrays ("EXPANNDS", version, n, m, k, dim, vectors, queries,
// n_answers x {count, ids}): 1M x 128 loads in a fraction of a second instead of minutes of JSON parsing.
static const char DS_MAGIC[8] = { 'E', 'X', 'P', 'A', 'N', 'N', 'D', 'S' };
static bool save_dataset_cache(const std::string& path, const dataset& ds) {
	std::ofstream f(path, std::ios::binary);
	if (!f) return false;
	const uint64_t hdr[5] = { 1, ds.n, ds.m, ds.k, ds.dim };
	f.write(DS_MAGIC, 8);
	f.write(reinterpret_cast<const char*>(hdr), sizeof(hdr));
	f.write(reinterpret_cast<const char*>(ds.vecs.data()), (std::streamsize)(ds.vecs.size() * 4));
	f.write(reinterpret_cast<const char*>(ds.queries.data()), (std::streamsize)(ds.queries.size() * 4));
	for (const auto& a : ds.answers) {
		const uint64_t c = a.size();
		f.write(reinterpret_cast<const char*>(&c), 8);
		for (size_t id : a) { const uint64_t v = id; f.write(reinterpret_cast<const char*>(&v), 8); }
	}
	return (bool)f;
}
static bool load_dataset_cache(const std::string& path, dataset& ds) {
	std::ifstream f(path, std::ios::binary);
	if (!f) return false;
	char magic[8];
	uint64_t hdr[5];
	if (!f.read(magic, 8) || std::memcmp(magic, DS_MAGIC, 8) != 0 || !f.read(reinterpret_cast<char*>(hdr), sizeof(hdr)) || hdr[0] != 1) return false;
	f.seekg(0, std::ios::end);
	const uint64_t fsize = (uint64_t)f.tellg();
	f.seekg(8 + sizeof(hdr), std::ios::beg);
	const uint64_t n = hdr[1], m = hdr[2], k = hdr[3], dim = hdr[4];
	if (dim == 0 || dim > (1u << 20) || n > fsize / (4 * dim) || m > fsize / (4 * dim) || (n + m) * dim * 4 + m * 8 > fsize) return false; // sizes vs file size
	ds.n = n; ds.m = m; ds.k = k; ds.dim = dim;
	ds.vecs.resize(n * dim);
	ds.queries.resize(m * dim);
	if (!f.read(reinterpret_cast<char*>(ds.vecs.data()), (std::streamsize)(n * dim * 4)) ||
	    !f.read(reinterpret_cast<char*>(ds.queries.data()), (std::streamsize)(m * dim * 4)))
		return false;
	ds.answers.assign(m, {});
	for (uint64_t q = 0; q < m; ++q) {
		uint64_t c;
		if (!f.read(reinterpret_cast<char*>(&c), 8) || c > n) return false;
		ds.answers[q].resize(c);
		for (uint64_t j = 0; j < c; ++j) {
			uint64_t v;
			if (!f.read(reinterpret_cast<char*>(&v), 8) || v >= n) return false;
			ds.answers[q][j] = (size_t)v;
		}
	}
	return true;
}

static dataset make_synthetic(size_t n, size_t m, size_t k, size_t d, const std::string& name) {
	// load_synethetic_uniform_sphere_points_no_cache (src/dataset_loader.h:10-41): iid N(0,1) components, base first
	// then queries; the reference seeds from random_device (src/randomgeometry.h:80-82) — fixed seed here.
	dataset ds;
	ds.name = name; ds.n = n; ds.m = m; ds.k = k; ds.dim = d;
	std::mt19937 gen(12345);
	std::normal_distribution<> nd(0, 1);
	auto fill = [&](std::vector<float>& out, size_t rows) {
		out.resize(rows * d);
		for (size_t i = 0; i < rows; ++i) {
			double n2;
			do {
				n2 = 0;
				for (size_t j = 0; j < d; ++j) { float v = (float)nd(gen); out[i * d + j] = v; n2 += (double)v * v; }
			} while (n2 < 1e-7); // src/randomgeometry.h:90-94
		}
	};
	fill(ds.vecs, n);
	fill(ds.queries, m);
	return ds;
}

static int g_gpus = 1; // --gpus: GPUs 0 .. gpus-1 behind every engine of the run

static void ground_truth(dataset& ds) { // src/dataset_loader.h:27-38: exhaustive search fills all_query_ans
	gpu_brute_force_engine_config gc;
	gc.gpus = g_gpus;
	gpu_brute_force_engine eng(gc);
	std::vector<uint64_t> ids, counts;
	for (size_t i = 0; i < ds.n; ++i) eng.store_vector(fvec(ds.vecs.begin() + i * ds.dim, ds.vecs.begin() + (i + 1) * ds.dim));
	eng.build();
	eng.query_k_batch(ds.queries.data(), ds.m, ds.k, ids, nullptr, &counts);
	ds.answers.assign(ds.m, {});
	for (size_t q = 0; q < ds.m; ++q)
		for (size_t j = 0; j < counts[q]; ++j) ds.answers[q].push_back(ids[q * ds.k + j]);
}

// basic_bench::get_benchmark_data (src/basic_bench.h:58-149): store all, build, m queries, recall vs the ground
// truth; the query phase uses the batched entry (param_list says so); `per_query` times the one-at-a-time loop of
// src/basic_bench.h:83-84 instead.
template <class Engine> static bench_data get_benchmark_data(const dataset& ds, Engine& eng, bool per_query) {
	bench_data bd;
	auto t0 = std::chrono::high_resolution_clock::now();
	for (size_t i = 0; i < ds.n; ++i) eng.store_vector(fvec(ds.vecs.begin() + i * ds.dim, ds.vecs.begin() + (i + 1) * ds.dim));
	eng.build();
	auto t1 = std::chrono::high_resolution_clock::now();
	std::vector<uint64_t> ids(ds.m * ds.k, UINT64_MAX), counts(ds.m, 0);
	std::vector<float> keys(ds.m * ds.k, 0.f);
	{ // warm-up (module load, buffer allocation) outside the timed region
		std::vector<uint64_t> wi;
		eng.query_k_batch(ds.queries.data(), std::min<size_t>(ds.m, 8), ds.k, wi);
	}
	auto q0 = std::chrono::high_resolution_clock::now();
	if (per_query) {
		for (size_t q = 0; q < ds.m; ++q) {
			auto ans = eng.query_k(fvec(ds.queries.begin() + q * ds.dim, ds.queries.begin() + (q + 1) * ds.dim), ds.k);
			counts[q] = ans.size();
			for (size_t j = 0; j < ans.size(); ++j) ids[q * ds.k + j] = ans[j];
		}
	} else
		eng.query_k_batch(ds.queries.data(), ds.m, ds.k, ids, &keys, &counts);
	auto q1 = std::chrono::high_resolution_clock::now();
	// bookkeeping outside the timed region (the reference keeps it inside, src/basic_bench.h:91-121)
	size_t hits = 0;
	double sd = 0, sd2 = 0;
	for (size_t q = 0; q < ds.m; ++q) {
		std::set<size_t> seen;
		for (size_t j = 0; j < counts[q]; ++j)
			if (!seen.insert(ids[q * ds.k + j]).second) std::cerr << "engine is buggy: duplicate id in answer" << std::endl; // :98-104
		for (size_t t : ds.answers[q]) hits += seen.count(t);
		if (counts[q]) { // distance to the best answer, recomputed on the host for the report only
			double d2 = 0;
			const float* x = &ds.vecs[ids[q * ds.k] * ds.dim];
			const float* y = &ds.queries[q * ds.dim];
			for (size_t j = 0; j < ds.dim; ++j) d2 += ((double)x[j] - y[j]) * ((double)x[j] - y[j]);
			sd += std::sqrt(d2);
			sd2 += d2;
		}
	}
	bd.time_per_query_ns = std::chrono::duration<double, std::nano>(q1 - q0).count() / ds.m;
	bd.time_to_build_ns = std::chrono::duration<double, std::nano>(t1 - t0).count();
	bd.average_distance = sd / ds.m;
	bd.average_squared_distance = sd2 / ds.m;
	bd.recall = double(hits) / (ds.m * ds.k);
	bd.engine_name = eng.name();
	bd.param_list = eng.param_list();
	bd.param_list["batched"] = per_query ? "0" : "1";
	return bd;
}

static void save_results(const std::vector<bench_data>& latest, const std::string& prefix) {
	// bench_data_manager::save (src/bench_data_manager.h:70-73): latest.json overwritten, all.json appended
	mkdirs(prefix + "data");
	auto dump = [&](const std::vector<std::string>& items, const std::string& file) {
		std::ofstream o(file);
		o << "[\n";
		for (size_t i = 0; i < items.size(); ++i) o << items[i] << (i + 1 < items.size() ? ",\n" : "\n");
		o << "]";
	};
	std::vector<std::string> cur;
	for (auto& b : latest) cur.push_back(to_json(b, 4));
	dump(cur, prefix + "data/latest.json");
	std::vector<std::string> all;
	{
		std::ifstream in(prefix + "data/all.json");
		std::stringstream ss;
		ss << in.rdbuf();
		std::string t = ss.str();
		size_t a = t.find('['), b = t.rfind(']');
		if (a != std::string::npos && b != std::string::npos && b > a) {
			std::string body = t.substr(a + 1, b - a - 1);
			if (body.find('{') != std::string::npos) all.push_back(body.substr(body.find_first_not_of(" \n\r\t"), body.find_last_not_of(" \n\r\t,") + 1 - body.find_first_not_of(" \n\r\t")));
		}
	}
	for (auto& c : cur) all.push_back(c);
	dump(all, prefix + "data/all.json");
}

int main(int argc, char** argv) {
	std::map<std::string, std::string> params;
	std::string config_file;
	for (int i = 1; i + 1 < argc; i += 2) { // --key value (src/main.cpp:32-50)
		std::string key = argv[i];
		if (key.rfind("--", 0) == 0) key = key.substr(2);
		if (key == "config") config_file = argv[i + 1]; else params[key] = argv[i + 1];
	}
	if (argc == 2 && (std::string(argv[1]) == "--help" || std::string(argv[1]) == "-h")) {
		std::cout << "usage: expann_gpu --config f.json [--dataset Synthetic|Sift1M] [--ds_name s] [--n N] [--m M] [--d D] [--k K]\n"
		             "                  [--engines bf,graph,quant] [--M0 32] [--per_query 0|1] [--out_prefix dir/] [--gpus G]\n"
		             "                  [--sift_dir datasets/sift/] [--cache 0|1] [--cache_file path]\n";
		return 0;
	}
	if (!config_file.empty()) {
		std::ifstream f(config_file);
		if (!f) { std::cerr << "cannot read " << config_file << std::endl; return 2; }
		std::stringstream ss;
		ss << f.rdbuf();
		for (auto& kv : parse_flat_json(ss.str())) if (!params.count(kv.first)) params[kv.first] = kv.second; // flags win (:40-44)
	}
	auto get = [&](const std::string& k, const std::string& def) { return params.count(k) ? params[k] : def; };
	const std::string dataset_kind = get("dataset", "Synthetic");
	const size_t n = std::stoul(get("n", "56000")), m = std::stoul(get("m", "400")), d = std::stoul(get("d", "128")), k = std::stoul(get("k", "10"));
	const std::string ds_name = get("ds_name", "synthetic_n_" + std::to_string(n));
	const std::string engines = get("engines", "bf,graph");
	const bool per_query = get("per_query", "0") == "1";
	const size_t M0 = std::stoul(get("M0", "32"));
	g_gpus = std::max(1, std::stoi(get("gpus", "1")));
	try {
		dataset ds;
		if (dataset_kind == "Sift1M") { // load_sift1m, src/dataset_loader.h:127-168 (paths of src/main.cpp:77-79)
			size_t rows = 0, dim = 0, qrows = 0, grows = 0, gdim = 0;
			std::vector<int32_t> gt;
			const std::string dir = get("sift_dir", "datasets/sift/");
			if (!tvecs_read<float>(dir + "sift_base.fvecs", ds.vecs, rows, dim, n) ||
			    !tvecs_read<float>(dir + "sift_query.fvecs", ds.queries, qrows, dim, m)) {
				std::cerr << "Sift1M files not found under " << dir << std::endl;
				return 2;
			}
			ds.name = ds_name; ds.n = rows; ds.m = qrows; ds.k = k; ds.dim = dim;
			if (rows == 1000000 && tvecs_read<int32_t>(dir + "sift_groundtruth.ivecs", gt, grows, gdim, qrows)) {
				ds.answers.assign(qrows, {});
				for (size_t q = 0; q < qrows; ++q) for (size_t j = 0; j < k && j < gdim; ++j) ds.answers[q].push_back(gt[q * gdim + j]); // :148-154
			} else
				ground_truth(ds);
		} else {
			// load_synethetic_uniform_sphere_points (src/dataset_loader.h:76-95): the cached dataset if there is one, else make
			// it (vectors, queries, exhaustive ground truth) and cache it; --cache 0 is the _no_cache form (:10-41)
			const std::string cache = "./data/synthetic_uniform_sphere_n" + std::to_string(n) + "_dim" + std::to_string(d) + "_m" +
			                          std::to_string(m) + "_k" + std::to_string(k) + ".dataset.bin";
			const std::string cache_path = get("cache_file", cache);
			const bool use_cache = get("cache", "0") == "1";
			if (use_cache && load_dataset_cache(cache_path, ds)) {
				ds.name = ds_name;
				std::printf("dataset cache: loaded %s (n=%zu m=%zu k=%zu dim=%zu)\n", cache_path.c_str(), ds.n, ds.m, ds.k, ds.dim);
			} else {
				ds = make_synthetic(n, m, k, d, ds_name);
				ground_truth(ds);
				if (use_cache) {
					mkdirs(cache_path.substr(0, cache_path.rfind('/') == std::string::npos ? 0 : cache_path.rfind('/')));
					if (save_dataset_cache(cache_path, ds)) std::printf("dataset cache: saved %s\n", cache_path.c_str());
					else std::cerr << "dataset cache: cannot write " << cache_path << std::endl; // the reference fails silently here (src/dataset_loader.h:65-76)
				}
			}
		}
		std::vector<bench_data> latest;
		auto report = [&](const bench_data& b) {
			std::printf("%-36s recall=%.4f  %.1f ns/query (%.0f QPS)  build %.2f s\n", b.engine_name.c_str(), b.recall,
			            b.time_per_query_ns, 1e9 / b.time_per_query_ns, b.time_to_build_ns * 1e-9);
			latest.push_back(b);
		};
		if (engines.find("bf") != std::string::npos) {
			gpu_brute_force_engine_config c;
			c.gpus = g_gpus;
			gpu_brute_force_engine e(c);
			report(get_benchmark_data(ds, e, per_query));
		}
		if (engines.find("graph") != std::string::npos) {
			// the reference's grid (src/bench_runner.h:133-172): ef_search_mult 1..6 x use_compression {0,1}; the graph is
			// built once and re-read through the index file like the reference's read_index/write_index pair
			const std::string index = get("out_prefix", "./data/" + ds_name + "/") + "index_M0_" + std::to_string(M0) + ".index";
			mkdirs(index.substr(0, index.rfind('/')));
			bool first = true;
			if (g_gpus > 1) {
				// several GPUs: the graph is built once on one GPU and written to the index file; every job of the grid then
				// reads it on all GPUs (replicas: the queries of a call are split between the GPUs, mode (A) of SURVEY 8(e))
				gpu_graph_engine_config c;
				c.M = M0 / 2; c.M0 = M0; c.ef_construction = 2 * M0; c.index_filename = index; c.write_index = true;
				gpu_graph_engine b(c);
				for (size_t i = 0; i < ds.n; ++i) b.store_vector(fvec(ds.vecs.begin() + i * ds.dim, ds.vecs.begin() + (i + 1) * ds.dim));
				b.build();
				first = false;
			}
			for (int comp = 0; comp <= (get("compression", "0") == "1" ? 1 : 0); ++comp)
				for (size_t efm = 1; efm <= 6; ++efm) {
					gpu_graph_engine_config c;
					c.M = M0 / 2; c.M0 = M0; c.ef_search_mult = efm; c.ef_construction = 2 * M0; c.use_compression = comp;
					c.index_filename = index; c.read_index = !first; c.write_index = first;
					c.search_width = std::stoul(get("search_width", "1"));
					c.gpus = g_gpus;
					gpu_graph_engine e(c);
					report(get_benchmark_data(ds, e, per_query));
					first = false;
				}
		}
		if (engines.find("quant") != std::string::npos) {
			gpu_quant_scan_engine_config c;
			c.quantizer = 1; c.kprime = 100; c.gpus = g_gpus;
			gpu_quant_scan_engine e(c);
			report(get_benchmark_data(ds, e, per_query));
		}
		save_results(latest, get("out_prefix", "./data/" + ds_name + "/")); // src/main.cpp:105-106
	} catch (const std::exception& ex) {
		std::cerr << "expann_gpu: " << ex.what() << std::endl;
		return 1;
	}
	return 0;
}
