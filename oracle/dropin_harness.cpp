// TEST INFRASTRUCTURE (not product code): the reference's OWN harness driving the GPU engines.
//
// Compiled against /root/reference by oracle/Makefile (target dropin, only where that tree exists): includes the
// reference's bench_runner.h / basic_bench.h / dataset_loader.h / bench_data_manager.h UNMODIFIED and the product
// header expann_b200/host/reference_dropin.hpp, whose engines derive from the real ann_engine<T, Derived>. It
//   1. makes a dataset with the reference's own loader (load_synethetic_uniform_sphere_points_no_cache,
//      src/dataset_loader.h:10-41: ground truth by the reference's CPU brute_force_engine),
//   2. lists  antitopo_engine<float> (CPU, builds + writes the index),  gpu_brute_force_engine<float>,
//      gpu_graph_engine<float> (reads the SAME index file)  in a JobTuple (src/bench_runner.h:116,131),
//   3. runs perform_and_store_benchmark_results (src/bench_runner.h:107-114), i.e. job::run ->
//      basic_bench::get_benchmark_data (src/basic_bench.h:58-149) per engine, one query per call,
//   4. lets the reference's bench_data_manager write <prefix>data/latest.json (src/bench_data_manager.h:70-73).
// usage: expann_dropin <n> <m> <k> <out_prefix>      (dimension = DIM = 128, the reference's compiled-in value)
#include <cstdlib>
#include <filesystem>
#include <iostream>

#include "bench_runner.h"
#include "reference_dropin.hpp"

int main(int argc, char** argv) {
	if (argc < 5) {
		std::cerr << "usage: expann_dropin n m k out_prefix" << std::endl;
		return 2;
	}
	const size_t n = std::strtoul(argv[1], nullptr, 10), m = std::strtoul(argv[2], nullptr, 10), k = std::strtoul(argv[3], nullptr, 10);
	const std::string prefix = argv[4];
	std::filesystem::create_directories(prefix + "index");
	dataset_loader<float> dsl;
	auto ds = dsl.load_synethetic_uniform_sphere_points_no_cache("dropin_synthetic", n, m, k, DIM);
	basic_bench<float, in_memory_test_dataset<float>> benchmarker(ds);
	size_t job_index = 0;
	// tuple order = execution order on the one worker thread: the CPU engine writes the index the GPU engine then reads
	JobTuple<antitopo_engine<float>, gpu_brute_force_engine<float>, gpu_graph_engine<float>> job_lists;
	const std::string index = prefix + "index/dropin_k16";
	const size_t M = 16;
	for (size_t mult : { 1, 2, 4 })
		ADD_JOB(antitopo_engine<float>, M, 2 * M, mult, 4 * M, 1, 0.5f, 0.0f, 0, false, false, index, mult != 1, mult == 1);
	ADD_JOB(gpu_brute_force_engine<float>, 1);
	for (size_t mult : { 1, 2, 4 })
		ADD_JOB(gpu_graph_engine<float>, M, 2 * M, mult, 4 * M, 1, 0.5f, 0.0f, 0, false, false, index, true, false);
	bench_data_manager bdm = std::apply(
	    [&](auto&... jobs) { return perform_and_store_benchmark_results(ds.name, 1, job_index, benchmarker, jobs...); }, job_lists);
	bdm.save(prefix);
	std::cout << "dropin: " << bdm.latest.size() << " results in " << prefix << "data/latest.json" << std::endl;
	return bdm.latest.size() == job_index ? 0 : 1;
}
