#!/bin/bash
# round-2 GPU call D: full test log (no -x), hit path with triple mask (default) vs per-lane v1, C5 code scan on tensor cores
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/pytest_d.log 2>&1; echo "pytest rc=$?"; tail -40 gpurun_out/pytest_d.log | cut -c1-200
V1=$PWD/expann_b200/libexpann_gpu_v1.so
for n in 125000 1000000; do
  N=$n TAG=sift_${n}_v5 timeout 120 python tools/time_tc.py 10000 2>&1 | tail -1
  EXPANN_GPU_LIB=$V1 N=$n TAG=sift_${n}_v1b timeout 120 python tools/time_tc.py 10000 2>&1 | tail -1
done
timeout 600 python bench.py --workload c5 --steps 5 --warmup 3 --no-cpu --no-extras > gpurun_out/bench_c5_quant.json 2> gpurun_out/bench_c5_quant.err; echo "bench c5 rc=$?"; tail -3 gpurun_out/bench_c5_quant.err; cut -c1-1800 gpurun_out/bench_c5_quant.json
