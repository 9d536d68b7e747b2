#!/bin/bash
# round-2 GPU call B: test suite, slow-path A/B in one session, the new bench.py end to end
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q 2>&1 | tail -25
V1=$PWD/expann_b200/libexpann_gpu_v1.so
for n in 125000 1000000; do
  for rep in 1 2; do
    N=$n TAG=sift_${n}_v2_$rep timeout 120 python tools/time_tc.py 10000 2>&1 | tail -1
    EXPANN_GPU_LIB=$V1 N=$n TAG=sift_${n}_v1_$rep timeout 120 python tools/time_tc.py 10000 2>&1 | tail -1
  done
done
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/bench_b.json 2> gpurun_out/bench_b.err; echo "bench rc=$?"; tail -3 gpurun_out/bench_b.err; cut -c1-1500 gpurun_out/bench_b.json
