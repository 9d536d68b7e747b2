#!/bin/bash
# round-2 GPU call A: test suite, slow-path A/B (v1 = all 32 values parked per entry, default = triple descent), one-term vs split format
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv,noheader
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -15
V1=$PWD/expann_b200/libexpann_gpu_v1.so
for n in 125000 250000 1000000; do
  N=$n TAG=sift_${n}_v2 timeout 120 python tools/time_tc.py 10000 2>&1 | tail -1
  EXPANN_GPU_LIB=$V1 N=$n TAG=sift_${n}_v1 timeout 120 python tools/time_tc.py 10000 2>&1 | tail -1
done
for real in 1 0; do
  GEN=deep D=96 N=2000000 REAL=$real TAG=deep2m_real$real timeout 200 python tools/time_tc.py 10000 2>&1 | tail -1
  GEN=gauss D=128 N=1000000 REAL=$real TAG=gauss1m_real$real timeout 200 python tools/time_tc.py 10000 2>&1 | tail -1
  GEN=gist D=960 N=200000 METRIC=1 REAL=$real TAG=gist200k_real$real timeout 200 python tools/time_tc.py 1000 2>&1 | tail -1
done
