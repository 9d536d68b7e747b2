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
# one GPU's share of C2 at 8 GPUs under the three decompositions: 8 row shards (125k rows x 10k queries, above),
# 4 row shards x 2 query groups, replicas (1M rows x 1250 queries)
N=250000 TAG=sift_250k_m5000 timeout 120 python tools/time_tc.py 5000 2>&1 | tail -1
N=500000 TAG=sift_500k_m2500 timeout 120 python tools/time_tc.py 2500 2>&1 | tail -1
N=1000000 TAG=sift_1m_m1250 timeout 120 python tools/time_tc.py 1250 2>&1 | tail -1
for real in 1 0; do
  GEN=deep D=96 N=2000000 REAL=$real TAG=deep2m_real$real timeout 200 python tools/time_tc.py 10000 2>&1 | tail -1
  GEN=gauss D=128 N=1000000 REAL=$real TAG=gauss1m_real$real timeout 200 python tools/time_tc.py 10000 2>&1 | tail -1
  GEN=gist D=960 N=200000 METRIC=1 REAL=$real TAG=gist200k_real$real timeout 200 python tools/time_tc.py 1000 2>&1 | tail -1
done
