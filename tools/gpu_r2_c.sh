#!/bin/bash
# round-2 GPU call C: full test log, warp-cooperative hit path (default) vs per-lane form (v1 library)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_c.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_c.log
V1=$PWD/expann_b200/libexpann_gpu_v1.so
for n in 125000 1000000; do
  N=$n TAG=sift_${n}_warp timeout 120 python tools/time_tc.py 10000 2>&1 | tail -1
  EXPANN_GPU_LIB=$V1 N=$n TAG=sift_${n}_lane timeout 120 python tools/time_tc.py 10000 2>&1 | tail -1
done
GEN=deep D=96 N=2000000 TAG=deep2m_warp timeout 200 python tools/time_tc.py 10000 2>&1 | tail -1
EXPANN_GPU_LIB=$V1 GEN=deep D=96 N=2000000 TAG=deep2m_lane timeout 200 python tools/time_tc.py 10000 2>&1 | tail -1
