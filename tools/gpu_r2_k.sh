#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_tensorcore.py tests/test_gpu_fuzz.py tests/test_gpu_bruteforce.py tests/test_gpu_multi.py -m gpu -q > gpurun_out/pytest_k.log 2>&1; echo "pytest rc=$?"; tail -12 gpurun_out/pytest_k.log | cut -c1-250
for tf in 1 0; do
  GEN=gist D=960 N=1000000 METRIC=1 TF32=$tf TAG=gist1m_tf$tf timeout 300 python tools/time_tc.py 1000 2>&1 | tail -1 | cut -c1-260
  GEN=gist D=960 N=1000000 METRIC=0 TF32=$tf TAG=gist1m_l2_tf$tf timeout 300 python tools/time_tc.py 1000 2>&1 | tail -1 | cut -c1-260
done
