#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_tensorcore.py tests/test_gpu_fuzz.py tests/test_gpu_bruteforce.py tests/test_gpu_multi.py tests/test_gpu_fullsize.py -m gpu -q > gpurun_out/pytest_q.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/pytest_q.log | cut -c1-250
for cfg in "1 1" "0 1" "1 0" "0 0"; do set -- $cfg
  echo "BOOT=$1 BUCKETS=$2"
  BOOT=$1 BUCKETS=$2 TAG=q_b$1_g$2 timeout 200 python tools/time_tc.py 300,1250,2500,5000,10000 2>&1 | grep "^{" | cut -c1-150
done
for cfg in "1 1" "0 1" "0 0"; do set -- $cfg
  echo "N=125000 BOOT=$1 BUCKETS=$2"
  N=125000 BOOT=$1 BUCKETS=$2 TAG=q125_b$1_g$2 timeout 200 python tools/time_tc.py 1250,10000 2>&1 | grep "^{" | cut -c1-150
done
