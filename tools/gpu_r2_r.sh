#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_tensorcore.py tests/test_gpu_fuzz.py tests/test_gpu_bruteforce.py tests/test_gpu_multi.py tests/test_gpu_fullsize.py -m gpu -q > gpurun_out/pytest_r.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/pytest_r.log | cut -c1-250
for boot in 1 0 1 0; do
  echo "C2 BOOT=$boot"
  BOOT=$boot TAG=r_b$boot timeout 200 python tools/time_tc.py 300,1250,2500,5000,10000 2>&1 | grep "^{" | cut -c1-150
done
for boot in 1 0; do
  echo "N=125000 BOOT=$boot"
  N=125000 BOOT=$boot TAG=r125_b$boot timeout 200 python tools/time_tc.py 1250,10000 2>&1 | grep "^{" | cut -c1-150
  echo "deep 2M x 96 (mode 3) BOOT=$boot"
  GEN=deep D=96 N=2000000 BOOT=$boot TAG=rdeep_b$boot timeout 300 python tools/time_tc.py 1250,10000 2>&1 | grep "^{" | cut -c1-150
  echo "gist 1M x 960 ip (mode 4) BOOT=$boot"
  GEN=gist D=960 N=1000000 METRIC=1 BOOT=$boot TAG=rgist_b$boot timeout 300 python tools/time_tc.py 1000 2>&1 | grep "^{" | cut -c1-150
done
