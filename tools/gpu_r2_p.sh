#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_tensorcore.py tests/test_gpu_fuzz.py tests/test_gpu_bruteforce.py tests/test_gpu_multi.py -m gpu -q -x > gpurun_out/pytest_p.log 2>&1; echo "pytest rc=$?"; tail -12 gpurun_out/pytest_p.log | cut -c1-250
for bt in 64 32 16; do for boot in 1 0; do
  echo "BOOT=$boot BOOTTILES=$bt"
  BOOT=$boot BOOTTILES=$bt TAG=p_b${boot}_t${bt} timeout 200 python tools/time_tc.py 300,1250,2500,5000,10000 2>&1 | grep "^{" | cut -c1-150
  if [ $boot = 0 ] && [ $bt != 64 ]; then break; fi
done; done
N=125000 BOOT=1 TAG=p125 timeout 200 python tools/time_tc.py 1250,10000 2>&1 | grep "^{" | cut -c1-150
N=125000 BOOT=0 TAG=p125b0 timeout 200 python tools/time_tc.py 1250,10000 2>&1 | grep "^{" | cut -c1-150
