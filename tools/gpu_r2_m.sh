#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
for boot in 0 2; do for ns in 0 7 14 22 29; do
  echo "BOOT=$boot NSPLIT=$ns"
  BOOT=$boot NSPLIT=$ns TAG=m_b${boot}_n${ns} timeout 200 python tools/time_tc.py 1250,2500,5000 2>&1 | grep "^{" | cut -c1-160
done; done
