#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
for bt in 64 128 256; do for ns in 0 14; do
  echo "BOOTTILES=$bt NSPLIT=$ns"
  BOOTTILES=$bt NSPLIT=$ns TAG=t_$bt_$ns timeout 200 python tools/time_tc.py 640,1250,2500,10000 2>&1 | grep "^{" | cut -c1-150
done; done
