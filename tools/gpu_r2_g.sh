#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
for dbg in 0 1 2 4; do
  echo "== mgpu_debug=$dbg"
  DEVICES=0,1 ROWG=0 MS=10000 OPTS=mgpu_debug=$dbg timeout 300 python tools/time_multi.py 2>&1 | grep -E "^multi|Error" | tail -2
done
echo "== tc_boot=0"
DEVICES=0,1 ROWG=0 MS=10000,8000 OPTS=tc_boot=0 timeout 300 python tools/time_multi.py 2>&1 | grep -E "^multi|Error" | tail -4
echo "== N=500000"
N=500000 DEVICES=0,1 ROWG=0 MS=10000 timeout 300 python tools/time_multi.py 2>&1 | grep -E "^multi|^single|Error" | tail -4
