#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
DEVICES=0,0 ROWG=0 timeout 300 python tools/time_multi.py 2>&1 | grep -v Warn | tail -12
DEVICES=0,0 ROWG=1 MS=10000 timeout 300 python tools/time_multi.py 2>&1 | grep multi | tail -4
DEVICES=0,0 ROWG=0 MS=10000 OPTS=tc_boot=0 timeout 300 python tools/time_multi.py 2>&1 | grep multi | tail -4
