#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 300 python tools/graph_inline_check.py 20000 128 > gpurun_out/u_small.log 2>&1; echo "rc small $?"
tail -20 gpurun_out/u_small.log
timeout 300 python tools/graph_inline_check.py 5000 96 > gpurun_out/u_d96.log 2>&1; echo "rc d96 $?"
tail -12 gpurun_out/u_d96.log
timeout 400 python tools/graph_inline_check.py 1000000 128 > gpurun_out/u_1m.log 2>&1; echo "rc 1m $?"
tail -20 gpurun_out/u_1m.log
