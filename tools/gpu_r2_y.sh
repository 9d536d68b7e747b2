#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_graph.py -x -q 2>&1 | tail -4
timeout 400 python tools/graph_inline_check.py 1000000 128 > gpurun_out/x_1m.log 2>&1; echo "rc 1m $?"; tail -1 gpurun_out/x_1m.log
