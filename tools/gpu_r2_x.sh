#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 400 python tools/graph_inline_check.py 1000000 128 > gpurun_out/x_1m.log 2>&1; echo "rc 1m $?"; tail -1 gpurun_out/x_1m.log
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_b.log 2>&1; echo "pytest rc $?"; tail -3 gpurun_out/pytest_gpu_b.log
timeout 600 ncu --set full --clock-control none --import-source on -k regex:graph_search_inline --launch-skip 60 -c 2 -f -o gpurun_out/r2_lat_inline python tools/profile_graph_latency.py 1 > gpurun_out/x_ncu.log 2>&1; echo "ncu rc $?"
