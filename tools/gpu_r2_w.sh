#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 600 ncu --set full --clock-control none --import-source on -k regex:graph_search_inline --launch-skip 60 -c 3 -f -o gpurun_out/r2_lat_inline python tools/profile_graph_latency.py 1 > gpurun_out/w_ncu.log 2>&1; echo "ncu rc $?"
tail -3 gpurun_out/w_ncu.log
