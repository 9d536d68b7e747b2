#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
TAG=s1250 python tools/time_tc.py 1250 > gpurun_out/plain_s.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:tc_filter_pair -s 8 -c 2 -f -o gpurun_out/prof_m1250 python tools/time_tc.py 1250 > gpurun_out/ncu_s.log 2>&1
echo rc=$?; tail -3 gpurun_out/ncu_s.log | cut -c1-200
