#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_d.log 2>&1; echo "pytest rc $?"; tail -3 gpurun_out/pytest_gpu_d.log
timeout 300 python tools/graph_crossover.py 2>&1 | tail -12 | head -11
