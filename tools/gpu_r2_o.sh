#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -x > gpurun_out/pytest_o_alone.log 2>&1; echo "alone rc=$?"; tail -15 gpurun_out/pytest_o_alone.log | cut -c1-300
timeout 1700 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_o_all.log 2>&1; echo "all rc=$?"; tail -8 gpurun_out/pytest_o_all.log | cut -c1-300
