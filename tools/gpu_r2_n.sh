#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_multi.py tests/test_gpu_bruteforce.py tests/test_gpu_tensorcore.py -m gpu -q -x > gpurun_out/pytest_n.log 2>&1; echo "pytest rc=$?"; tail -12 gpurun_out/pytest_n.log | cut -c1-250
DEVICES=0,1 MS=10000,5000 timeout 300 python tools/time_multi.py 2>&1 | grep "ms/step" | cut -c1-200
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 200 --warmup 5 --no-extras > gpurun_out/bench_n2.json 2> gpurun_out/bench_n2.err; echo rc=$?; cut -c1-400 gpurun_out/bench_n2.json
