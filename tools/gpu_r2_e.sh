#!/bin/bash
# round-2 GPU call E (N GPUs): multi-device handle on real devices -- tests, the driver's torchrun launch of bench.py, CLI --gpus
cd "$(dirname "$0")/.."
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name --format=csv,noheader | head -8
nvidia-smi topo -m 2>/dev/null | head -12
timeout 600 python -m pytest tests/test_gpu_multi.py tests/test_gpu_cli.py -m gpu -q > gpurun_out/pytest_e_$N.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_e_$N.log | cut -c1-200
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/bench_e_$N.json 2> gpurun_out/bench_e_$N.err; echo "bench rc=$?"; tail -3 gpurun_out/bench_e_$N.err; cut -c1-2500 gpurun_out/bench_e_$N.json
timeout 300 expann_b200/bin/expann_gpu --n 200000 --m 2000 --d 128 --k 10 --engines bf,graph --gpus $N --out_prefix /tmp/cli_e/ 2>&1 | tail -9
