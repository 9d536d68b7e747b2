#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python -m pytest tests/test_gpu_multi.py -m gpu -x -q 2>&1 | tail -3
N=$(nvidia-smi -L | wc -l)
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/m_bench_$N.json 2> gpurun_out/m_bench_$N.err; echo "bench rc $?"; tail -c 400 gpurun_out/m_bench_$N.err
python - <<PY
import json
d=json.loads(open("gpurun_out/m_bench_$N.json").read().strip().splitlines()[-1])
print({k:d[k] for k in ("metric","value","ms_per_step","n_gpus") if k in d})
print("e2e", d["e2e"]["value"])
r=d["roofline"]
print({k:(r[k].get("value") if isinstance(r[k],dict) else None) for k in r if k in ("c4","c5","row_sharded","graph")})
print(json.dumps(r.get("c4"))[:600])
PY
