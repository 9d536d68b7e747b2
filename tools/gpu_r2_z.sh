#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python -m pytest tests/test_gpu_dropin.py tests/test_gpu_graph.py tests/test_abi.py -m gpu -x -q 2>&1 | tail -3
python bench.py > gpurun_out/z_bench.json 2> gpurun_out/z_bench.err; echo "bench rc $?"; tail -c 600 gpurun_out/z_bench.err
python - <<'PY'
import json
d=json.loads(open("gpurun_out/z_bench.json").read().strip().splitlines()[-1])
print({k:d[k] for k in ("metric","value","ms_per_step","gpu_launches") if k in d})
print("e2e", d.get("e2e"))
g=d["roofline"].get("graph",{})
print("graph", {k:g.get(k) for k in ("at_recall_0.95",)} )
PY
