#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_c.log 2>&1; echo "pytest rc $?"; tail -3 gpurun_out/pytest_gpu_c.log
python bench.py > gpurun_out/z_bench.json 2> gpurun_out/z_bench.err; echo "bench rc $?"; tail -c 300 gpurun_out/z_bench.err
python - <<'PY'
import json
d=json.loads(open("gpurun_out/z_bench.json").read().strip().splitlines()[-1])
print({k:d[k] for k in ("metric","value","ms_per_step","gpu_launches") if k in d})
g=d["roofline"].get("graph",{})
c=g.get("at_recall_0.95",{})
print("graph", {k:c.get(k) for k in ("ef","recall","qps","kernel_ms","gather_GBs","e2e_qps","single_query_call")})
print("sweep", [(r["ef"], round(r["recall"],4), round(r["qps"])) for r in g.get("sweep",[])])
PY
