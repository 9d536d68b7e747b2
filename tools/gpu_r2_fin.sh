#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_fin.log 2>&1; echo "pytest rc $?"; tail -3 gpurun_out/pytest_gpu_fin.log
python bench.py > gpurun_out/fin_bench.json 2> gpurun_out/fin_bench.err; echo "bench rc $?"
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/fin_bench_ref.json 2>> gpurun_out/fin_bench.err; echo "ref rc $?"; cut -c1-300 gpurun_out/fin_bench_ref.json
python - <<'PY'
import json
d=json.loads(open("gpurun_out/fin_bench.json").read().strip().splitlines()[-1])
print({k:d[k] for k in ("metric","value","ms_per_step","gpu_launches") if k in d}, "e2e", d["e2e"]["value"], "roofline frac", d["roofline"]["frac"])
c=d["roofline"]["graph"]["at_recall_0.95"]
print("graph", {k:c.get(k) for k in ("ef","recall","qps","e2e_qps")}, {k:round(v,1) for k,v in c["single_query_call"].items() if k.endswith("per_call")})
print("hbm_scan", [(r.get("queries"), round(r.get("frac",0),3)) for r in d["roofline"]["hbm_scan"].get("per_qtile", [])] if isinstance(d["roofline"].get("hbm_scan"),dict) else None)
print("c5", {k:d["roofline"]["c5"].get(k) for k in list(d["roofline"].get("c5",{}))[:0]})
PY
