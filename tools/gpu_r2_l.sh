#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 600 python - <<'PY' > gpurun_out/c3_codes.log 2>&1
import json, sys
sys.path.insert(0, "tools")
import bench_extras
r = bench_extras.graph_c3(0, 10, 3, cpu=False)
json.dump(r, open("gpurun_out/c3_codes.json", "w"), indent=1)
for row in r["sweep"]: print("fp32", {k: (round(v, 4) if isinstance(v, float) else v) for k, v in row.items()})
print("chosen", r["at_recall_0.95"])
print("compressed", json.dumps(r["compressed"])[:3000])
PY
tail -20 gpurun_out/c3_codes.log | cut -c1-600
