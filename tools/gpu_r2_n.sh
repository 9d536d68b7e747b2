#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
cat > /tmp/ncu_blocks.py <<'PY'
import sys
sys.path.insert(0, '.')
import numpy as np, torch, expann_b200
from expann_b200 import datasets
base, q = datasets.make("sift", 1000000, 10000, 128, variant="lowrank")
e = expann_b200.GraphEngine(); e.store_many_vectors(base); e.build_graph(M=16, M0=32, K=64); e.build(); e.set_ef_search(40)
dq = torch.from_numpy(q).cuda(); keys = torch.empty((10000, 10), dtype=torch.int64, device="cuda")
st = torch.cuda.current_stream().cuda_stream
for _ in range(6): e.query_device(dq.data_ptr(), 10000, 10, keys.data_ptr(), st)
torch.cuda.synchronize(); e.close()
PY
timeout 600 ncu --set full --clock-control none --import-source on -k regex:graph_search_blocks --launch-skip 3 -c 1 -f -o gpurun_out/r2_graph_blocks python /tmp/ncu_blocks.py > gpurun_out/n_ncu.log 2>&1; echo "ncu rc $?"; tail -2 gpurun_out/n_ncu.log
