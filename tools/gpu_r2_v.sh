#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
N=${N:-1000000}
timeout 400 python - > gpurun_out/v.log 2>&1 <<PY
import sys, time
sys.path.insert(0, '.')
import numpy as np, expann_b200
from expann_b200 import datasets
base, q = datasets.make("sift", $N, 512, 128, variant="lowrank")
e = expann_b200.GraphEngine(); e.set_option("graph_inline", 1); e.store_many_vectors(base); e.build_graph(M=16, M0=32, K=64); e.build()
for ef in (40,):
    e.set_ef_search(ef)
    for i in range(200): e.query_k_batch(q[i:i+1], 10)
    t = time.perf_counter()
    for i in range(400): e.query_k_batch(q[i:i+1], 10)
    print("ef", ef, (time.perf_counter() - t) / 400 * 1e6, "us per call")
e.close()
PY
echo rc $?; grep -v "^GIPROF" gpurun_out/v.log | tail -5; grep "^GI" gpurun_out/v.log | tail -9
