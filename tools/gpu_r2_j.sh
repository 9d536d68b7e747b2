#!/bin/bash
# round-2 GPU call J (1 GPU): small-batch filter variants (one GPU's share of C2 at 8 GPUs with the queries split), C3 at the reference's grid
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
for boot in 0 2; do for ns in 0 14 20 40; do
  echo "boot=$boot nsplit=$ns"; N=1000000 BOOT=$boot NSPLIT=$ns TAG=m1250_b${boot}_n$ns timeout 120 python tools/time_tc.py 1250,2500 2>&1 | tail -2 | cut -c1-200
done; done
echo "== C3 at M=60 M0=120"
timeout 900 python - <<'PY'
import sys, time, json, numpy as np, torch
sys.path.insert(0, '.')
import expann_b200
from expann_b200 import datasets
n, m, k, d = 1_000_000, 10_000, 10, 128
base, q = datasets.make("sift", n, m, d, variant="lowrank")
bf = expann_b200.BruteForceEngine(); bf.store_many_vectors(base); bf.build(); truth = bf.query_k_batch(q, k)[0]; bf.close()
out = {}
for (M, M0, K) in ((16, 32, 64), (60, 120, 127)):
    e = expann_b200.GraphEngine()
    e.store_many_vectors(base)
    t0 = time.perf_counter()
    try:
        e.build_graph(M=M, M0=M0, K=K); e.build()
    except Exception as ex:
        print("build failed", M, M0, K, ex); continue
    bs = time.perf_counter() - t0
    dq = torch.from_numpy(q).cuda(); keys = torch.empty((m, k), dtype=torch.int64, device="cuda"); st = torch.cuda.current_stream().cuda_stream
    rows = []
    for ef in (10, 15, 20, 30, 40, 60):
        e.set_ef_search(ef)
        ids = e.query_k_batch(q, k)[0]
        rec = float(np.mean([len(set(a.tolist()) & set(b.tolist())) / k for a, b in zip(ids, truth)]))
        for _ in range(2): e.query_device(dq.data_ptr(), m, k, keys.data_ptr(), st)
        torch.cuda.synchronize(); e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s0 = e.stats(); e0.record()
        for _ in range(5): e.query_device(dq.data_ptr(), m, k, keys.data_ptr(), st)
        e1.record(); torch.cuda.synchronize(); s1 = e.stats()
        rows.append(dict(ef=ef, recall=rec, qps=m * 5 / (e0.elapsed_time(e1) * 1e-3), distcomps=(s1["distcomps"] - s0["distcomps"]) / (5 * m)))
        print(M, M0, K, rows[-1], flush=True)
    out[f"M{M}_M0{M0}_K{K}"] = dict(build_s=bs, sweep=rows)
    e.close()
json.dump(out, open("gpurun_out/r2_graph_refgrid.json", "w"), indent=1)
PY
