"""GPU: the reference's call pattern -- one query per call (src/basic_bench.h:82-126) -- through expann_gpu_query."""
import os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import expann_b200
from expann_b200 import datasets
n, d, k = 1_000_000, 128, 10
base, q = datasets.make("sift", n, 2000, d, variant="lowrank")
for name in ("bf", "graph"):
    if name == "bf":
        e = expann_b200.BruteForceEngine(); e.store_many_vectors(base); e.build()
    else:
        e = expann_b200.GraphEngine(); e.store_many_vectors(base); e.build_graph(M=16, M0=32, K=64); e.build(); e.set_ef_search(40)
    hq = torch.from_numpy(q).pin_memory().numpy()
    out = (torch.zeros((1, k), dtype=torch.int64).pin_memory().numpy().view(np.uint64), torch.zeros((1, k), dtype=torch.float32).pin_memory().numpy(),
           torch.zeros((1,), dtype=torch.int64).pin_memory().numpy().view(np.uint64))
    for mode in ("pageable", "pinned"):
        for i in range(50): e.query_k_batch(q[i:i + 1], k)
        t0 = time.perf_counter()
        for i in range(2000):
            if mode == "pinned": e.query_k_batch(hq[i:i + 1], k, out=out)
            else: e.query_k_batch(q[i:i + 1], k)
        dt = (time.perf_counter() - t0) / 2000
        print({"engine": name, "buffers": mode, "us_per_query": round(dt * 1e6, 1), "qps_sequential": round(1 / dt)}, flush=True)
    e.close()
