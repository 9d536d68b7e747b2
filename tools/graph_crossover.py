"""per-call time of the graph engine by batch size: one CTA per query (blocks) against one warp per query (blocks)"""
import os, sys, time, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import expann_b200
from expann_b200 import datasets
base, q = datasets.make("sift", 1000000, 4096, 128, variant="lowrank")
e = expann_b200.GraphEngine(); e.store_many_vectors(base); e.build_graph(M=16, M0=32, K=64); e.build(); e.set_ef_search(40)
k = 10
hq = torch.from_numpy(q).pin_memory().numpy()
out = {}
for m in (1, 8, 32, 64, 128, 148, 296, 444, 592, 1024, 2048):
    outp = (torch.empty((m, k), dtype=torch.int64).pin_memory().numpy().view(np.uint64), torch.empty((m, k), dtype=torch.float32).pin_memory().numpy(),
            torch.empty((m,), dtype=torch.int64).pin_memory().numpy().view(np.uint64))
    row = {}
    for name, cta in (("cta", 1 << 20), ("warp", 0)):
        e.set_option("graph_cta_batch", cta)
        for i in range(10): e.query_k_batch(hq[:m], k, out=outp)
        R = 100
        t = time.perf_counter()
        for i in range(R):
            o = (i * m) % (4096 - m + 1)
            e.query_k_batch(hq[o:o + m], k, out=outp)
        row[name] = (time.perf_counter() - t) / R * 1e6
    out[m] = row
    print(m, {a: round(b, 1) for a, b in row.items()}, flush=True)
print(json.dumps(out))
e.close()
