"""GPU: K5 call time for small batches, one-warp-per-query form against the one-CTA-per-query form
(option graph_cta_batch), through expann_gpu_query with page-locked buffers. 1M x 128 low-rank, ef = 40."""
import os, sys, time, json
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import expann_b200
from expann_b200 import datasets
n, d, k = 1_000_000, 128, 10
base, q = datasets.make("sift", n, 4096, d, variant="lowrank")
e = expann_b200.GraphEngine(); e.store_many_vectors(base); e.build_graph(M=16, M0=32, K=64); e.build(); e.set_ef_search(40)
hq = torch.from_numpy(q).pin_memory().numpy()
res = []
for m in (1, 2, 4, 8, 16, 32, 64, 128, 256, 512, 1024):
    out = (torch.zeros((m, k), dtype=torch.int64).pin_memory().numpy().view(np.uint64), torch.zeros((m, k), dtype=torch.float32).pin_memory().numpy(),
           torch.zeros((m,), dtype=torch.int64).pin_memory().numpy().view(np.uint64))
    row = {"m": m}
    ref = None
    for form, opt in (("warp", 0), ("cta", 1 << 20)):
        e.set_option("graph_cta_batch", opt)
        reps = 400 if m <= 64 else 100
        for i in range(20): e.query_k_batch(hq[:m], k, out=out)
        t0 = time.perf_counter()
        for i in range(reps):
            o = (i * m) % (4096 - m + 1)
            e.query_k_batch(hq[o:o + m], k, out=out)
        dt = (time.perf_counter() - t0) / reps
        row[form + "_us_per_call"] = round(dt * 1e6, 1)
        ids = e.query_k_batch(hq[:m], k)[0].copy()
        if ref is None: ref = ids
        else: row["same_ids"] = bool(np.array_equal(ref, ids))
    res.append(row)
    print(json.dumps(row), flush=True)
os.makedirs("gpurun_out", exist_ok=True)
json.dump(res, open("gpurun_out/graph_latency.json", "w"), indent=1)
e.close()
