"""GPU: K5 call time for small batches through expann_gpu_query — one-warp-per-query form against the
one-CTA-per-query latency form (option graph_cta_batch) with one or two vertices expanded per hop (graph_cta_width);
recall@10 and distance evaluations per query of each form. 1M x 128 low-rank SIFT-shape, GPU-built graph."""
import os, sys, time, json
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import expann_b200
from expann_b200 import datasets
n, d, k = 1_000_000, 128, 10
base, q = datasets.make("sift", n, 4096, d, variant="lowrank")
bf = expann_b200.BruteForceEngine(); bf.store_many_vectors(base); bf.build()
truth = bf.query_k_batch(q, k)[0].copy(); bf.close()
e = expann_b200.GraphEngine(); e.store_many_vectors(base); e.build_graph(M=16, M0=32, K=64); e.build()
hq = torch.from_numpy(q).pin_memory().numpy()
FORMS = (("warp", 0, 1), ("cta", 1 << 20, 1), ("cta2", 1 << 20, 2))
def recall(ids): return float(np.mean([len(set(a.tolist()) & set(b.tolist())) / k for a, b in zip(ids, truth)]))
res = {"recall": [], "latency": []}
for ef in (10, 20, 30, 40, 60, 100):
    e.set_ef_search(ef)
    row = {"ef": ef}
    for form, opt, width in FORMS:
        e.set_option("graph_cta_batch", opt); e.set_option("graph_cta_width", width)
        dc0 = e.stats()["distcomps"]
        ids = e.query_k_batch(hq, k)[0]
        row[form] = {"recall": round(recall(ids), 4), "distcomps_per_query": round((e.stats()["distcomps"] - dc0) / len(hq), 1)}
    res["recall"].append(row); print(json.dumps(row), flush=True)
for ef in (40, 30):
    e.set_ef_search(ef)
    for m in (1, 8, 32, 128, 256, 512):
        out = (torch.zeros((m, k), dtype=torch.int64).pin_memory().numpy().view(np.uint64), torch.zeros((m, k), dtype=torch.float32).pin_memory().numpy(),
               torch.zeros((m,), dtype=torch.int64).pin_memory().numpy().view(np.uint64))
        row = {"ef": ef, "m": m}
        for form, opt, width in FORMS:
            e.set_option("graph_cta_batch", opt); e.set_option("graph_cta_width", width)
            reps = 400 if m <= 64 else 100
            for i in range(20): e.query_k_batch(hq[:m], k, out=out)
            t0 = time.perf_counter()
            for i in range(reps):
                o = (i * m) % (4096 - m + 1)
                e.query_k_batch(hq[o:o + m], k, out=out)
            row[form + "_us_per_call"] = round((time.perf_counter() - t0) / reps * 1e6, 1)
        res["latency"].append(row); print(json.dumps(row), flush=True)
os.makedirs("gpurun_out", exist_ok=True)
json.dump(res, open("gpurun_out/graph_latency.json", "w"), indent=1)
e.close()
