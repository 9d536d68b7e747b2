"""GPU: graph search QPS at ef = 20 / 40 / 80 for each residency setting (graph_occ = CTAs of 4 warps per SM)."""
import os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import expann_b200
from expann_b200 import datasets
n, m, k = int(os.environ.get("N", 1_000_000)), 10000, 10
base, q = datasets.make("sift", n, m, 128, variant="lowrank")
bf = expann_b200.BruteForceEngine(); bf.store_many_vectors(base); bf.build()
truth = bf.query_k_batch(q, k)[0]; bf.close()
dq = torch.from_numpy(q).cuda(); keys = torch.empty((m, k), dtype=torch.int64, device="cuda")
st = torch.cuda.current_stream().cuda_stream
path = "/tmp/occ_graph.idx"
for occ in [int(x) for x in os.environ.get("OCC", "2,4,6,8").split(",")]:
    e = expann_b200.GraphEngine(use_compression=bool(int(os.environ.get("COMP", 0))))
    e.set_option("graph_occ", occ)
    e.store_many_vectors(base)
    e.build_graph(M=16, M0=32, K=64)
    e.build()
    for ef in (20, 40, 80):
        e.set_ef_search(ef)
        ids = e.query_k_batch(q, k)[0]
        rec = float(np.mean([len(set(a.tolist()) & set(b.tolist())) / k for a, b in zip(ids, truth)]))
        for _ in range(2): e.query_device(dq.data_ptr(), m, k, keys.data_ptr(), st)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(5): e.query_device(dq.data_ptr(), m, k, keys.data_ptr(), st)
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 5
        print({"occ": occ, "ef": ef, "recall": round(rec, 4), "ms": round(ms, 3), "qps": round(m / ms * 1e3)}, flush=True)
    e.close()
