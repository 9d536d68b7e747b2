"""GPU: build a graph on the GPU for the C3 shape and sweep the beam width: recall@10, QPS, distcomps/query.
usage: graph_sweep.py [N] [M0] [K] [compressed]"""
import json, os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import expann_b200
from expann_b200 import datasets
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
M0 = int(sys.argv[2]) if len(sys.argv) > 2 else 32
K = int(sys.argv[3]) if len(sys.argv) > 3 else 64
comp = bool(int(sys.argv[4])) if len(sys.argv) > 4 else False
m, k = 10000, 10
base, q = datasets.make("sift", n, m, 128, variant="lowrank")
bf = expann_b200.BruteForceEngine(); bf.store_many_vectors(base); bf.build()
truth = bf.query_k_batch(q, k)[0]; bf.close()
e = expann_b200.GraphEngine(use_compression=comp)
e.store_many_vectors(base)
t0 = time.perf_counter(); e.build_graph(M=M0 // 2, M0=M0, K=K); t_build = time.perf_counter() - t0
e.build()
print(f"n={n} M0={M0} K={K} build {t_build:.1f}s", flush=True)
dq = torch.from_numpy(q).cuda(); keys = torch.empty((m, k), dtype=torch.int64, device="cuda")
st = torch.cuda.current_stream().cuda_stream
rows = []
for ef in (10, 15, 20, 30, 40, 60, 80, 120, 200):
    e.set_ef_search(ef)
    ids = e.query_k_batch(q, k)[0]
    rec = float(np.mean([len(set(a.tolist()) & set(b.tolist())) / k for a, b in zip(ids, truth)]))
    dc0 = e.stats()["distcomps"]
    for _ in range(2): e.query_device(dq.data_ptr(), m, k, keys.data_ptr(), st)
    torch.cuda.synchronize()
    dc1 = e.stats()["distcomps"]
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(3): e.query_device(dq.data_ptr(), m, k, keys.data_ptr(), st)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 3
    dcq = (dc1 - dc0) / (2 * m)
    rows.append(dict(ef=ef, recall=rec, qps=m / ms * 1e3, ms=ms, distcomps_per_query=dcq, gather_GBs=dcq * (128 if comp else 512) * m / ms / 1e6))
    print(rows[-1], flush=True)
json.dump(dict(n=n, M0=M0, K=K, compressed=comp, build_s=t_build, sweep=rows), open(f"gpurun_out/graph_sweep_n{n}_M{M0}_K{K}_c{int(comp)}.json", "w"))
