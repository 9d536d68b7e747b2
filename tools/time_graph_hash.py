"""GPU: batch form of the graph search (one warp per query) with the visited set in shared memory (graph_hash = 1) against
the per-vertex bytes in HBM (0): QPS, recall and distance evaluations per query over a beam-width sweep, C3 data."""
import os, sys, json
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import expann_b200
from expann_b200 import datasets
n, m, k, d = 1_000_000, 10_000, 10, 128
base, q = datasets.make("sift", n, m, d, variant="lowrank")
bf = expann_b200.BruteForceEngine(); bf.store_many_vectors(base); bf.build(); truth = bf.query_k_batch(q, k)[0]; bf.close()
rows = []
for comp in (False, True):
    e = expann_b200.GraphEngine(use_compression=comp); e.store_many_vectors(base); e.build_graph(M=16, M0=32, K=64); e.build()
    dq = torch.from_numpy(q).cuda(); keys = torch.empty((m, k), dtype=torch.int64, device="cuda"); st = torch.cuda.current_stream().cuda_stream
    for ef in (20, 40, 60, 120):
        e.set_ef_search(ef); ref = None
        for hsh in (0, 1):
            e.set_option("graph_hash", hsh)
            ids = e.query_k_batch(q, k)[0]
            rec = float(np.mean([len(set(a.tolist()) & set(b.tolist())) / k for a, b in zip(ids, truth)]))
            for _ in range(2): e.query_device(dq.data_ptr(), m, k, keys.data_ptr(), st)
            torch.cuda.synchronize(); s0 = e.stats(); e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(5): e.query_device(dq.data_ptr(), m, k, keys.data_ptr(), st)
            e1.record(); torch.cuda.synchronize(); s1 = e.stats()
            same = True if ref is None else bool(np.array_equal(ref, ids))
            ref = ids
            rows.append(dict(compressed=comp, ef=ef, graph_hash=hsh, recall=rec, qps=m * 5 / (e0.elapsed_time(e1) * 1e-3),
                             distcomps=(s1["distcomps"] - s0["distcomps"]) / (5 * m), same_ids_as_hbm_bytes=same))
            print(rows[-1], flush=True)
    e.close()
os.makedirs("gpurun_out", exist_ok=True); json.dump(rows, open("gpurun_out/r2_graph_hash.json", "w"), indent=1)
