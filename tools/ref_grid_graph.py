"""C3 at the reference's OWN grid (M = 60, M0 = 120, ef_construction = 480, src/bench_runner.h:132-162) on a graph BUILT BY THE
UNMODIFIED REFERENCE (oracle/_ref, sequential CPU insert), handed over through its index file: recall@10 and time per query of
the reference engine (one thread, its own loop) and of the GPU engine (batch and one query per call) on the SAME graph.
usage: ref_grid_graph.py [n] [m]"""
import os, sys, time, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import expann_b200
from expann_b200 import datasets
from oracle import pyoracle

n = int(sys.argv[1]) if len(sys.argv) > 1 else 20000
m = int(sys.argv[2]) if len(sys.argv) > 2 else 2000
k = 10
base, q = datasets.make("sift", n, m, 128, seed=11, variant="lowrank")
ref = pyoracle.Reference(128)
G = ref.graph(M=60, M0=120, ef_construction=480)
G.store(base)
t = time.perf_counter(); G.build(); build_s = time.perf_counter() - t
path = "/tmp/expann_ref_grid.index"
G.save(path)
bf = expann_b200.BruteForceEngine(); bf.store_many_vectors(base); bf.build(); truth = bf.query_k_batch(q, k)[0]; bf.close()
def recall(ids): return float(np.mean([len(set(a.tolist()) & set(b.tolist())) / k for a, b in zip(ids, truth)]))
e = expann_b200.GraphEngine(); e.set_option("graph_inline", 1); e.load_index(path); e.build()
out = {"workload": "SIFT-shape synthetic (low-rank, u8-valued) %dx128, %d queries, graph built by the unmodified reference at its shipped grid M=60 M0=120 ef_construction=480" % (n, m),
       "reference_build_s": build_s, "info": G.info(), "sweep": []}
for ef in (10, 20, 40, 80, 160):
    G.set_ef(ef)
    G.query(q[:100], k)
    t = time.perf_counter(); rid, _ = G.query(q, k); ref_us = (time.perf_counter() - t) / m * 1e6
    e.set_ef_search(ef)
    ids = e.query_k_batch(q, k)[0]
    for _ in range(3): e.query_k_batch(q, k)
    t = time.perf_counter()
    for _ in range(10): e.query_k_batch(q, k)
    batch_us = (time.perf_counter() - t) / 10 / m * 1e6
    for i in range(50): e.query_k_batch(q[i:i + 1], k)
    t = time.perf_counter()
    for i in range(500): e.query_k_batch(q[i:i + 1], k)
    one_us = (time.perf_counter() - t) / 500 * 1e6
    same = float(np.mean([set(a.tolist()) == set(b.tolist()) for a, b in zip(ids, rid)]))
    row = {"ef": ef, "reference_recall": recall(rid), "gpu_recall": recall(ids), "same_result_sets": same, "reference_us_per_query_1thread": ref_us,
           "gpu_us_per_query_in_a_batch_of_%d_host_call" % m: batch_us, "gpu_us_per_one_query_call": one_us}
    out["sweep"].append(row)
    print(row, flush=True)
print(json.dumps(out))
e.close()
