"""One-warp-per-query graph search over the neighbourhood blocks (graph_inline_batch) against the row form: identical keys,
and QPS of device-resident batches. usage: graph_blocks_batch_check.py [n] [d] [m]"""
import os, sys, time, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import expann_b200
from expann_b200 import datasets

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1000000
d = int(sys.argv[2]) if len(sys.argv) > 2 else 128
m = int(sys.argv[3]) if len(sys.argv) > 3 else 10000
base, q = datasets.make("sift", n, m, d, variant="lowrank")
e = expann_b200.GraphEngine(); e.set_option("graph_inline", 1)
if os.environ.get("GRAPH_OCC"): e.set_option("graph_occ", int(os.environ["GRAPH_OCC"]))
e.store_many_vectors(base); e.build_graph(M=16, M0=32, K=64); e.build()
dq = torch.from_numpy(q).cuda(); k = 10
keys = {0: torch.empty((m, k), dtype=torch.int64, device="cuda"), 1: torch.empty((m, k), dtype=torch.int64, device="cuda")}
st = torch.cuda.current_stream().cuda_stream
out = {}
for ef in (10, 20, 40, 80, 200):
    e.set_ef_search(ef)
    row = {}
    for b in (0, 1):
        e.set_option("graph_inline_batch", b)
        dc0 = e.stats()["distcomps"]
        e.query_device(dq.data_ptr(), m, k, keys[b].data_ptr(), st)
        torch.cuda.synchronize()
        row["dc%d" % b] = e.stats()["distcomps"] - dc0
        for _ in range(3): e.query_device(dq.data_ptr(), m, k, keys[b].data_ptr(), st)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10): e.query_device(dq.data_ptr(), m, k, keys[b].data_ptr(), st)
        e1.record(); torch.cuda.synchronize()
        row["qps%d" % b] = m / (e0.elapsed_time(e1) / 10) * 1e3
    row["identical"] = bool(torch.equal(keys[0], keys[1])) and row["dc0"] == row["dc1"]
    out[ef] = row
    print(ef, row, flush=True)
print(json.dumps(out))
e.close()
