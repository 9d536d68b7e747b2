"""GPU: time the scan (K1) + merge (K7) for several query counts / tiles on the C2 base (1M x 128)."""
import json, os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import expann_b200
from expann_b200 import datasets

n, d, k = int(os.environ.get("N", 1_000_000)), int(os.environ.get("D", 128)), 10
base, q = datasets.make("sift", n, 10000, d, variant="iid")
e = expann_b200.BruteForceEngine()
e.store_many_vectors(base); e.build()
dq = torch.from_numpy(q).cuda()
keys = torch.empty((10000, k), dtype=torch.int64, device="cuda")
st = torch.cuda.current_stream().cuda_stream
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
rows = []
for m, qt, ns in [(1,1,0),(4,4,0),(8,8,0),(8,8,296),(8,8,592),(8,4,0),(16,8,0),(64,8,0),(256,8,0),(1024,8,0),(10000,8,0),(10000,4,0),(10000,8,4),(10000,8,8)]:
    e.set_option("qtile", qt); e.set_option("nsplit", ns)
    for _ in range(3): e.query_device(dq.data_ptr(), m, k, keys.data_ptr(), st)
    torch.cuda.synchronize()
    ts = []
    for _ in range(5):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); e.query_device(dq.data_ptr(), m, k, keys.data_ptr(), st); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ms = float(np.median(ts)); s = e.stats()
    passes = -(-m // qt)
    rows.append(dict(m=m, qtile=qt, nsplit=int(s["last_nsplit"]), ms=ms, qps=m / ms * 1e3, alg_GBs=passes * n * d * 4 / ms / 1e6))
    print(rows[-1], flush=True)
json.dump(rows, open("gpurun_out/sweep_scan.json", "w"))
