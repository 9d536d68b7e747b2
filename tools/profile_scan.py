"""GPU: a few scan launches for ncu. usage: profile_scan.py M QTILE NSPLIT [ITERS]"""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import expann_b200
from expann_b200 import datasets
m, qt, ns = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
iters = int(sys.argv[4]) if len(sys.argv) > 4 else 3
n, d, k = int(os.environ.get("N", 1_000_000)), int(os.environ.get("D", 128)), 10
base, q = datasets.make("sift", n, max(m, 16), d, variant="iid")
e = expann_b200.BruteForceEngine(); e.store_many_vectors(base); e.build()
e.set_option("qtile", qt); e.set_option("nsplit", ns); e.set_option("tc_pair", int(os.environ.get("PAIR", 1))); e.set_option("tc_boot", int(os.environ.get("BOOT", 1))); e.set_option("tc_min_batch", int(os.environ.get("TCMIN", 5)))
dq = torch.from_numpy(q).cuda(); keys = torch.empty((max(m, 16), k), dtype=torch.int64, device="cuda")
for _ in range(iters):
    e.query_device(dq.data_ptr(), m, k, keys.data_ptr(), torch.cuda.current_stream().cuda_stream)
torch.cuda.synchronize()
print("ok", e.stats())
