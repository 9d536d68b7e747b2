"""GPU: small-batch scans (m = 1, 4, 8) back to back for an ncu launch list."""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import expann_b200
from expann_b200 import datasets
n, d, k = int(os.environ.get("N", 1_000_000)), int(os.environ.get("D", 128)), 10
base, q = datasets.make("sift", n, 16, d, variant="iid")
e = expann_b200.BruteForceEngine(); e.set_option("tc", 0); e.store_many_vectors(base); e.build()
dq = torch.from_numpy(q).cuda(); keys = torch.empty((16, k), dtype=torch.int64, device="cuda")
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for m in (1, 4, 8, 16):
    for _ in range(3):
        flush.zero_()
        e.query_device(dq.data_ptr(), m, k, keys.data_ptr(), torch.cuda.current_stream().cuda_stream)
torch.cuda.synchronize()
print("ok", e.stats())
