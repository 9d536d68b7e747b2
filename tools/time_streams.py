"""GPU: a single-device brute-force handle called on ONE stream against the same calls alternating over TWO streams (the
handle keeps two sets of work buffers, csrc/api.cu scratch_for_stream): ms per call."""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import expann_b200
from expann_b200 import datasets
n, d, k = int(os.environ.get("N", 1_000_000)), 128, 10
base, q = datasets.make("sift", n, 10000, d, variant="iid")
e = expann_b200.BruteForceEngine(); e.store_many_vectors(base); e.build()
dq = torch.from_numpy(q).cuda()
keys = [torch.empty((10000, k), dtype=torch.int64, device="cuda") for _ in range(2)]
streams = [torch.cuda.Stream(), torch.cuda.Stream()]
main = torch.cuda.current_stream()
for m in [int(x) for x in os.environ.get("MS", "1250,2500,5000,10000").split(",")]:
    for mode in ("one", "two"):
        def run(reps):
            for i in range(reps):
                s = streams[i & 1] if mode == "two" else streams[0]
                e.query_device(dq.data_ptr(), m, k, keys[i & 1].data_ptr(), s.cuda_stream)
        run(6); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(main)
        for s in streams: s.wait_stream(main)
        run(40)
        for s in streams: main.wait_stream(s)
        e1.record(main); torch.cuda.synchronize()
        print("m", m, mode, "stream(s): ms/call %.4f" % (e0.elapsed_time(e1) / 40), flush=True)
