"""GPU: per-step time of a multi-device handle against the single-device handle on the same data (device list from
DEVICES, e.g. "0,0" = two shards on one GPU, "0,1" = two GPUs); ROWG = row groups, M = queries."""
import os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import expann_b200
from expann_b200 import datasets
n, d, k = int(os.environ.get("N", 1_000_000)), 128, 10
devs = [int(x) for x in os.environ.get("DEVICES", "0,0").split(",")]
rowg = int(os.environ.get("ROWG", 0))
base, q = datasets.make("sift", n, 10000, d, variant="iid")
dev0 = torch.device("cuda", devs[0])
torch.cuda.set_device(dev0)
dq = torch.from_numpy(q).to(dev0)
keys = torch.empty((10000, k), dtype=torch.int64, device=dev0)
use_torch_stream = int(os.environ.get("TSTREAM", 1))
side = torch.cuda.Stream(device=dev0)
for label, kw in (("single", {}), ("multi", dict(devices=devs, row_groups=rowg))):
    e = expann_b200.BruteForceEngine(**kw); e.store_many_vectors(base); e.build()
    for opt in os.environ.get("OPTS", "").split(","):
        if opt:
            nm, v = opt.split("=")
            if label == "multi" or not nm.startswith("mgpu"):
                e.set_option(nm, int(v))
    for m in [int(x) for x in os.environ.get("MS", "10000,5000").split(",")]:
        for sname, st in (("default", torch.cuda.current_stream(dev0)), ("side", side)):
            with torch.cuda.stream(st):
                for _ in range(3): e.query_device(dq.data_ptr(), m, k, keys.data_ptr(), st.cuda_stream)
                torch.cuda.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                t0 = time.perf_counter(); e0.record(st)
                for _ in range(20): e.query_device(dq.data_ptr(), m, k, keys.data_ptr(), st.cuda_stream)
                t1 = time.perf_counter(); e1.record(st); torch.cuda.synchronize(); t2 = time.perf_counter()
            stt = e.stats()
            print(label, devs if label == "multi" else "", "m", m, "stream", sname, "ms/step %.3f" % (e0.elapsed_time(e1) / 20), "host enqueue ms/step %.3f" % ((t1 - t0) * 50),
                  "wall %.3f" % ((t2 - t0) * 50), "rg", stt["row_groups"], "qg", stt["query_groups"], flush=True)
    e.close()
