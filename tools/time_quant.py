"""GPU: C5 code scan (K6) timing against the row-split count."""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import expann_b200
n, d, m, k = int(os.environ.get("N", 1_000_000)), 960, int(os.environ.get("M", 1000)), 10
g = torch.Generator(device="cuda"); g.manual_seed(1)
base = torch.rand((n, d), generator=g, device="cuda").cpu().numpy(); q = torch.rand((m, d), generator=g, device="cuda").cpu().numpy()
e = expann_b200.QuantScanEngine(quantizer=1, kprime=100, metric=1)
e.store_many_vectors(base); e.build()
dq = torch.from_numpy(q).cuda(); keys = torch.empty((m, k), dtype=torch.int64, device="cuda")
st = torch.cuda.current_stream().cuda_stream
ref = None
for ns in [int(x) for x in os.environ.get("NSPLIT", "0,1,2,4,8").split(",")]:
    e.set_option("nsplit", ns)
    for _ in range(2): e.query_device(dq.data_ptr(), m, k, keys.data_ptr(), st)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(3): e.query_device(dq.data_ptr(), m, k, keys.data_ptr(), st)
    e1.record(); torch.cuda.synchronize()
    same = True if ref is None else bool((ref == keys).all())
    if ref is None: ref = keys.clone()
    print({"nsplit_opt": ns, "nsplit": e.stats()["last_nsplit"], "ms": e0.elapsed_time(e1) / 3, "same_as_first": same}, flush=True)
