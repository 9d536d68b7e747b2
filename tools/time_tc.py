"""GPU: time the large-batch brute-force path (K3 filter + select + K4 rerank) on the C2 base; prints per-m timings."""
import json, os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import expann_b200
from expann_b200 import datasets
n, d, k = int(os.environ.get("N", 1_000_000)), int(os.environ.get("D", 128)), int(os.environ.get("K", 10))
gen, variant = os.environ.get("GEN", "sift"), os.environ.get("VARIANT", "iid")
ms_list = [int(x) for x in (sys.argv[1] if len(sys.argv) > 1 else "10000").split(",")]
base, q = datasets.make(gen, n, max(ms_list), d, variant=variant)
e = expann_b200.BruteForceEngine(metric=int(os.environ.get("METRIC", 0))); e.set_option("tc_real", int(os.environ.get("REAL", 1))); e.set_option("tc_tf32", int(os.environ.get("TF32", 1))); e.store_many_vectors(base); e.build()
e.set_option("tc_v1", int(os.environ.get("V1", 0))); e.set_option("nsplit", int(os.environ.get("NSPLIT", 0)))
e.set_option("tc_min_batch", int(os.environ.get("TCMIN", 5))); e.set_option("tc_pair", int(os.environ.get("PAIR", 1))); e.set_option("tc_boot", int(os.environ.get("BOOT", 1))); e.set_option("tc_boot_tiles", int(os.environ.get("BOOTTILES", 128)))
dq = torch.from_numpy(q).cuda(); keys = torch.empty((max(ms_list), k), dtype=torch.int64, device="cuda")
st = torch.cuda.current_stream().cuda_stream
rows = []
for m in ms_list:
    for _ in range(3): e.query_device(dq.data_ptr(), m, k, keys.data_ptr(), st)
    torch.cuda.synchronize(); e.set_timing(True); s0 = e.stats(); ts = []
    for _ in range(5):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); e.query_device(dq.data_ptr(), m, k, keys.data_ptr(), st); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    s1 = e.stats(); e.set_timing(False)
    hot = (s1["hot_time_ns"] - s0["hot_time_ns"]) / 1e6 / max(1, s1["hot_timed_launches"] - s0["hot_timed_launches"])
    kd = (d + 31) // 32 * 32 if s1["tc_mode"] == 4 else (d + 63) // 64 * 64 * (3 if s1["tc_mode"] == 2 else 1)
    row = dict(m=m, ms=float(np.median(ts)), qps=m / float(np.median(ts)) * 1e3, hot_ms=hot, tc_mode=int(s1["tc_mode"]),
               tflops=2.0 * m * n * kd / (hot * 1e-3) / 1e12, qtile=int(s1["last_qtile"]), nsplit=int(s1["last_nsplit"]),
               fallbacks=int(s1["tc_fallbacks"]), n=n, d=d, gen=gen, lib=os.path.basename(os.environ.get("EXPANN_GPU_LIB", "default")))
    rows.append(row); print(row, flush=True)
os.makedirs("gpurun_out", exist_ok=True)
json.dump(rows, open("gpurun_out/time_tc_%s.json" % os.environ.get("TAG", "x"), "w"))
