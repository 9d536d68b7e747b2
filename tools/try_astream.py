"""GPU: streamed-A pair form of K3 (long rows: d > 128 real-valued) against the SIMT scan of the same library; then timing."""
import os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import expann_b200
def run(base, q, k, tc=1, metric=0, astream=1):
    e = expann_b200.BruteForceEngine(metric=metric); e.set_option("tc", tc); e.set_option("tc_astream", astream)
    e.store_many_vectors(base); e.build()
    out = e.query_k_batch(q, k); st = e.stats(); e.close(); return out, st
rng = np.random.default_rng(3)
cases = [(20011, 300, 128, "u8", 2), (50000, 300, 96, "real", 2), (50000, 300, 960, "unit", 1), (40000, 100, 256, "real", 1),
         (30000, 600, 512, "u8", 1), (3000, 20, 960, "unit", 1)]
if os.environ.get("CASES", "1") == "1":
    for (n, m, d, kind, astream) in cases:
        if kind == "u8": base = rng.integers(0, 256, (n, d)).astype(np.float32); q = rng.integers(0, 256, (m, d)).astype(np.float32)
        elif kind == "real": base = rng.standard_normal((n, d)).astype(np.float32); q = rng.standard_normal((m, d)).astype(np.float32)
        else: base = rng.random((n, d), dtype=np.float32); q = rng.random((m, d), dtype=np.float32)
        for metric in (0, 1):
            (i1, d1, c1), s1 = run(base, q, 10, metric=metric, astream=astream)
            (i0, d0, c0), s0 = run(base, q, 10, tc=0, metric=metric)
            ok = np.array_equal(i1, i0) and np.array_equal(d1, d0)
            print(n, m, d, kind, "metric", metric, "ok" if ok else "MISMATCH", "mode", s1["tc_mode"], "tc_launches", s1["tc_launches"], "fallbacks", s1["tc_fallbacks"],
                  "qtile", s1["last_qtile"], "nsplit", s1["last_nsplit"], flush=True)
            if not ok:
                bad = np.where((i1 != i0).any(1))[0]
                print("  mismatching queries:", len(bad), bad[:10], i1[bad[0]] if len(bad) else "", i0[bad[0]] if len(bad) else "", flush=True)
# timing: GIST-shape 1M x 960 IP, 1000 queries
n, d, m = int(os.environ.get("N", 1_000_000)), int(os.environ.get("D", 960)), int(os.environ.get("M", 1000))
metric = int(os.environ.get("METRIC", 1))
g = torch.Generator(device="cuda"); g.manual_seed(1)
base = torch.rand((n, d), generator=g, device="cuda").cpu().numpy(); q = torch.rand((m, d), generator=g, device="cuda").cpu().numpy()
for tc in (1, 0):
    e = expann_b200.BruteForceEngine(metric=metric); e.set_option("tc", tc); e.store_many_vectors(base); e.build()
    dq = torch.from_numpy(q).cuda(); keys = torch.empty((m, 10), dtype=torch.int64, device="cuda")
    stm = torch.cuda.current_stream().cuda_stream
    for _ in range(2): e.query_device(dq.data_ptr(), m, 10, keys.data_ptr(), stm)
    torch.cuda.synchronize(); e.set_timing(True); st0 = e.stats()
    t0 = torch.cuda.Event(enable_timing=True); t1 = torch.cuda.Event(enable_timing=True)
    reps = 5 if tc else 2
    t0.record()
    for _ in range(reps): e.query_device(dq.data_ptr(), m, 10, keys.data_ptr(), stm)
    t1.record(); torch.cuda.synchronize(); st = e.stats()
    ms = t0.elapsed_time(t1) / reps
    hot = (st["hot_time_ns"] - st0["hot_time_ns"]) / 1e6 / max(1, st["hot_timed_launches"] - st0["hot_timed_launches"])
    kd = (d + 63) // 64 * 64 * (3 if st["tc_mode"] == 2 else 1)
    print({"tc": tc, "n": n, "d": d, "m": m, "ms": ms, "qps": m / ms * 1e3, "hot_ms": hot, "mode": st["tc_mode"], "tflops": 2.0 * m * n * kd / (hot * 1e-3) / 1e12 if tc else None,
           "fallbacks": st["tc_fallbacks"], "qtile": st["last_qtile"], "nsplit": st["last_nsplit"]}, flush=True)
    if tc: k1 = keys.clone()
    else: print("same ids as SIMT:", bool((k1 == keys).all()))
    e.close()
