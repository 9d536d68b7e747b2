"""GPU: randomized differential test of the tensor-core path (all forms) against the SIMT scan of the same library."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import expann_b200
seed = int(os.environ.get("SEED", 1)); budget = float(os.environ.get("BUDGET", 240))
rng = np.random.default_rng(seed)
t0 = time.time(); it = 0; bad = 0
while time.time() - t0 < budget:
    it += 1
    d = int(rng.choice([8, 20, 32, 64, 96, 100, 128, 200, 256, 384, 512, 960]))
    n = int(rng.choice([1, 7, 255, 256, 257, 1000, 5000, 33000, 70000, 200000]))
    if n * d > 40_000_000: n = 40_000_000 // d
    m = int(rng.choice([5, 9, 16, 100, 128, 129, 255, 256, 257, 600, 1500, 9000 if n * d < 10_000_000 else 300]))
    k = int(rng.choice([1, 2, 5, 10, 16]))
    kind = rng.choice(["u8", "small", "real", "unit", "dup", "sorted"])
    metric = int(rng.integers(0, 2))
    if kind == "u8": base = rng.integers(0, 256, (n, d)).astype(np.float32); q = rng.integers(0, 256, (m, d)).astype(np.float32)
    elif kind == "small": base = rng.integers(-8, 9, (n, d)).astype(np.float32); q = rng.integers(-8, 9, (m, d)).astype(np.float32)
    elif kind == "real": base = rng.standard_normal((n, d)).astype(np.float32); q = rng.standard_normal((m, d)).astype(np.float32)
    elif kind == "unit": base = rng.random((n, d), dtype=np.float32); q = rng.random((m, d), dtype=np.float32)
    elif kind == "dup":
        pool = rng.integers(0, 256, (max(1, n // 50), d)).astype(np.float32); base = pool[rng.integers(0, len(pool), n)]; q = pool[rng.integers(0, len(pool), m)].copy()
    else:
        base = rng.integers(0, 256, (n, d)).astype(np.float32); c = np.rint(base.mean(0)); base = np.ascontiguousarray(base[np.argsort(-((base - c) ** 2).sum(1))])
        q = np.tile(c, (m, 1)).astype(np.float32)
    opts = {"tc_pair": int(rng.integers(0, 2)), "tc_boot": int(rng.choice([0, 1, 2])), "tc_astream": int(rng.choice([1, 1, 2]))}
    res = []
    for tc in (1, 0):
        e = expann_b200.BruteForceEngine(metric=metric); e.set_option("tc", tc)
        if tc:
            for o, v in opts.items(): e.set_option(o, v)
        e.store_many_vectors(base); e.build()
        out = e.query_k_batch(q, k); st = e.stats(); e.close(); res.append((out, st))
    (a, sa), (b, sb) = res
    ok = np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1]) and np.array_equal(a[2], b[2])
    print(it, "ok" if ok else "MISMATCH", dict(n=n, m=m, d=d, k=k, kind=str(kind), metric=metric, **opts), "mode", sa["tc_mode"], "tcl", sa["tc_launches"], "fb", sa["tc_fallbacks"], "qt", sa["last_qtile"], flush=True)
    if not ok:
        bad += 1
        w = np.where((a[0] != b[0]).any(1) | (a[1] != b[1]).any(1))[0]
        print("   rows:", len(w), w[:5], a[0][w[0]], b[0][w[0]], a[1][w[0]], b[1][w[0]], flush=True)
print("done", it, "cases,", bad, "mismatches")
