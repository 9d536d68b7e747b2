"""GPU: pair form of K3 (tc_pair=1) against the SIMT scan of the same library; small, then C2 timing."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import expann_b200
def run(base, q, k, pair, tc=1, metric=0):
    e = expann_b200.BruteForceEngine(metric=metric); e.set_option("tc", tc); e.set_option("tc_pair", pair)
    e.store_many_vectors(base); e.build()
    out = e.query_k_batch(q, k); st = e.stats(); e.close(); return out, st
rng = np.random.default_rng(3)
for (n, m, d, real) in [(20011, 777, 128, False), (300001, 1300, 128, False), (150000, 700, 96, True)]:
    base = rng.standard_normal((n, d)).astype(np.float32) if real else rng.integers(0, 256, (n, d)).astype(np.float32)
    q = rng.standard_normal((m, d)).astype(np.float32) if real else rng.integers(0, 256, (m, d)).astype(np.float32)
    for metric in (0, 1):
        (i1, d1, c1), s1 = run(base, q, 10, 1, metric=metric)
        (i0, d0, c0), s0 = run(base, q, 10, 0, tc=0, metric=metric)
        ok = np.array_equal(i1, i0) and np.array_equal(d1, d0)
        print(n, m, d, "metric", metric, "pair ok" if ok else "PAIR MISMATCH", "tc_launches", s1["tc_launches"], "fallbacks", s1["tc_fallbacks"],
              "qtile", s1["last_qtile"], "nsplit", s1["last_nsplit"], flush=True)
        if not ok:
            bad = np.where((i1 != i0).any(1))[0]
            print("  mismatching queries:", len(bad), bad[:10], i1[bad[0]], i0[bad[0]], flush=True)
