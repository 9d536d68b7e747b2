"""GPU: bounded randomized differential test of the tensor-core path (every form of K3: single CTA, pair, seeded,
streamed query operand; operand formats exact / one-term / split) against the SIMT scan of the same library, which the
other tests pin to the oracle. Fixed seed; stops after a time budget (tools/fuzz_tc.py is the open-ended version)."""
import time

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_fuzz_tc_vs_simt():
    import expann_b200
    rng = np.random.default_rng(20261105)
    t0, cases, paths = time.time(), 0, set()
    while time.time() - t0 < 40.0 and cases < 400:
        d = int(rng.choice([8, 20, 32, 64, 96, 100, 128, 200, 256, 384, 512, 960]))
        n = int(rng.choice([1, 7, 255, 256, 257, 1000, 5000, 33000, 70000]))
        if n * d > 10_000_000:
            n = 10_000_000 // d
        m = int(rng.choice([5, 9, 16, 100, 128, 129, 255, 256, 257, 600, 1500]))
        k = int(rng.choice([1, 2, 5, 10, 16]))
        kind = str(rng.choice(["u8", "small", "signed", "real", "unit", "dup", "sorted"]))
        metric = int(rng.integers(0, 2))
        if kind == "u8":
            base = rng.integers(0, 256, (n, d)).astype(np.float32); q = rng.integers(0, 256, (m, d)).astype(np.float32)
        elif kind == "small":
            base = rng.integers(-8, 9, (n, d)).astype(np.float32); q = rng.integers(-8, 9, (m, d)).astype(np.float32)
        elif kind == "signed":
            base = rng.integers(-255, 256, (n, d)).astype(np.float32); q = rng.integers(-255, 256, (m, d)).astype(np.float32)
        elif kind == "real":
            base = rng.standard_normal((n, d)).astype(np.float32); q = rng.standard_normal((m, d)).astype(np.float32)
        elif kind == "unit":
            base = rng.random((n, d), dtype=np.float32); q = rng.random((m, d), dtype=np.float32)
        elif kind == "dup":
            pool = rng.integers(0, 256, (max(1, n // 50), d)).astype(np.float32)
            base = pool[rng.integers(0, len(pool), n)]; q = pool[rng.integers(0, len(pool), m)].copy()
        else:
            base = rng.integers(0, 256, (n, d)).astype(np.float32)
            c = np.rint(base.mean(0))
            base = np.ascontiguousarray(base[np.argsort(-((base - c) ** 2).sum(1))])
            q = np.tile(c, (m, 1)).astype(np.float32)
        opts = {"tc_pair": int(rng.integers(0, 2)), "tc_boot": int(rng.choice([0, 1, 2])), "tc_astream": int(rng.choice([1, 1, 2])),
                "tc_real": int(rng.integers(0, 2)), "tc_tf32": int(rng.integers(0, 2))}
        res = []
        for tc in (1, 0):
            e = expann_b200.BruteForceEngine(metric=metric)
            e.set_option("tc", tc)
            if tc:
                for o, v in opts.items():
                    e.set_option(o, v)
            e.store_many_vectors(base); e.build()
            out = e.query_k_batch(q, k); st = e.stats(); e.close()
            res.append((out, st))
        (a, sa), (b, sb) = res
        tag = dict(n=n, m=m, d=d, k=k, kind=kind, metric=metric, **opts, mode=sa["tc_mode"], fb=sa["tc_fallbacks"])
        assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1]) and np.array_equal(a[2], b[2]), tag
        paths.add((int(sa["tc_mode"]), int(sa["last_qtile"])))
        cases += 1
    assert cases >= 20, cases
    assert len({p[0] for p in paths}) >= 2, paths   # more than one operand format was exercised
