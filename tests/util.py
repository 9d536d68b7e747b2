"""shared helpers for the parity tests"""
import os

import numpy as np

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
REL_TOL = 1e-5  # north_star: fp32 distances within 1e-5 relative


def gold(name):
    z = np.load(os.path.join(GOLD, name + ".npz"))
    out = {k: z[k] for k in z.files}
    for k in ("base", "q"):
        if k in out and out[k].dtype == np.uint8:
            out[k] = out[k].astype(np.float32)
    return out


def key64(base, q, metric="l2"):
    """fp64 ranking key of every (query, row) pair"""
    b, qq = base.astype(np.float64), q.astype(np.float64)
    if metric == "ip":
        return -(qq @ b.T)
    return ((qq[:, None, :] - b[None, :, :]) ** 2).sum(-1)


def assert_topk_parity(ids, dists, exp_ids, exp_dists, base, q, metric="l2", exact=False, tag=""):
    """ids must equal the expected ids in order. On real-valued data a mismatch is admissible only if the two
    candidates are a near-tie: their fp64 keys agree within REL_TOL relative. Returns the
    number of admitted near-tie swaps. exact=True: bit-identical ids and distances required."""
    ids, exp_ids = np.asarray(ids), np.asarray(exp_ids)
    assert ids.shape == exp_ids.shape, (tag, ids.shape, exp_ids.shape)
    if exact:
        assert np.array_equal(ids, exp_ids), f"{tag}: ids differ"
        if dists is not None and exp_dists is not None:
            assert np.array_equal(np.asarray(dists), np.asarray(exp_dists)), f"{tag}: distances differ"
        return 0
    swaps = 0
    bad = np.argwhere(ids != exp_ids)
    if len(bad):
        for qi, j in bad:
            a, b = int(ids[qi, j]), int(exp_ids[qi, j])
            assert a != 2**64 - 1 and b != 2**64 - 1, f"{tag}: padding mismatch q={qi} j={j}"
            ka = key64(base[a:a + 1], q[qi:qi + 1], metric)[0, 0]
            kb = key64(base[b:b + 1], q[qi:qi + 1], metric)[0, 0]
            scale = max(abs(ka), abs(kb), 1e-30)
            assert abs(ka - kb) <= REL_TOL * scale, f"{tag}: q={qi} j={j} ids {a} vs {b}: keys {ka} vs {kb} not a near-tie"
            swaps += 1
        # as sets, each query's results must agree up to near-ties at the k boundary; checked above element-wise
    if dists is not None and exp_dists is not None:
        d, e = np.asarray(dists, np.float64), np.asarray(exp_dists, np.float64)
        fin = np.isfinite(e)
        assert np.array_equal(np.isfinite(d), fin), f"{tag}: padding differs"
        rel = np.abs(d[fin] - e[fin]) / np.maximum(np.abs(e[fin]), 1e-30)
        assert rel.max(initial=0.0) <= REL_TOL, f"{tag}: distance rel err {rel.max()}"
    return swaps


WIDE_DIMS = (64, 128, 832, 960)


def wide_primitive_inputs():
    """seeded inputs of tests/golden/primitives_wide.npz (256 pairs per dimension; rows of extreme values: 1e15-scale,
    denormal-scale squares, u8-valued, identical vectors, zeros; largest u8 / int8 differences; int16 differences past the
    reference's 16-bit mullo range). tests/golden/make_golden_wide.py runs the reference on exactly these arrays."""
    NP = 256
    rng = np.random.default_rng(2026)
    out = {}
    for d in WIDE_DIMS:
        a = rng.standard_normal((NP, d)).astype(np.float32) * 3
        b = rng.standard_normal((NP, d)).astype(np.float32) * 3
        a[200:216] *= np.float32(1e15); b[200:216] *= np.float32(1e15)
        a[216:232] *= np.float32(1e-20); b[216:232] *= np.float32(1e-20)
        a[232:248] = rng.integers(0, 256, (16, d)); b[232:248] = rng.integers(0, 256, (16, d))
        b[248:252] = a[248:252]
        a[252:256] = 0; b[254:256] = 0
        out[f"a{d}"], out[f"b{d}"] = a, b
        qi = rng.integers(0, 256, (NP, d)).astype(np.float32) + rng.random((NP, d)).astype(np.float32) * 0.9
        ci = rng.integers(0, 256, (NP, d)).astype(np.uint8)
        qi[240:248] = 255.0; ci[240:244] = 0; ci[244:248] = 255
        qi[248:256] = 0.0; ci[248:252] = 255
        out[f"cq{d}"], out[f"cc{d}"] = qi, ci
        i8a = rng.integers(-128, 128, (NP, d)).astype(np.int8); i8b = rng.integers(-128, 128, (NP, d)).astype(np.int8)
        i8a[240:248] = 127; i8b[240:248] = -128
        out[f"i8a{d}"], out[f"i8b{d}"] = i8a, i8b
        i16a = rng.integers(-300, 300, (NP, d)).astype(np.int16); i16b = rng.integers(-300, 300, (NP, d)).astype(np.int16)
        i16a[240:248] = 181; i16b[240:248] = -181
        out[f"i16a{d}"], out[f"i16b{d}"] = i16a, i16b
    return out
