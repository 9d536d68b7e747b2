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
