"""CPU: pin the oracle (oracle/oracle.c) against golden vectors minted from the UNMODIFIED reference
(tests/golden/make_golden.py) and, when oracle/_ref is present on this host, against the reference live."""
import numpy as np
import pytest

from tests.util import assert_topk_parity, gold

DIMS = (16, 64, 96, 128, 256, 960)


def test_primitives_f32_bit_exact_vs_reference_avx(oracle):
    g = gold("primitives")
    for d in DIMS:
        a, b = g[f"a{d}"], g[f"b{d}"]
        got = np.array([oracle.dist2(a[i], b[i]) for i in range(8)], np.float32)
        assert np.array_equal(got, g[f"avx_dist2_{d}"]), d   # src/distance.h:136-148
        if d % 128 == 0:
            assert np.array_equal(got, g[f"avx128_dist2_{d}"]), d  # src/distance.h:86-111
        got = np.array([oracle.dot(a[i], b[i]) for i in range(8)], np.float32)
        assert np.array_equal(got, g[f"avx_dot_{d}"]), d     # src/distance.h:181-190
        # Eigen reductions (src/vec.h:139-142,109-111) use another summation tree: 1e-5 relative
        np.testing.assert_allclose(np.array([oracle.dist2(a[i], b[i]) for i in range(8)]), g[f"eigen_dist2_{d}"], rtol=1e-5)


def test_primitives_integer_bit_exact(oracle):
    g = gold("primitives")
    for d in (64, 128, 256, 960):
        for i in range(8):
            assert oracle.dist2_u8(g[f"cq{d}"][i], g[f"cc{d}"][i]) == g[f"compressed_{d}"][i]  # antitopo_engine.h:38-61
            assert oracle.int_dist("orc_dist2_i8_asref", np.ascontiguousarray(g[f"i8a{d}"][i]),
                                   np.ascontiguousarray(g[f"i8b{d}"][i])) == g[f"i8_asref_{d}"][i]  # distance.h:29-53
            assert oracle.int_dist("orc_dist2_i16_asref", np.ascontiguousarray(g[f"i16a{d}"][i]),
                                   np.ascontiguousarray(g[f"i16b{d}"][i])) == g[f"i16_asref_{d}"][i]  # distance.h:14-27
            a, b = g[f"i8a{d}"][i].astype(np.int64), g[f"i8b{d}"][i].astype(np.int64)
            assert oracle.int_dist("orc_dist2_i8", np.ascontiguousarray(g[f"i8a{d}"][i]),
                                   np.ascontiguousarray(g[f"i8b{d}"][i])) == int(((a - b) ** 2).sum())


@pytest.mark.parametrize("name,exact", [("bf_ties_64x16", True), ("bf_tiny_5x16", False), ("bf_u8_3000x128", True),
                                        ("bf_u8dup_1500x128", True), ("bf_gauss_3000x128_dim128", False),
                                        ("bf_gauss_3000x128_dyn", False), ("bf_deep_2000x96", False),
                                        ("bf_gist_600x960", False), ("bf_k100_3000x128", False)])
def test_bruteforce_vs_golden(oracle, name, exact):
    g = gold(name)
    ids, dists, counts = oracle.bf_query(g["base"], g["q"], int(g["k"]))
    assert np.array_equal(counts, g["counts"])
    assert_topk_parity(ids, dists, g["ids"], g["dists"], g["base"], g["q"], exact=exact, tag=name)


def test_topk_t_semantics(oracle):
    """src/topk_t.h:24-35: bounded max-heap on (dist,id), dedupe of ids currently held, strict '>' admission"""
    rng = np.random.default_rng(3)
    t = oracle.topk(5)
    for _ in range(200):
        d, v = float(rng.integers(0, 20)), int(rng.integers(0, 30))
        before = dict(zip(*[x.tolist() for x in t.to_combined()][::-1]))  # id -> dist
        good = t.consider(d, v)
        exp_good = (v not in before) and (len(before) < 5 or max(before.values()) > d)
        assert good == exp_good
        assert t.size() <= 5
    dd, ii = t.to_combined()
    assert np.all(np.diff(dd) >= 0)
    assert len(set(ii.tolist())) == len(ii)
    assert t.worst_val() == dd[-1]


def test_graph_vs_golden(oracle):
    """oracle graph search on the reference-built index reproduces the reference's own results id for id
    (u8-valued data: exact arithmetic; libstdc++ heap behaviour restated)"""
    import os
    from tests.util import GOLD
    g = gold("graph_u8_2000")
    G = oracle.graph_load(os.path.join(GOLD, "graph_u8_2000.index"))
    assert [G.start, G.max_layer, G.n, G.M, G.M0] == g["info"].tolist()
    for ef in (10, 20, 40, 80):
        ids, _, _, dc = G.query(g["q"], 10, ef, compressed=False)
        assert np.array_equal(ids, g[f"ids_ef{ef}"]), ef
        assert dc == int(g[f"distcomps_ef{ef}"]), ef
        idc, _, _, _ = G.query(g["q"], 10, ef, compressed=True)
        assert np.array_equal(idc, g[f"ids_comp_ef{ef}"]), ef


def test_live_reference_matches_oracle(oracle, reference_dyn):
    """only where oracle/_ref exists (build container): random shapes, reference vs restatement"""
    rng = np.random.default_rng(5)
    for d in (32, 96, 128):
        base = rng.integers(0, 256, (700, d)).astype(np.float32)
        q = rng.integers(0, 256, (5, d)).astype(np.float32)
        i1, d1, _ = oracle.bf_query(base, q, 7)
        i2, d2, _ = reference_dyn.bf(base).query(q, 7)
        assert np.array_equal(i1, i2) and np.array_equal(d1, d2)


def test_primitives_wide_goldens(oracle):
    """256 pairs per dimension incl. d = 832 and rows of extreme values (tests/golden/make_golden_wide.py): the oracle's
    restatements of the reference kernels are bit-exact against what the UNMODIFIED reference computed"""
    from tests.util import WIDE_DIMS, wide_primitive_inputs
    g, inp = gold("primitives_wide"), wide_primitive_inputs()
    for d in WIDE_DIMS:
        a, b = inp[f"a{d}"], inp[f"b{d}"]
        got = np.array([oracle.dist2(a[i], b[i]) for i in range(256)], np.float32)
        assert np.array_equal(got, g[f"avx_dist2_{d}"]), d                   # src/distance.h:112-148
        if d % 128 == 0:
            assert np.array_equal(got, g[f"avx128_dist2_{d}"]), d            # src/distance.h:86-111
        got = np.array([oracle.dot(a[i], b[i]) for i in range(256)], np.float32)
        assert np.array_equal(got, g[f"avx_dot_{d}"]), d                     # src/distance.h:181-190
        for i in range(256):
            assert oracle.dist2_u8(inp[f"cq{d}"][i], inp[f"cc{d}"][i]) == g[f"compressed_{d}"][i]          # antitopo_engine.h:38-61
            assert oracle.int_dist("orc_dist2_i8_asref", np.ascontiguousarray(inp[f"i8a{d}"][i]),
                                   np.ascontiguousarray(inp[f"i8b{d}"][i])) == g[f"i8_asref_{d}"][i]       # distance.h:29-53 as written
            assert oracle.int_dist("orc_dist2_i16_asref", np.ascontiguousarray(inp[f"i16a{d}"][i]),
                                   np.ascontiguousarray(inp[f"i16b{d}"][i])) == g[f"i16_asref_{d}"][i]     # distance.h:14-27 as written
        # the reference's 16-bit mullo wraps for |diff| > 181 (rows 240-247: diff 362): as-written differs from the true value
        t = ((inp[f"i16a{d}"][240].astype(np.int64) - inp[f"i16b{d}"][240].astype(np.int64)) ** 2).sum()
        assert int(g[f"i16_asref_{d}"][240]) != int(t)
