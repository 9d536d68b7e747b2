"""GPU parity: distance primitives, quantizers (a15/a16), K6 compressed scan + K4 rerank."""
import numpy as np
import pytest

from tests.util import assert_topk_parity, gold

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ex():
    import expann_b200
    return expann_b200


def test_primitives_vs_reference_goldens(ex, oracle):
    g = gold("primitives")
    for d in (16, 64, 96, 128, 256, 960):
        a, b = g[f"a{d}"], g[f"b{d}"]
        np.testing.assert_allclose(ex.pair_distances(0, a, b), g[f"avx_dist2_{d}"], rtol=1e-5)   # src/distance.h:136-148
        np.testing.assert_allclose(ex.pair_distances(0, a, b), g[f"eigen_dist2_{d}"], rtol=1e-5)  # src/vec.h:139-142
        scale = np.linalg.norm(a, axis=1) * np.linalg.norm(b, axis=1)
        assert np.all(np.abs(ex.pair_distances(1, a, b) - g[f"avx_dot_{d}"]) <= 1e-5 * scale)    # src/distance.h:181-190
    for d in (64, 128, 256, 960):
        # integer kernels: bit-exact. u8 compressed as the reference computes it (src/antitopo_engine.h:38-61)
        assert np.array_equal(ex.pair_distances(2, g[f"cq{d}"], g[f"cc{d}"]), g[f"compressed_{d}"])
        # int8 / int16: the mathematically correct value (the reference kernels are wrong for negative
        # differences / overflow in 16 bits, src/distance.h:14-53; deliberate deviation, oracle orc_dist2_i8/i16)
        a8, b8 = g[f"i8a{d}"], g[f"i8b{d}"]
        exp8 = ((a8.astype(np.int64) - b8.astype(np.int64)) ** 2).sum(1).astype(np.int32)
        assert np.array_equal(ex.pair_distances(3, a8, b8), exp8)
        a16, b16 = g[f"i16a{d}"], g[f"i16b{d}"]
        exp16 = np.array([oracle.int_dist("orc_dist2_i16", np.ascontiguousarray(a16[i]), np.ascontiguousarray(b16[i]))
                          for i in range(8)], np.int32)
        assert np.array_equal(ex.pair_distances(4, a16, b16), exp16)
    # exact-arithmetic data: fp32 kernels bit-identical to the reference's AVX and Eigen results
    rng = np.random.default_rng(1)
    a = rng.integers(0, 256, (16, 128)).astype(np.float32)
    b = rng.integers(0, 256, (16, 128)).astype(np.float32)
    exp = np.array([oracle.dist2(a[i], b[i]) for i in range(16)], np.float32)
    assert np.array_equal(ex.pair_distances(0, a, b), exp)


@pytest.mark.parametrize("metric", [0, 1])
@pytest.mark.parametrize("d", [128, 96, 960])
def test_u8_scan_vs_oracle(ex, oracle, metric, d):
    """quantizer_simple<u8> codes: integer scores are exact => the k' survivors and the re-ranked top-k are
    bit-identical to the oracle's compressed scan (orc_scan_u8_query)"""
    rng = np.random.default_rng(d + metric)
    n = 6000 if d < 900 else 3000
    base = rng.integers(0, 256, (n, d)).astype(np.float32)
    q = rng.integers(0, 256, (11, d)).astype(np.float32)
    e = ex.QuantScanEngine(metric=metric, kprime=40)
    e.store_many_vectors(base)
    e.build()
    ids, dists, counts = e.query_k_batch(q, 10)
    e.close()
    codes = oracle.quant_simple_u8(base)
    oi, od = oracle.scan_u8_query(base, codes, q, 10, 40, metric=metric, threads=8)
    exact = d * 255 * 255 < 2 ** 24
    assert_topk_parity(ids, dists, oi, od, base, q, metric="ip" if metric else "l2", exact=exact, tag=f"u8 d{d} m{metric}")
    # lossless codes => equal to the plain fp32 brute force as well
    bi, bd, _ = oracle.bf_query(base, q, 10, metric=metric, threads=8)
    assert_topk_parity(ids, dists, bi, bd, base, q, metric="ip" if metric else "l2", exact=exact, tag="vs bf")


def test_ranged_q8(ex, oracle):
    """quantizer_ranged_q8 fit/encode follow src/quantizer.h:196-227 (oracle restatement, parity unpinned: the
    reference never instantiates it); recall of scan+rerank vs exact search on real-valued data"""
    from expann_b200 import datasets
    base, q = datasets.make("gist", 20000, 32, 960)
    e = ex.QuantScanEngine(quantizer=ex.QUANT_RANGED_Q8, metric=1, kprime=100)
    e.store_many_vectors(base)
    e.build()
    s, o = e.quantizer_params()
    _, os_, oo = oracle.quant_ranged_q8(base)
    assert s == os_ and o == oo
    ids, dists, counts = e.query_k_batch(q, 10)
    bi, bd, _ = oracle.bf_query(base, q, 10, metric=1, threads=8)
    rec = np.mean([len(set(a.tolist()) & set(b.tolist())) / 10 for a, b in zip(ids, bi)])
    assert rec >= 0.95, rec
    # returned keys are exact fp32 keys of the returned ids, sorted
    assert np.all(np.diff(dists, axis=1) >= 0)
    for i in (0, 31):
        ref = -(base[ids[i].astype(np.int64)].astype(np.float64) @ q[i].astype(np.float64))
        np.testing.assert_allclose(dists[i], ref, rtol=1e-5)
    e.close()
    # L2 on gaussian data
    base, q = datasets.make("gauss", 20000, 32, 128)
    e = ex.QuantScanEngine(quantizer=ex.QUANT_RANGED_Q8, metric=0, kprime=100)
    e.store_many_vectors(base)
    e.build()
    ids, dists, _ = e.query_k_batch(q, 10)
    bi, bd, _ = oracle.bf_query(base, q, 10, threads=8)
    rec = np.mean([len(set(a.tolist()) & set(b.tolist())) / 10 for a, b in zip(ids, bi)])
    assert rec >= 0.9, rec
    e.close()


def _quant_run(ex, base, q, k, kprime, quantizer, metric, tc, min_batch=None, nsplit=None):
    e = ex.QuantScanEngine(quantizer=quantizer, metric=metric, kprime=kprime)
    e.set_option("quant_tc", tc)
    if min_batch is not None:
        e.set_option("quant_tc_min_batch", min_batch)
    if nsplit is not None:
        e.set_option("nsplit", nsplit)
    e.store_many_vectors(base)
    e.build()
    out = e.query_k_batch(q, k)
    st = e.stats()
    e.close()
    return out + (st,)


@pytest.mark.parametrize("metric", [0, 1])
@pytest.mark.parametrize("n,m,d,kprime", [(20011, 300, 128, 100), (20011, 300, 128, 10), (9000, 130, 960, 100), (5000, 64, 96, 16),
                                          (70000, 700, 512, 128), (300, 40, 128, 100)])
def test_code_scan_on_tensor_cores_equals_simt(ex, oracle, metric, n, m, d, kprime):
    """K6 as a tensor-core candidate filter (tcgen05.mma.kind::i8, csrc/tc_i8.cu) must leave exactly the SIMT scan's k'
    best codes: results of the two forms bit-identical (ids, fp32 keys, counts) for quantizer_simple<u8> codes, whose
    scores are the reference's dist2_compressed integers (src/antitopo_engine.h:38-61), on resident (d <= 512) and
    streamed (d = 960) query tiles, k' <= 16 (exact per-stream lists) and k' > 16 (verified union of per-stream lists);
    and against the oracle's compressed scan"""
    rng = np.random.default_rng(n + d + kprime + metric)
    base = rng.integers(0, 256, (n, d)).astype(np.float32)
    q = rng.integers(0, 256, (m, d)).astype(np.float32)
    a = _quant_run(ex, base, q, 10, kprime, 0, metric, tc=1)
    b = _quant_run(ex, base, q, 10, kprime, 0, metric, tc=0)
    # (a base too small for enough row streams per query at k' > 16 stays on the SIMT scan: the 300-row case)
    assert a[3]["tc_launches"] == (0 if n == 300 else 1) and a[3]["tc_fallbacks"] == 0 and b[3]["tc_launches"] == 0
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1]) and np.array_equal(a[2], b[2])
    if n <= 20011:
        codes = oracle.quant_simple_u8(base)
        oi, od = oracle.scan_u8_query(base, codes, q, 10, kprime, metric=metric, threads=16)
        exact = d * 255 * 255 < 2 ** 24
        assert_topk_parity(a[0], a[1], oi, od, base, q, metric="ip" if metric else "l2", exact=exact, tag=f"code tc d{d} m{metric}")


def test_code_scan_tensor_cores_ranged_q8_ties_and_redo(ex):
    """ranged q8 codes (7-bit, many score ties: the k'-th place is decided by id) through both forms; and rows ordered so
    that one row range owns far more than 16 of a query's best 100 codes: the select kernel's verification must raise the
    flag and the queued SIMT redo must deliver the same answer (counted in tc_fallbacks)"""
    from expann_b200 import datasets
    base, q = datasets.make("gist", 30000, 200, 960)
    for metric in (0, 1):
        a = _quant_run(ex, base, q, 10, 100, 1, metric, tc=1)
        b = _quant_run(ex, base, q, 10, 100, 1, metric, tc=0)
        assert a[3]["tc_launches"] == 1 and a[3]["tc_fallbacks"] == 0
        assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])
    rng = np.random.default_rng(9)
    n, d = 60000, 128
    base = rng.integers(0, 256, (n, d)).astype(np.float32)
    c = rng.integers(100, 156, d).astype(np.float32)
    base[1000:1400] = np.clip(c + rng.integers(-3, 4, (400, d)), 0, 255)     # 400 near-duplicates of one point, contiguous rows
    q = np.clip(c + rng.integers(-2, 3, (64, d)), 0, 255).astype(np.float32)
    a = _quant_run(ex, base, q, 10, 100, 0, 0, tc=1, nsplit=4)
    b = _quant_run(ex, base, q, 10, 100, 0, 0, tc=0)
    assert a[3]["tc_launches"] == 1 and a[3]["tc_fallbacks"] == 1
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])


def test_code_scan_small_batches_stay_on_the_simt_scan(ex):
    rng = np.random.default_rng(10)
    base = rng.integers(0, 256, (8000, 128)).astype(np.float32)
    q = rng.integers(0, 256, (31, 128)).astype(np.float32)
    a = _quant_run(ex, base, q, 10, 40, 0, 0, tc=1)
    assert a[3]["tc_launches"] == 0 and a[3]["scan_launches"] == 1
    b = _quant_run(ex, base, q, 10, 40, 0, 0, tc=1, min_batch=8)
    assert b[3]["tc_launches"] == 1 and b[3]["tc_fallbacks"] == 0
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])


def test_primitives_wide_goldens(ex, oracle):
    """the GPU primitives against 256 reference-minted pairs per dimension (d = 64, 128, 832, 960) incl. rows of extreme
    values: fp32 kernels within 1e-5 relative of the reference's AVX-512 and Eigen results (another summation order),
    dist2_compressed bit-exact, int8 / int16 equal to the mathematically correct value"""
    from tests.util import WIDE_DIMS, wide_primitive_inputs
    g, inp = gold("primitives_wide"), wide_primitive_inputs()
    for d in WIDE_DIMS:
        a, b = inp[f"a{d}"], inp[f"b{d}"]
        got = ex.pair_distances(0, a, b)
        for ref in (g[f"avx_dist2_{d}"], g[f"eigen_dist2_{d}"]):
            assert np.all(np.abs(got.astype(np.float64) - ref) <= 1e-5 * np.abs(ref.astype(np.float64)) + 1e-45)
        assert np.all(got[248:252] == 0) and np.all(got[254:256] == 0)           # identical vectors / zero vectors
        assert np.array_equal(got[232:248], g[f"avx_dist2_{d}"][232:248]) or d * 255 * 255 >= 2 ** 24   # u8-valued rows: exact
        dot = ex.pair_distances(1, a, b)
        scale = np.linalg.norm(a.astype(np.float64), axis=1) * np.linalg.norm(b.astype(np.float64), axis=1)
        assert np.all(np.abs(dot.astype(np.float64) - g[f"avx_dot_{d}"]) <= 1e-5 * scale + 1e-45)
        assert np.array_equal(ex.pair_distances(2, inp[f"cq{d}"], inp[f"cc{d}"]), g[f"compressed_{d}"])
        a8, b8 = inp[f"i8a{d}"], inp[f"i8b{d}"]
        assert np.array_equal(ex.pair_distances(3, a8, b8), ((a8.astype(np.int64) - b8.astype(np.int64)) ** 2).sum(1).astype(np.int32))
        a16, b16 = inp[f"i16a{d}"], inp[f"i16b{d}"]
        assert np.array_equal(ex.pair_distances(4, a16, b16), ((a16.astype(np.int64) - b16.astype(np.int64)) ** 2).sum(1).astype(np.int32))
