"""GPU parity at BASELINE.json's full sizes, through a property that does not need the CPU oracle to finish: the
tensor-core paths (K3 exact / one-term formats, K6-TC code filter) and the SIMT scans of the same library are independent
implementations that the small-size tests pin to the oracle; here they must agree bit for bit (ids, fp32 keys, counts) on
the full C2 base (1M x 128, u8-valued), on a Deep-shape base (2M x 96, real-valued: the C4 shard of an 8-GPU run holds
1.25M rows) and on the full C5 base (1M x 960, ranged-q8 codes), on a query sample the SIMT scan finishes in
milliseconds. Also: sortedness of every result list, ids unique and in range, keys equal to the fp64 distance of the
returned id within 1e-5 relative."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _properties(ids, keys, base, q, metric):
    n = len(base)
    assert ids.max() < n and np.all(np.diff(keys, axis=1) >= 0)
    for i in range(0, len(q), max(1, len(q) // 16)):
        assert len(set(ids[i].tolist())) == ids.shape[1]
        x = base[ids[i].astype(np.int64)].astype(np.float64)
        ref = -(x @ q[i].astype(np.float64)) if metric else ((x - q[i].astype(np.float64)) ** 2).sum(1)
        np.testing.assert_allclose(keys[i], ref, rtol=1e-5, atol=1e-30)


def _bf(ex, base, q, k, metric, tc):
    e = ex.BruteForceEngine(metric=metric)
    e.set_option("tc", tc)
    e.store_many_vectors(base); e.build()
    out = e.query_k_batch(q, k); st = e.stats(); e.close()
    return out, st


def test_c2_full_base_tensor_core_equals_simt():
    import expann_b200 as ex
    from expann_b200 import datasets
    base, q = datasets.make("sift", 1_000_000, 10_000, 128, variant="iid")
    (a, sa) = _bf(ex, base, q, 10, 0, 1)            # all 10 000 queries through K3 (pair form, seeded)
    assert sa["tc_mode"] == 1 and sa["tc_launches"] == 2 and sa["tc_fallbacks"] == 0
    sub = np.arange(0, 10_000, 13)[:768]
    (b, sb) = _bf(ex, base, q[sub], 10, 0, 0)        # SIMT scan of a query sample
    assert sb["tc_launches"] == 0
    assert np.array_equal(a[0][sub], b[0]) and np.array_equal(a[1][sub], b[1]) and np.array_equal(a[2][sub], b[2])
    _properties(a[0], a[1], base, q, 0)


def test_deep_shape_one_term_format_equals_simt():
    import expann_b200 as ex
    from expann_b200 import datasets
    base, q = datasets.make("deep", 2_000_000, 10_000, 96)
    (a, sa) = _bf(ex, base, q, 10, 0, 1)
    assert sa["tc_mode"] == 3 and sa["tc_fallbacks"] == 0
    sub = np.arange(0, 10_000, 17)[:512]
    (b, sb) = _bf(ex, base, q[sub], 10, 0, 0)
    assert np.array_equal(a[0][sub], b[0]) and np.array_equal(a[1][sub], b[1])
    _properties(a[0], a[1], base, q, 0)


def test_c5_full_base_code_filter_equals_simt_code_scan():
    import expann_b200 as ex
    from expann_b200 import datasets
    base, q = datasets.make("gist", 1_000_000, 1_000, 960)
    res = []
    for tc in (1, 0):
        e = ex.QuantScanEngine(quantizer=1, kprime=100, metric=1)
        e.set_option("quant_tc", tc)
        e.store_many_vectors(base); e.build()
        qq = q if tc else q[:256]
        res.append((e.query_k_batch(qq, 10), e.stats())); e.close()
    (a, sa), (b, sb) = res
    assert sa["tc_launches"] == 1 and sa["tc_fallbacks"] == 0 and sb["tc_launches"] == 0
    assert np.array_equal(a[0][:256], b[0]) and np.array_equal(a[1][:256], b[1])
    _properties(a[0], a[1], base, q, 1)
