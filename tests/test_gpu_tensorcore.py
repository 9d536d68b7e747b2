"""GPU parity: K3 tensor-core candidate filter (tcgen05) + K4 fp32 rerank must give the SAME answer as the exact scan.
Large query batches (m >= tc_min_batch) take this path; smaller ones the SIMT scan."""
import numpy as np
import pytest

from tests.util import assert_topk_parity

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ex():
    import expann_b200
    return expann_b200


def run(ex, base, q, k, metric=0, tc=1, id_offset=0, v1=0, pair=1, boot=None, real=None, tf32=None):
    e = ex.BruteForceEngine(metric=metric, id_offset=id_offset)
    e.set_option("tc", tc)
    if tf32 is not None:
        e.set_option("tc_tf32", tf32)  # 1 (default): tf32 on the fp32 rows (format 4) is tried between the one-term format and the split
    if real is not None:
        e.set_option("tc_real", real)  # real-valued rows: 1 (default) one bf16 term per value (mode 3), 0 two-term split (mode 2)
    e.set_option("tc_pair", pair)  # 1 (default): clusters of two CTAs, tcgen05.mma.cta_group::2; 0: single-CTA form
    if boot is not None:
        e.set_option("tc_boot", boot)  # 2: seeding launch whatever the batch size / mode (default: exact mode, >= 8192 queries)
    e.set_option("tc_v1", v1)   # 1: force the chunked first form of K3 (otherwise used for k > 16 only)
    e.store_many_vectors(base)
    e.build()
    out = e.query_k_batch(q, k)
    st = e.stats()
    e.close()
    return out + (st,)


@pytest.mark.parametrize("metric", [0, 1])
def test_exact_mode_u8_bit_identical(ex, oracle, metric):
    rng = np.random.default_rng(1 + metric)
    base = rng.integers(0, 256, (20011, 128)).astype(np.float32)   # ragged n (not a multiple of the 256-row tile)
    q = rng.integers(0, 256, (777, 128)).astype(np.float32)        # ragged m (not a multiple of the 128-query tile)
    ids, dists, counts, st = run(ex, base, q, 10, metric=metric)
    assert st["tc_mode"] == 1 and st["tc_launches"] > 0 and st["tc_fallbacks"] == 0
    oi, od, oc = oracle.bf_query(base, q, 10, metric=metric, threads=16)
    assert np.array_equal(counts, oc)
    assert_topk_parity(ids, dists, oi, od, base, q, exact=True, tag=f"tc exact metric{metric}")
    # and identical to the SIMT path of the same library
    ids2, dists2, _, st2 = run(ex, base, q, 10, metric=metric, tc=0)
    assert st2["tc_launches"] == 0
    assert np.array_equal(ids, ids2) and np.array_equal(dists, dists2)


@pytest.mark.parametrize("real", [0, 1])
@pytest.mark.parametrize("d", [128, 96, 20])
def test_split_mode_real_valued(ex, oracle, d, real):
    """real-valued rows in both operand formats: the two-term split (mode 2) and one bf16 term per value (mode 3, the
    default: a third of the flops, slack 2^-6 |q| max|x|)"""
    rng = np.random.default_rng(d)
    base = rng.standard_normal((30000, d)).astype(np.float32)
    q = rng.standard_normal((600, d)).astype(np.float32)
    for metric in (0, 1):
        ids, dists, counts, st = run(ex, base, q, 10, metric=metric, real=real)
        assert st["tc_mode"] == (3 if real else 2) and st["tc_launches"] > 0 and st["tc_fallbacks"] == 0
        oi, od, oc = oracle.bf_query(base, q, 10, metric=metric, threads=16)
        swaps = assert_topk_parity(ids, dists, oi, od, base, q, metric="ip" if metric else "l2", tag=f"tc split d{d}")
        assert swaps <= 4


def test_k100_and_duplicates(ex, oracle):
    rng = np.random.default_rng(5)
    pool = rng.integers(0, 256, (300, 64)).astype(np.float32)
    base = pool[rng.integers(0, 300, 9000)]                       # heavy exact duplicates: ties broken by id
    q = rng.integers(0, 256, (520, 64)).astype(np.float32)
    for k in (1, 10, 100):
        ids, dists, counts, st = run(ex, base, q, k)
        assert st["tc_launches"] > 0
        oi, od, _ = oracle.bf_query(base, q, k, threads=16)
        assert_topk_parity(ids, dists, oi, od, base, q, exact=True, tag=f"k{k}")


def test_adversarial_order_falls_back(ex, oracle):
    """rows sorted from far to near: every row beats the running threshold, the candidate buffers overflow, the
    batch is redone by the SIMT scan — the answer must still be exact"""
    rng = np.random.default_rng(6)
    base = rng.integers(0, 256, (40000, 32)).astype(np.float32)
    c = np.rint(base.mean(0))
    order = np.argsort(-((base - c) ** 2).sum(1))
    base = np.ascontiguousarray(base[order])
    q = np.ascontiguousarray(np.tile(c, (512, 1)) + rng.integers(-2, 3, (512, 32))).astype(np.float32)
    ids, dists, counts, st = run(ex, base, q, 10)
    oi, od, _ = oracle.bf_query(base, q, 10, threads=16)
    assert_topk_parity(ids, dists, oi, od, base, q, exact=True, tag="adversarial")
    assert st["tc_launches"] > 0


def test_redo_behind_device_flag(ex, oracle):
    """bf16-exact integer base, queries that are NOT bf16-exact: the device-side check raises the batch's flag and the
    SIMT redo queued behind it (ScanParams::gate) runs -- no host round trip; counted in tc_fallbacks. A clean batch on
    the same handle afterwards must not redo."""
    rng = np.random.default_rng(12)
    base = rng.integers(0, 256, (30000, 128)).astype(np.float32)
    q = rng.integers(0, 256, (300, 128)).astype(np.float32)
    e = ex.BruteForceEngine()
    e.store_many_vectors(base); e.build()
    qf = q + np.float32(0.25)
    ids, dists, counts = e.query_k_batch(qf, 10)
    st = e.stats()
    assert st["tc_mode"] == 1 and st["tc_launches"] == 1 and st["tc_fallbacks"] == 1
    oi, od, _ = oracle.bf_query(base, qf, 10, threads=16)
    assert_topk_parity(ids, dists, oi, od, base, qf, tag="redo")
    ids2, dists2, _ = e.query_k_batch(q, 10)
    st2 = e.stats()
    assert st2["tc_launches"] == 2 and st2["tc_fallbacks"] == 1
    oi2, od2, _ = oracle.bf_query(base, q, 10, threads=16)
    assert np.array_equal(ids2, oi2) and np.array_equal(dists2, od2)
    e.close()


def test_shard_offset_and_small_batches(ex, oracle):
    rng = np.random.default_rng(7)
    base = rng.integers(0, 256, (5000, 128)).astype(np.float32)
    q = rng.integers(0, 256, (600, 128)).astype(np.float32)
    ids, dists, _, st = run(ex, base, q, 10, id_offset=1000)
    oi, od, _ = oracle.bf_query(base, q, 10, threads=16)
    assert np.array_equal(ids, oi + np.uint64(1000)) and np.array_equal(dists, od)
    ids, dists, _, st = run(ex, base, q[:4], 10)        # below tc_min_batch (5): SIMT path
    assert st["tc_launches"] == 0 and np.array_equal(ids, oi[:4])
    ids, dists, _, st = run(ex, base, q[:5], 10)        # smallest batch on the tensor-core path
    assert st["tc_launches"] == 1 and np.array_equal(ids, oi[:5]) and np.array_equal(dists, od[:5])
    ids, dists, _, st = run(ex, base, q[:100], 10)      # small batch on the tensor-core path: many cold row splits
    assert st["tc_launches"] in (1, 2) and np.array_equal(ids, oi[:100]) and np.array_equal(dists, od[:100])


def test_first_form_still_exact(ex, oracle):
    """the chunked (host-refreshed threshold) form of K3 stays available for k > 16; force it for k = 10"""
    rng = np.random.default_rng(8)
    base = rng.integers(0, 256, (20011, 128)).astype(np.float32)
    q = rng.integers(0, 256, (777, 128)).astype(np.float32)
    ids, dists, counts, st = run(ex, base, q, 10, v1=1)
    assert st["tc_launches"] > 1 and st["tc_fallbacks"] == 0
    oi, od, _ = oracle.bf_query(base, q, 10, threads=16)
    assert_topk_parity(ids, dists, oi, od, base, q, exact=True, tag="tc v1")


@pytest.mark.parametrize("k", [1, 10, 16])
def test_second_form_many_splits(ex, k):
    """larger case: several row splits and query blocks, shared bounds between splits; must equal the SIMT scan of the
    same library bit for bit (which the other tests pin to the oracle)"""
    rng = np.random.default_rng(9)
    base = rng.integers(0, 256, (300_001, 128)).astype(np.float32)
    q = rng.integers(0, 256, (1300, 128)).astype(np.float32)
    ids, dists, counts, st = run(ex, base, q, k, boot=0)
    assert st["tc_launches"] == 1 and st["tc_fallbacks"] == 0 and st["last_nsplit"] > 1
    ids3, dists3, counts3, st3 = run(ex, base, q, k)   # with the seeding launch (several row splits + seed merge; the default)
    assert st3["tc_launches"] == 2 and st3["tc_fallbacks"] == 0
    assert np.array_equal(ids, ids3) and np.array_equal(dists, dists3) and np.array_equal(counts, counts3)
    ids2, dists2, counts2, st2 = run(ex, base, q, k, tc=0)
    assert st2["tc_launches"] == 0
    assert np.array_equal(ids, ids2) and np.array_equal(dists, dists2) and np.array_equal(counts, counts2)


def test_second_form_real_valued_splits(ex):
    rng = np.random.default_rng(10)
    base = rng.standard_normal((150_000, 96)).astype(np.float32)
    base /= np.linalg.norm(base, axis=1, keepdims=True)
    q = rng.standard_normal((700, 96)).astype(np.float32)
    for metric in (0, 1):
        ids2, dists2, _, _ = run(ex, base, q, 10, metric=metric, tc=0)
        for real in (0, 1):
            ids, dists, _, st = run(ex, base, q, 10, metric=metric, real=real)
            assert st["tc_mode"] == (3 if real else 2) and st["tc_launches"] in (1, 2) and st["tc_fallbacks"] == 0
            assert np.array_equal(ids, ids2) and np.array_equal(dists, dists2)


@pytest.mark.parametrize("real", [False, True])
def test_single_cta_form_still_exact(ex, real):
    """the single-CTA second form (tc_pair = 0; also what batches of <= 128 queries use) against the pair form"""
    rng = np.random.default_rng(11)
    if real:
        base = rng.standard_normal((90_000, 96)).astype(np.float32)
        q = rng.standard_normal((600, 96)).astype(np.float32)
    else:
        base = rng.integers(0, 256, (90_000, 128)).astype(np.float32)
        q = rng.integers(0, 256, (600, 128)).astype(np.float32)
    a = run(ex, base, q, 10, pair=0)
    b = run(ex, base, q, 10, pair=1)
    assert a[3]["tc_launches"] in (1, 2) and b[3]["tc_launches"] in (1, 2) and a[3]["tc_fallbacks"] == 0 and b[3]["tc_fallbacks"] == 0
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])
    c = run(ex, base, q[:100], 10)   # <= 128 queries: single-CTA form, many cold splits
    assert np.array_equal(c[0], a[0][:100]) and np.array_equal(c[1], a[1][:100])


@pytest.mark.parametrize("n,m", [(100, 129), (256, 256), (257, 257), (32768, 130), (33000, 511), (70001, 1025)])
def test_pair_form_edges(ex, n, m):
    """pair form at its boundaries: a cluster whose second CTA holds 1 query, bases smaller than / equal to / one over a
    256-row tile, the smallest base that gets a seeding launch (128 tiles); every case against the SIMT scan, both
    metrics, integer-valued and real-valued rows"""
    rng = np.random.default_rng(n + m)
    for real in (False, True):
        if real:
            base = rng.standard_normal((n, 64)).astype(np.float32)
            q = rng.standard_normal((m, 64)).astype(np.float32)
        else:
            base = rng.integers(0, 256, (n, 128)).astype(np.float32)
            q = rng.integers(0, 256, (m, 128)).astype(np.float32)
        for metric in (0, 1):
            a = run(ex, base, q, 10, metric=metric, boot=2)
            b = run(ex, base, q, 10, metric=metric, tc=0)
            assert a[3]["tc_launches"] == (2 if n >= 32768 else 1) and a[3]["last_qtile"] == 256 and b[3]["tc_launches"] == 0
            assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1]) and np.array_equal(a[2], b[2])


def test_pair_form_seeded_ties_and_overflow(ex):
    """seeding launch + main launch on a base of heavy duplicates (ties at the k-th value, broken by id) and on rows
    ordered far-to-near (candidate buffers overflow, the batch falls back to the SIMT scan): both must equal the SIMT
    result bit for bit"""
    rng = np.random.default_rng(21)
    pool = rng.integers(0, 256, (2000, 128)).astype(np.float32)
    base = pool[rng.integers(0, 2000, 150_000)]
    q = rng.integers(0, 256, (300, 128)).astype(np.float32)
    a = run(ex, base, q, 10, boot=2)
    b = run(ex, base, q, 10, tc=0)
    assert a[3]["tc_launches"] == 2
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])
    base2 = rng.integers(0, 256, (150_000, 32)).astype(np.float32)
    c = np.rint(base2.mean(0))
    base2 = np.ascontiguousarray(base2[np.argsort(-((base2 - c) ** 2).sum(1))])
    q2 = np.ascontiguousarray(np.tile(c, (300, 1)) + rng.integers(-2, 3, (300, 32))).astype(np.float32)
    a = run(ex, base2, q2, 10, boot=2)
    b = run(ex, base2, q2, 10, tc=0)
    assert a[3]["tc_launches"] >= 1
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])


@pytest.mark.parametrize("n,m,d,kind,astream", [(20011, 300, 128, "u8", 2), (50000, 300, 96, "real", 2), (50000, 300, 960, "unit", 1),
                                               (40000, 100, 256, "real", 1), (30000, 600, 512, "u8", 1), (3000, 20, 960, "unit", 1)])
def test_streamed_query_operand(ex, oracle, n, m, d, kind, astream):
    """pair form with the query operand streamed through the ring (rows too long for a resident query tile: GIST d = 960,
    d = 256 real-valued, d = 512 integers whose dot products pass 2^24 -> split mode), and forced (tc_astream = 2) on
    shapes the resident form also takes. Against the SIMT scan bit for bit, and against the oracle."""
    rng = np.random.default_rng(n + d)
    if kind == "u8":
        base = rng.integers(0, 256, (n, d)).astype(np.float32); q = rng.integers(0, 256, (m, d)).astype(np.float32)
    elif kind == "real":
        base = rng.standard_normal((n, d)).astype(np.float32); q = rng.standard_normal((m, d)).astype(np.float32)
    else:
        base = rng.random((n, d), dtype=np.float32); q = rng.random((m, d), dtype=np.float32)
    for metric in (0, 1):
        e = ex.BruteForceEngine(metric=metric)
        e.set_option("tc_astream", astream)
        e.set_option("tc_real", 0)   # the two-term split: the format whose rows are too long for a resident query tile
        e.store_many_vectors(base); e.build()
        ids, dists, counts = e.query_k_batch(q, 10)
        st = e.stats(); e.close()
        assert st["tc_launches"] >= 1 and st["tc_fallbacks"] == 0 and st["last_qtile"] == 256
        ids2, dists2, counts2, st2 = run(ex, base, q, 10, metric=metric, tc=0)
        assert st2["tc_launches"] == 0
        assert np.array_equal(ids, ids2) and np.array_equal(dists, dists2) and np.array_equal(counts, counts2)
    if d == 960 and n == 50000:
        oi, od, _ = oracle.bf_query(base, q, 10, metric=1, threads=16)
        swaps = assert_topk_parity(ids, dists, oi, od, base, q, metric="ip", tag="streamed d960 ip")
        assert swaps <= 4


def test_streamed_exact_mode_long_rows_and_seeding(ex):
    """streamed query operand in the bf16-EXACT mode (d = 960 of small integers: dot products stay below 2^24, 15 k-blocks
    per tile), with and without the seeding launch; against the SIMT scan"""
    rng = np.random.default_rng(77)
    base = rng.integers(0, 64, (40_000, 960)).astype(np.float32)
    q = rng.integers(0, 64, (300, 960)).astype(np.float32)
    ref = run(ex, base, q, 10, tc=0)
    for boot in (0, 2):
        a = run(ex, base, q, 10, boot=boot)
        assert a[3]["tc_mode"] == 1 and a[3]["tc_launches"] == (2 if boot else 1) and a[3]["tc_fallbacks"] == 0 and a[3]["last_qtile"] == 256
        assert np.array_equal(a[0], ref[0]) and np.array_equal(a[1], ref[1])


def test_small_batch_select_many_candidates(ex):
    """block select of small batches: k = 16 on clustered data, where far more keys than the fast path parks (2048) sit
    under the shared bound, must take the general path and still be exact"""
    rng = np.random.default_rng(78)
    centers = rng.integers(0, 256, (8, 64)).astype(np.float32)
    base = np.clip(centers[rng.integers(0, 8, 400_000)] + rng.integers(-1, 2, (400_000, 64)), 0, 255).astype(np.float32)
    q = np.clip(centers[rng.integers(0, 8, 20)] + rng.integers(-1, 2, (20, 64)), 0, 255).astype(np.float32)
    a = run(ex, base, q, 16)
    b = run(ex, base, q, 16, tc=0)
    assert a[3]["tc_launches"] == 1
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])


@pytest.mark.parametrize("amp", [100, 200, 230, 250])
def test_signed_integer_rows(ex, oracle, amp):
    """signed integer-valued rows (int8-style embeddings): the accumulator of the exact mode runs through
    sum(-2 q_i x_i) + ||x||^2 and reaches 2 d amax^2 + ||x||^2_max with mixed signs -- the exact-mode gate
    (tc_check_exact_kernel and the build-time decision) must count that, or fall back. Whatever path is taken the
    answer must equal the SIMT scan bit for bit and the oracle."""
    rng = np.random.default_rng(300 + amp)
    base = rng.integers(-amp, amp + 1, (40_000, 128)).astype(np.float32)
    q = rng.integers(-amp, amp + 1, (400, 128)).astype(np.float32)
    # a few rows / queries at the extremes of the range: largest norms and dot products of either sign
    base[:64] = np.where(rng.random((64, 128)) < 0.5, -amp, amp)
    q[:16] = base[:16]
    q[16:32] = -base[16:32]
    for metric in (0, 1):
        a = run(ex, base, q, 10, metric=metric)
        b = run(ex, base, q, 10, metric=metric, tc=0)
        assert a[3]["tc_launches"] >= 1 and b[3]["tc_launches"] == 0
        assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1]) and np.array_equal(a[2], b[2])
        oi, od, _ = oracle.bf_query(base, q, 10, metric=metric, threads=16)
        assert_topk_parity(a[0], a[1], oi, od, base, q, metric="ip" if metric else "l2", tag=f"signed amp{amp} metric{metric}")
    if amp <= 100:   # 3 d amax^2 < 2^24: stays in the exact mode without a redo
        assert a[3]["tc_mode"] == 1 and a[3]["tc_fallbacks"] == 0


def test_operand_format_moves_on_after_a_redo(ex, oracle):
    """format 3 (one bf16 term per value) passes the build probe on this base, but the QUERIES carry a large component
    orthogonal to the data: the slack (2^-6 |q| max|x|) then dwarfs the spread of the distances, every row survives, the
    batch is redone by the SIMT scan (exact answer) and the handle moves to tf32 (format 4: slack 5x tighter, still too
    wide here: redone again) and then to the two-term split (format 2), where the third batch runs without a redo"""
    rng = np.random.default_rng(31)
    n, d = 60_000, 64
    base = np.zeros((n, d), np.float32)
    base[:, :32] = rng.standard_normal((n, 32))
    q = np.zeros((300, d), np.float32)
    q[:, :32] = rng.standard_normal((300, 32))
    q[:, 32:] = 200.0
    e = ex.BruteForceEngine()
    e.store_many_vectors(base); e.build()
    assert e.stats()["tc_mode"] == 3
    ref = run(ex, base, q, 10, tc=0)
    seen = []
    for step in range(4):               # (each host call synchronises: the flag word of a batch has arrived before the next one)
        a = e.query_k_batch(q, 10)
        st = e.stats()
        seen.append((int(st["tc_mode"]), int(st["tc_fallbacks"])))
        assert np.array_equal(a[0], ref[0]) and np.array_equal(a[1], ref[1]), seen
    assert seen[0] == (3, 1) and seen[1] == (4, 2) and seen[2] == (2, 2) and seen[3] == (2, 2), seen
    e.close()


def test_build_probe_picks_the_operand_format(ex):
    """build() probes the one-term format with the first rows of the base as queries: data whose distances are too
    concentrated for its slack (GIST-like d = 960; u8 values at d = 512) goes on to tf32 on the fp32 rows (format 4, probed
    the same way; the two-term split only if that fails too), Deep- and Gaussian-like data stays in the one-term format; no
    user batch is redone either way. With tc_tf32 = 0 the second choice is the split."""
    rng = np.random.default_rng(55)
    cases = [("unit960", rng.random((30000, 960), dtype=np.float32), 4), ("u8_512", rng.integers(0, 256, (30000, 512)).astype(np.float32), 4),
             ("gauss128", rng.standard_normal((30000, 128)).astype(np.float32), 3)]
    x = rng.standard_normal((30000, 96)).astype(np.float32)
    cases.append(("deep96", x / np.linalg.norm(x, axis=1, keepdims=True), 3))
    for name, base, mode in cases:
        q = base[rng.integers(0, len(base), 300)] + (0.01 * rng.standard_normal((300, base.shape[1]))).astype(np.float32)
        a = run(ex, base, q, 10)
        assert a[3]["tc_mode"] == mode and a[3]["tc_launches"] >= 1 and a[3]["tc_fallbacks"] == 0, (name, a[3])
        b = run(ex, base, q, 10, tc=0)
        assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1]), name
        if mode == 4:
            c = run(ex, base, q, 10, tf32=0)
            assert c[3]["tc_mode"] == 2 and c[3]["tc_fallbacks"] == 0 and np.array_equal(c[0], b[0]) and np.array_equal(c[1], b[1]), name
            c = run(ex, base, q[:40], 10)        # a small batch in format 4: still the pair form
            assert c[3]["tc_mode"] == 4 and c[3]["last_qtile"] == 256 and np.array_equal(c[0], b[0][:40]) and np.array_equal(c[1], b[1][:40])
