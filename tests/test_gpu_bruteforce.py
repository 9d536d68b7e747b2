"""GPU parity: K1/K2 (scan) + K7 (merge) through the C ABI vs the oracle / the reference-minted goldens."""
import numpy as np
import pytest

from tests.util import assert_topk_parity, gold

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ex():
    import expann_b200
    return expann_b200


def run_bf(ex, base, q, k, metric=0, qtile=0, id_offset=0, chunked=False):
    e = ex.BruteForceEngine(metric=metric, qtile=qtile, id_offset=id_offset)
    e.set_option("tc", 0)  # this file pins the SIMT scan (K1/K2); the tensor-core path has tests/test_gpu_tensorcore.py
    if chunked:  # mix of single-vector store_vector calls and bulk adds
        e.store_vector(base[0])
        e.store_many_vectors(base[1:7])
        if len(base) > 7:
            e.store_many_vectors(base[7:])
    else:
        e.store_many_vectors(base)
    e.build()
    out = e.query_k_batch(q, k)
    st = e.stats()
    e.close()
    return out + (st,)


GOLDENS = [("bf_ties_64x16", True), ("bf_tiny_5x16", False), ("bf_u8_3000x128", True), ("bf_u8dup_1500x128", True),
           ("bf_gauss_3000x128_dim128", False), ("bf_gauss_3000x128_dyn", False), ("bf_deep_2000x96", False),
           ("bf_gist_600x960", False), ("bf_k100_3000x128", False)]


@pytest.mark.parametrize("name,exact", GOLDENS)
def test_golden(ex, name, exact):
    g = gold(name)
    ids, dists, counts, st = run_bf(ex, g["base"], g["q"], int(g["k"]))
    assert st["scan_launches"] == 1
    assert np.array_equal(counts, g["counts"])
    assert_topk_parity(ids, dists, g["ids"], g["dists"], g["base"], g["q"], exact=exact, tag=name)


@pytest.mark.parametrize("qtile", [1, 4, 8])
@pytest.mark.parametrize("metric", [0, 1])
def test_vs_oracle_qtiles(ex, oracle, qtile, metric):
    rng = np.random.default_rng(100 + qtile)
    base = rng.integers(0, 256, (20011, 128)).astype(np.float32)  # ragged n: not a multiple of the 512-row stage
    q = rng.integers(0, 256, (13, 128)).astype(np.float32)        # ragged m: not a multiple of the query tile
    ids, dists, counts, st = run_bf(ex, base, q, 10, metric=metric, qtile=qtile, chunked=True)
    oi, od, oc = oracle.bf_query(base, q, 10, metric=metric, threads=8)
    assert st["last_qtile"] == qtile and st["last_nsplit"] >= 2
    assert np.array_equal(counts, oc)
    # u8-valued, d=128: every partial sum is an integer < 2^24 => bit-exact ids and keys
    assert_topk_parity(ids, dists, oi, od, base, q, exact=True, tag=f"qt{qtile} metric{metric}")


@pytest.mark.parametrize("d", [4, 20, 96, 100, 256, 960])
def test_dims(ex, oracle, d):
    rng = np.random.default_rng(d)
    base = rng.standard_normal((5000, d)).astype(np.float32)
    q = rng.standard_normal((9, d)).astype(np.float32)
    for metric in (0, 1):
        ids, dists, counts, _ = run_bf(ex, base, q, 10, metric=metric)
        oi, od, oc = oracle.bf_query(base, q, 10, metric=metric, threads=8)
        swaps = assert_topk_parity(ids, dists, oi, od, base, q, metric="ip" if metric else "l2", tag=f"d{d}")
        assert swaps <= 2


def test_k_edges(ex, oracle):
    rng = np.random.default_rng(9)
    base = rng.integers(0, 256, (700, 64)).astype(np.float32)
    q = rng.integers(0, 256, (5, 64)).astype(np.float32)
    for k in (1, 2, 31, 32, 33, 100, 128):
        ids, dists, counts, _ = run_bf(ex, base, q, k)
        oi, od, oc = oracle.bf_query(base, q, k)
        assert_topk_parity(ids, dists, oi, od, base, q, exact=True, tag=f"k{k}")
    # k > n: fewer than k results, padded (src/brute_force_engine.h:33-45 returns only n ids)
    ids, dists, counts, _ = run_bf(ex, base[:20], q, 32)
    assert np.all(counts == 20) and np.all(ids[:, 20:] == np.uint64(2**64 - 1)) and np.all(np.isinf(dists[:, 20:]))
    oi, od, _ = oracle.bf_query(base[:20], q, 32)
    assert np.array_equal(ids, oi) and np.array_equal(dists, od)


def test_single_query_api(ex, oracle):
    rng = np.random.default_rng(10)
    base = rng.integers(0, 256, (3000, 128)).astype(np.float32)
    e = ex.BruteForceEngine()
    for v in base[:50]:
        e.store_vector(v)           # src/ann_engine.h:22-24, one vector at a time
    e.store_many_vectors(base[50:])
    e.build()
    q = rng.integers(0, 256, (4, 128)).astype(np.float32)
    oi, _, _ = oracle.bf_query(base, q, 10)
    for i in range(4):
        assert e.query_k(q[i], 10) == [int(x) for x in oi[i]]
    assert e.name().startswith("Brute-Force") and isinstance(e.param_list(), dict)
    e.close()


def test_errors(ex):
    e = ex.BruteForceEngine()
    with pytest.raises(ex.ExpannError):
        e.build()                               # empty engine (reference asserts n > 0)
    e.store_many_vectors(np.zeros((4, 8), np.float32))
    with pytest.raises(ex.ExpannError):
        e.query_k_batch(np.zeros((1, 8), np.float32), 3)  # query before build
    with pytest.raises(ex.ExpannError):
        e.store_many_vectors(np.zeros((4, 9), np.float32))  # dimension mismatch
    e.build()
    with pytest.raises(ex.ExpannError):
        e.query_k_batch(np.zeros((1, 8), np.float32), 129)  # k beyond the supported list size
    ids, _, counts = e.query_k_batch(np.zeros((0, 8), np.float32), 3)
    assert ids.shape == (0, 3)
    e.close()


def test_shard_merge_is_partition_independent(ex, oracle):
    """top-k of a union == merge of per-shard top-k; ids carry the shard offset"""
    import torch
    rng = np.random.default_rng(12)
    base = rng.integers(0, 256, (9000, 128)).astype(np.float32)
    q = rng.integers(0, 256, (33, 128)).astype(np.float32)
    k, G = 10, 3
    oi, od, _ = oracle.bf_query(base, q, k, threads=8)
    bounds = [0, 2000, 2512, 9000]
    dq = torch.from_numpy(q).cuda()
    gathered = torch.empty((G, 33, k), dtype=torch.int64, device="cuda")
    engines = []
    for g in range(G):
        e = ex.BruteForceEngine(id_offset=bounds[g])
        e.store_many_vectors(base[bounds[g]:bounds[g + 1]])
        e.build()
        e.query_device(dq.data_ptr(), 33, k, gathered[g].data_ptr(), torch.cuda.current_stream().cuda_stream)
        engines.append(e)
    out = torch.empty((33, k), dtype=torch.int64, device="cuda")
    ids = torch.empty((33, k), dtype=torch.int64, device="cuda")
    dist = torch.empty((33, k), dtype=torch.float32, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    engines[0].merge_device(gathered.data_ptr(), 33, G, k, k, 33 * k, k, out.data_ptr(), st)
    engines[0].decode_device(out.data_ptr(), 33, k, ids.data_ptr(), dist.data_ptr(), 0, st)
    torch.cuda.synchronize()
    assert np.array_equal(ids.cpu().numpy().astype(np.uint64), oi)
    assert np.array_equal(dist.cpu().numpy(), od)
    for e in engines:
        e.close()


def test_large_property_1m(ex, oracle):
    """BASELINE C2 size (1M x 128, k=10): size-independent properties + spot parity on a query subset"""
    from expann_b200 import datasets
    base, q = datasets.make("sift", 1_000_000, 256, 128, variant="iid")
    e = ex.BruteForceEngine()
    e.store_many_vectors(base)
    e.build()
    ids, dists, counts = e.query_k_batch(q, 10)
    assert np.all(counts == 10)
    assert np.all(np.diff(dists, axis=1) >= 0)                       # sorted nearest first
    tie = np.diff(dists, axis=1) == 0
    assert np.all(np.diff(ids.astype(np.int64), axis=1)[tie] > 0)    # ties broken by id
    for i in range(256):                                             # no duplicate ids
        assert len(set(ids[i].tolist())) == 10
    # returned distances are the true distances of the returned ids (exact: u8-valued data)
    for i in range(0, 256, 17):
        ref = ((base[ids[i].astype(np.int64)] - q[i]) ** 2).sum(1, dtype=np.float64)
        assert np.array_equal(ref.astype(np.float32), dists[i])
    # idempotence + tile-independence: same answer one query at a time (QT=1 path) as in the batch
    for i in (0, 100, 255):
        assert e.query_k(q[i], 10) == [int(x) for x in ids[i]]
    # spot parity against the oracle on 8 queries
    oi, od, _ = oracle.bf_query(base, q[:8], 10, threads=8)
    assert np.array_equal(ids[:8], oi) and np.array_equal(dists[:8], od)
    # a planted exact duplicate of a query must come back first with distance 0
    e.close()


def test_host_call_with_page_locked_buffers(ex):
    """expann_gpu_query with page-locked caller buffers (copied by the copy engines directly) and with pageable ones
    (through the handle's staging buffers) must return the same thing"""
    import torch
    rng = np.random.default_rng(31)
    base = rng.integers(0, 256, (30_000, 128)).astype(np.float32)
    q = rng.integers(0, 256, (333, 128)).astype(np.float32)
    e = ex.BruteForceEngine()
    e.store_many_vectors(base); e.build()
    ids, dists, counts = e.query_k_batch(q, 10)
    hq = torch.from_numpy(q).pin_memory().numpy()
    out = (torch.zeros((333, 10), dtype=torch.int64).pin_memory().numpy().view(np.uint64),
           torch.zeros((333, 10), dtype=torch.float32).pin_memory().numpy(),
           torch.zeros((333,), dtype=torch.int64).pin_memory().numpy().view(np.uint64))
    r = e.query_k_batch(hq, 10, out=out)
    assert r[0] is out[0]
    assert np.array_equal(out[0], ids) and np.array_equal(out[1], dists) and np.array_equal(out[2], counts)
    # mixed: pinned queries, pageable outputs
    ids2, dists2, counts2 = e.query_k_batch(hq, 10)
    assert np.array_equal(ids2, ids) and np.array_equal(dists2, dists)
    with pytest.raises(ValueError):
        e.query_k_batch(q, 10, out=(out[0][:, :5], out[1], out[2]))
    e.close()
