"""GPU parity: K5 graph beam search through the C ABI on reference-built graphs.
Bar (north_star): recall@10 >= the reference's at every beam width on the SAME graph. On u8-valued data the
arithmetic is exact, so result sets are also compared id for id and the distance-evaluation counter
(RECORD_STATS num_distcomps, src/antitopo_engine.h:125-129) against the reference's."""
import os

import numpy as np
import pytest

from tests.util import GOLD, gold

pytestmark = pytest.mark.gpu


def recall(ids, truth):
    return float(np.mean([len(set(a.tolist()) & set(b.tolist())) / len(b) for a, b in zip(ids, truth)]))


@pytest.fixture(scope="module")
def ex():
    import expann_b200
    return expann_b200


@pytest.mark.parametrize("compressed", [False, True])
def test_golden_index(ex, compressed):
    g = gold("graph_u8_2000")
    e = ex.GraphEngine(use_compression=compressed)
    e.load_index(os.path.join(GOLD, "graph_u8_2000.index"))
    e.build()
    q = g["q"].astype(np.float32)
    for ef in (10, 20, 40, 80):
        e.set_ef_search(ef)
        dc0 = e.stats()["distcomps"]
        ids, dists, counts = e.query_k_batch(q, 10)
        dc = e.stats()["distcomps"] - dc0
        ref_ids = g[f"ids_comp_ef{ef}" if compressed else f"ids_ef{ef}"]
        r_gpu, r_ref = recall(ids, g["truth"]), recall(ref_ids, g["truth"])
        assert np.all(counts == 10)
        assert r_gpu >= r_ref - 1e-9, (ef, r_gpu, r_ref)
        for i in range(len(q)):
            assert len(set(ids[i].tolist())) == 10                       # no duplicate ids (src/basic_bench.h:98-104)
        assert np.all(np.diff(dists, axis=1) >= 0)
        if not compressed:
            # exact arithmetic + same algorithm: same result SETS up to distance ties at the k-th place
            same = np.mean([set(a.tolist()) == set(b.tolist()) for a, b in zip(ids, ref_ids)])
            assert same >= 0.9, (ef, same)
            ref_dc = int(g[f"distcomps_ef{ef}"])
            assert abs(dc - ref_dc) <= 0.02 * ref_dc, (ef, dc, ref_dc)
    e.close()


def test_live_reference_graph(ex, oracle, reference_128):
    """build a graph with the unmodified reference on this host, hand it over as CSR arrays, compare recall"""
    from expann_b200 import datasets
    base, q = datasets.make("sift", 6000, 64, 128, seed=77, variant="lowrank")
    G = reference_128.graph(M=12, M0=24, ef_construction=48)
    G.store(base)
    G.build()
    path = "/tmp/expann_test_graph.index"
    G.save(path)
    og = oracle.graph_load(path)
    truth, _, _ = oracle.bf_query(base, q, 10, threads=8)
    e = ex.GraphEngine()
    e.load_index(path)
    e.build()
    for ef in (10, 30, 60):
        G.set_ef(ef)
        rid, _ = G.query(q, 10)
        e.set_ef_search(ef)
        ids, dists, counts = e.query_k_batch(q, 10)
        assert recall(ids, truth) >= recall(rid, truth) - 1e-9, ef
        # the returned keys are the true fp32 distances of the returned ids
        for i in (0, 17, 63):
            ref = ((base[ids[i].astype(np.int64)] - q[i]) ** 2).sum(1, dtype=np.float64).astype(np.float32)
            assert np.array_equal(ref, dists[i])
    e.close()


def test_graph_errors(ex):
    e = ex.GraphEngine()
    e.store_many_vectors(np.zeros((8, 32), np.float32))
    with pytest.raises(ex.ExpannError):
        e.build()  # no graph loaded
    with pytest.raises(ex.ExpannError):
        e.load_index("/nonexistent/file.index")
    e.close()


def test_gpu_built_graph_roundtrip(ex, oracle, reference_128):
    """GPU graph construction + index writer (row 2): the graph built on the GPU is written in
    the reference's serialize format, the UNMODIFIED reference deserializes it and searches it on the CPU; on that
    same graph the GPU search must reach the reference's recall at every beam width, and the graph must be good."""
    from expann_b200 import datasets
    base, q = datasets.make("sift", 8000, 64, 128, seed=31, variant="lowrank")
    truth, _, _ = oracle.bf_query(base, q, 10, threads=8)
    e = ex.GraphEngine()
    e.store_many_vectors(base)
    e.build_graph(M=12, M0=24, K=48)
    e.build()
    path = "/tmp/expann_gpu_built.index"
    e.save_index(path)
    G = reference_128.graph(M=12, M0=24, ef_construction=48)
    G.load(path, 128)
    info = G.info()
    assert info["n"] == 8000 and info["M0"] == 24
    og = oracle.graph_load(path)
    degs = np.array([len(og.neighbours(v, 0)) for v in range(0, 8000, 97)])
    assert degs.max() <= 24 and degs.mean() >= 6
    last = 0.0
    for ef in (10, 20, 40, 80):
        G.set_ef(ef)
        rid, _ = G.query(q, 10)
        e.set_ef_search(ef)
        ids, dists, counts = e.query_k_batch(q, 10)
        r_gpu, r_ref = recall(ids, truth), recall(rid, truth)
        assert r_gpu >= r_ref - 1e-9, (ef, r_gpu, r_ref)
        same = np.mean([set(a.tolist()) == set(b.tolist()) for a, b in zip(ids, rid)])
        assert same >= 0.9, (ef, same)
        last = r_gpu
    assert last >= 0.95, last
    e.close()


def test_latency_form_real_valued(ex):
    """one-CTA-per-query form of K5 (calls with few queries) against the one-warp-per-query form on the same graph,
    real-valued rows (d = 96, unit norm): same visit order and same per-vertex summation order; the forms differ only
    where a neighbour's distance EQUALS the beam's bound (row-at-once admission), which real-valued data almost never
    produces => identical ids / distances / distance-evaluation counts for (nearly) every query. Covers more queries
    than CTA slots, the 8-bit visited-epoch wrap and one query per call (src/basic_bench.h:82-126)."""
    from expann_b200 import datasets
    base, q = datasets.make("deep", 20000, 700, 96, seed=5)
    e = ex.GraphEngine()
    e.store_many_vectors(base)
    e.build_graph(M=12, M0=24, K=48)
    e.build()
    for ef in (1, 10, 40, 100, 200):
        k = 10 if ef >= 10 else 1
        e.set_ef_search(ef)
        e.set_option("graph_cta_batch", 0)
        dc0 = e.stats()["distcomps"]
        ids_w, dists_w, counts_w = e.query_k_batch(q, k)
        dc1 = e.stats()["distcomps"]
        e.set_option("graph_cta_batch", 1 << 20)
        e.set_option("graph_cta_width", 1)
        ids_c, dists_c, counts_c = e.query_k_batch(q, k)
        dc2 = e.stats()["distcomps"]
        same = np.all(ids_w == ids_c, axis=1) & np.all(dists_w == dists_c, axis=1)
        assert same.mean() >= 0.99, (ef, same.mean())
        assert np.array_equal(counts_w, counts_c), ef
        assert abs((dc2 - dc1) - (dc1 - dc0)) <= 0.01 * (dc1 - dc0), ef
        for i in (0, 7):
            i1, d1, c1 = e.query_k_batch(q[i:i + 1], k)   # small host call: the search decodes its own results
            assert np.array_equal(i1[0], ids_c[i]) and np.array_equal(d1[0], dists_c[i]) and c1[0] == counts_c[i]
    # two vertices expanded per hop (option graph_cta_width = 2): a different visit order, so the bar is the
    # engine's own — recall@10 >= the one-at-a-time search's at every beam width, no duplicates, ascending distances
    truth = np.argsort(((base[None, :, :] - q[:64, None, :]) ** 2).sum(-1), axis=1, kind="stable")[:, :10]
    for ef in (10, 20, 40, 100):
        e.set_ef_search(ef)
        e.set_option("graph_cta_width", 1)
        ids_1, _, _ = e.query_k_batch(q[:64], 10)
        e.set_option("graph_cta_width", 2)
        ids_2, dists_2, counts_2 = e.query_k_batch(q[:64], 10)
        assert recall(ids_2, truth) >= recall(ids_1, truth) - 0.005, (ef, recall(ids_2, truth), recall(ids_1, truth))
        assert np.all(counts_2 == 10) and all(len(set(r.tolist())) == 10 for r in ids_2)
        assert np.all(np.diff(dists_2, axis=1) >= 0)
        ref = ((base[ids_2[5].astype(np.int64)] - q[5]) ** 2).sum(1, dtype=np.float64).astype(np.float32)
        assert np.allclose(ref, dists_2[5], rtol=1e-5)
    # fewer vertices reachable than k: unused slots are id = 2^64-1, dist = +inf, count < k (include/expann_gpu.h)
    e.set_ef_search(3)
    e.set_option("graph_cta_width", 1)
    i1, d1, c1 = e.query_k_batch(q[:1], 10)
    i2, d2, c2 = e.query_k_batch(q[:400], 10)
    assert c1[0] == 3 and np.all(i1[0, 3:] == np.uint64(2**64 - 1)) and np.all(np.isinf(d1[0, 3:]))
    assert np.array_equal(i1[0], i2[0]) and np.array_equal(d1[0], d2[0]) and c2[0] == 3
    e.close()


@pytest.mark.parametrize("width", [1, 2])
@pytest.mark.parametrize("compressed", [False, True])
def test_latency_form_golden_index(ex, compressed, width):
    """the latency form on the reference-built golden graph (u8-valued rows: distance ties are common, so the forms
    may order tied vertices differently): recall@10 >= the reference's at every beam width, same bar as
    test_golden_index; result sets equal to the warp form's for most queries."""
    g = gold("graph_u8_2000")
    q = g["q"].astype(np.float32)
    e = ex.GraphEngine(use_compression=compressed)
    e.load_index(os.path.join(GOLD, "graph_u8_2000.index"))
    e.build()
    for ef in (10, 20, 40, 80):
        e.set_ef_search(ef)
        e.set_option("graph_cta_batch", 0)
        ids_w, _, _ = e.query_k_batch(q, 10)
        e.set_option("graph_cta_batch", 1 << 20)
        e.set_option("graph_cta_width", width)
        ids, dists, counts = e.query_k_batch(q, 10)
        ref_ids = g[f"ids_comp_ef{ef}" if compressed else f"ids_ef{ef}"]
        assert np.all(counts == 10)
        assert recall(ids, g["truth"]) >= recall(ref_ids, g["truth"]) - 1e-9, ef
        assert all(len(set(r.tolist())) == 10 for r in ids)
        assert np.all(np.diff(dists, axis=1) >= 0)
        if width == 1:  # same visit order as the warp form
            same = np.mean([set(a.tolist()) == set(b.tolist()) for a, b in zip(ids, ids_w)])
            assert same >= 0.9, (ef, same)
    e.close()


def test_corrupt_index_files_are_rejected(ex, tmp_path):
    """a truncated or corrupt index must come back as an error code through the C ABI (no exception, no abort, no
    out-of-range id aliasing a valid vertex): graph_load_index validates every count against the file size"""
    from expann_b200._lib import ExpannError, ERR_IO
    good = open(os.path.join(GOLD, "graph_u8_2000.index"), "rb").read()
    cases = {
        "empty": b"",
        "header_only": good[:40],
        "truncated_vectors": good[:len(good) // 3],
        "truncated_edges": good[:len(good) - 1000],
    }
    # header layout (src/antitopo_engine.h:931-991): 4 x u64, bool has_ef [u64], 2 x u64, 2 x f32, u64, 2 x bool, u64 max_layer, u64 n
    has_ef = good[32]
    off_n = 33 + (8 if has_ef else 0) + 16 + 8 + 8 + 2 + 8
    huge_n = bytearray(good); huge_n[off_n:off_n + 8] = (2**62).to_bytes(8, "little"); cases["huge_n"] = bytes(huge_n)
    n = int.from_bytes(good[off_n:off_n + 8], "little")
    d = int.from_bytes(good[off_n + 8:off_n + 16], "little")
    assert n == 2000 and d == 128
    huge_d = bytearray(good); huge_d[off_n + 8:off_n + 16] = (2**40).to_bytes(8, "little"); cases["huge_d"] = bytes(huge_d)
    # first adjacency record: after the vectors, u64 n, u64 layers, u64 deg, then {f32 len, u64 id}
    off_v = off_n + 8 + n * (8 + 4 * d) + 8
    deg_off = off_v + 8
    bad_id = bytearray(good); bad_id[deg_off + 8 + 4:deg_off + 8 + 12] = (2**32 + 5).to_bytes(8, "little"); cases["id_aliases_mod_2_32"] = bytes(bad_id)
    bad_deg = bytearray(good); bad_deg[deg_off:deg_off + 8] = (2**50).to_bytes(8, "little"); cases["huge_degree"] = bytes(bad_deg)
    bad_nl = bytearray(good); bad_nl[off_v:off_v + 8] = (2**33).to_bytes(8, "little"); cases["huge_layer_count"] = bytes(bad_nl)
    for name, blob in cases.items():
        p = tmp_path / (name + ".index")
        p.write_bytes(blob)
        e = ex.GraphEngine()
        with pytest.raises(ExpannError) as ei:
            e.load_index(str(p))
        assert ei.value.code == ERR_IO, (name, ei.value)
        e.close()
    e = ex.GraphEngine()          # and the intact file still loads
    e.load_index(os.path.join(GOLD, "graph_u8_2000.index"))
    e.build()
    e.close()


def test_load_index_after_vectors_of_another_dimension(ex):
    """vectors stored before load_index are replaced by the file's (read_index, src/antitopo_engine.h:312-313); the HBM
    buffer sized for the old dimension must not be reused: d = 32 rows first, then the d = 128 golden index"""
    g = gold("graph_u8_2000")
    q = g["q"].astype(np.float32)
    ref = ex.GraphEngine()
    ref.load_index(os.path.join(GOLD, "graph_u8_2000.index")); ref.build(); ref.set_ef_search(40)
    want = ref.query_k_batch(q, 10)
    ref.close()
    for d0 in (32, 100):    # smaller d_pad (buffer overflow before the fix) / same d_pad, smaller d (stale padding columns)
        e = ex.GraphEngine()
        e.store_many_vectors(np.full((5000, d0), 7.0, np.float32))
        e.load_index(os.path.join(GOLD, "graph_u8_2000.index"))
        e.build(); e.set_ef_search(40)
        got = e.query_k_batch(q, 10)
        e.close()
        assert np.array_equal(got[0], want[0]) and np.array_equal(got[1], want[1])


@pytest.mark.parametrize("n,d,M,M0,nq", [(6000, 128, 12, 32, 48), (3000, 96, 8, 24, 40), (2500, 100, 8, 48, 32),
                                         (1500, 256, 8, 16, 16), (800, 20, 4, 8, 16)])
def test_inline_blocks_identical(ex, n, d, M, M0, nq):
    """the latency form on neighbourhood blocks (graph_inline, csrc/graph_inline.cuh) returns what the other forms
    return, bit for bit — ids, distances, counts and the distance-evaluation counter — whether the visited set is the
    shared-memory bitmap (1) or the bytes in HBM (2); rows of more than 32 neighbours, d that is not a multiple of 32
    and d above 128 take the general path of the scorers"""
    from expann_b200 import datasets
    base, q = datasets.make("sift", n, nq, d, seed=5, variant="lowrank")
    eng = {}
    for inl in (0, 1, 2):
        e = ex.GraphEngine()
        e.set_option("graph_inline", inl)
        e.store_many_vectors(base)
        e.build_graph(M=M, M0=M0, K=2 * M0)
        e.build()
        eng[inl] = e
    for ef in (1, 10, 33, 100, 250):
        k = min(10, ef)
        res = {}
        for inl, e in eng.items():
            e.set_ef_search(ef)
            dc0 = e.stats()["distcomps"]
            one = [e.query_k_batch(q[i:i + 1], k) for i in range(nq)]          # one CTA, one query per call
            dc = e.stats()["distcomps"] - dc0
            few = e.query_k_batch(q, k)                                          # one CTA per query, nq CTAs
            res[inl] = (np.concatenate([o[0] for o in one]), np.concatenate([o[1] for o in one]),
                        np.concatenate([o[2] for o in one]), few, dc)
        for inl in (1, 2):
            for a, b in zip(res[0][:3], res[inl][:3]):
                assert np.array_equal(a, b), (ef, inl)
            for a, b in zip(res[0][3], res[inl][3]):
                assert np.array_equal(a, b), (ef, inl)
            assert res[0][4] == res[inl][4], (ef, inl, res[0][4], res[inl][4])
        # and the one-warp-per-query form agrees with all of them: over adjacency rows (engine 0), over the blocks (engine 1,
        # graph_search_blocks_kernel), and over adjacency rows of an engine that has blocks (graph_inline_batch = 0)
        for inl, batch in ((0, 1), (1, 1), (1, 0)):
            eng[inl].set_option("graph_cta_batch", 0)
            eng[inl].set_option("graph_inline_batch", batch)
            dc0 = eng[inl].stats()["distcomps"]
            warp = eng[inl].query_k_batch(q, k)
            dc = eng[inl].stats()["distcomps"] - dc0
            eng[inl].set_option("graph_cta_batch", -1)
            eng[inl].set_option("graph_inline_batch", 1)
            for a, b in zip(res[1][3], warp):
                assert np.array_equal(a, b), (ef, inl, batch)
            assert dc == res[0][4], (ef, inl, batch, dc, res[0][4])
    for e in eng.values():
        e.close()


@pytest.mark.parametrize("n,d,maxdeg,updeg,layers", [(900, 64, 200, 40, 3), (300, 128, 33, 8, 2), (64, 96, 5, 3, 1), (1, 128, 0, 0, 1)])
def test_inline_blocks_handmade_graphs(ex, n, d, maxdeg, updeg, layers):
    """hand-made CSR graphs (random edges; nothing about them is good, the forms only have to agree): bottom rows of up
    to 200 neighbours and upper rows of more than 32 (the scorers' general path, several rounds per row), ragged
    degrees, vertices without neighbours, a graph without upper layers, a one-vertex graph, k above what is reachable"""
    rng = np.random.default_rng(n + d)
    base = rng.integers(0, 256, (n, d)).astype(np.float32)
    q = rng.integers(0, 256, (24, d)).astype(np.float32)
    nl = np.ones(n, np.uint32)
    if layers > 1:
        top = rng.permutation(n)[:max(2, n // 6)]
        nl[top] = rng.integers(2, layers + 1, top.size).astype(np.uint32)
    start = int(np.argmax(nl))
    offsets, edges = [0], []
    for v in range(n):
        for l in range(int(nl[v])):
            pool = np.arange(n) if l == 0 else np.nonzero(nl > l)[0]
            pool = pool[pool != v]
            cap = maxdeg if l == 0 else updeg
            deg = 0 if pool.size == 0 or cap == 0 else int(rng.integers(0 if v % 7 == 3 else 1, min(cap, pool.size) + 1))
            edges += rng.choice(pool, deg, replace=False).tolist()
            offsets.append(len(edges))
    eng = {}
    for inl in (0, 1, 2):
        e = ex.GraphEngine()
        e.set_option("graph_inline", inl)
        e.store_many_vectors(base)
        e.set_graph(nl, np.array(offsets, np.uint64), np.array(edges if edges else [0], np.uint32)[:len(edges)], start, int(nl.max()))
        e.build()
        eng[inl] = e
    for ef, k in ((1, 1), (7, 5), (40, 10), (256, 100)):
        res = {}
        for inl, e in eng.items():
            e.set_ef_search(ef)
            one = [e.query_k_batch(q[i:i + 1], k) for i in range(len(q))]
            res[inl] = [np.concatenate([o[j] for o in one]) for j in range(3)] + list(e.query_k_batch(q, k))
        e = eng[1]
        e.set_option("graph_cta_batch", 0)
        res["warp"] = res[1][:3] + list(e.query_k_batch(q, k))
        e.set_option("graph_cta_batch", -1)
        for name in (1, 2, "warp"):
            for a, b in zip(res[0], res[name]):
                assert np.array_equal(a, b), (ef, k, name)
    for e in eng.values():
        e.close()


def test_inline_blocks_several_queries_per_cta(ex):
    """more queries than resident CTA slots: every CTA of the block form runs several queries one after the other (its
    rendezvous counters, visited bitmap and scorer state carry over); results equal the one-warp-per-query form's"""
    from expann_b200 import datasets
    base, q = datasets.make("sift", 5000, 64, 128, seed=9, variant="lowrank")
    rng = np.random.default_rng(3)
    qq = np.ascontiguousarray(q[rng.integers(0, 64, 3000)] + rng.integers(0, 3, (3000, 128)).astype(np.float32))
    for inl in (1, 2):
        e = ex.GraphEngine()
        e.set_option("graph_inline", inl)
        e.set_option("graph_occ", 2)          # 2 x 148 CTAs x 4 slots = 1184 < 3000 queries
        e.store_many_vectors(base)
        e.build_graph(M=8, M0=24, K=48)
        e.build()
        for ef in (10, 40):
            e.set_ef_search(ef)
            e.set_option("graph_cta_batch", 10 ** 6)
            a = e.query_k_batch(qq, 10)
            e.set_option("graph_cta_batch", 0)
            b = e.query_k_batch(qq, 10)
            for x, y in zip(a, b):
                assert np.array_equal(x, y), (inl, ef)
        e.close()


@pytest.mark.parametrize("n,d,M,M0", [(6000, 128, 12, 32), (3000, 96, 8, 24), (2500, 100, 8, 48), (1500, 256, 8, 16)])
def test_inline_code_blocks_identical(ex, n, d, M, M0):
    """compressed graphs (use_compression: the bottom layer scored on 8-bit codes, fp32 re-rank of the ef survivors): the
    block form over CODE blocks returns what the row forms return, bit for bit, with either visited-set form"""
    from expann_b200 import datasets
    base, q = datasets.make("sift", n, 40, d, seed=6, variant="lowrank")
    eng = {}
    for inl in (0, 1, 2):
        e = ex.GraphEngine(use_compression=True)
        e.set_option("graph_inline", inl)
        e.store_many_vectors(base)
        e.build_graph(M=M, M0=M0, K=2 * M0)
        e.build()
        eng[inl] = e
    for ef in (1, 10, 40, 130):
        k = min(10, ef)
        res = {}
        for inl, e in eng.items():
            e.set_ef_search(ef)
            dc0 = e.stats()["distcomps"]
            one = [e.query_k_batch(q[i:i + 1], k) for i in range(len(q))]
            dc = e.stats()["distcomps"] - dc0
            res[inl] = [np.concatenate([o[j] for o in one]) for j in range(3)] + list(e.query_k_batch(q, k)) + [dc]
        e = eng[0]
        e.set_option("graph_cta_batch", 0)
        warp = list(e.query_k_batch(q, k))
        e.set_option("graph_cta_batch", -1)
        for inl in (1, 2):
            for a, b in zip(res[0][:6], res[inl][:6]):
                assert np.array_equal(a, b), (ef, inl)
            assert res[0][6] == res[inl][6], (ef, inl)
            for a, b in zip(res[inl][3:6], warp):
                assert np.array_equal(a, b), (ef, inl)
    for e in eng.values():
        e.close()
