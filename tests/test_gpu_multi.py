"""GPU parity: ONE handle over several GPUs (expann_gpu_create_multi, include/expann_gpu.h; csrc/multi.cu) must return
exactly what the single-device handle returns, for every decomposition G = query groups x row shards.

The fan-out / gather / merge logic is the same whether the shards sit on different GPUs or on one: on a one-GPU box the
device list repeats ordinal 0 ("virtual shards"), on a multi-GPU box the real ordinals are used as well (peer stores over
NVLink into the first GPU's gather buffer)."""
import os

import numpy as np
import pytest

from tests.util import GOLD, gold

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ex():
    import expann_b200
    return expann_b200


def device_lists():
    import torch
    cnt = torch.cuda.device_count()
    lists = [("virtual2", [0, 0]), ("virtual3", [0, 0, 0]), ("virtual8", [0] * 8)]
    for g in (2, 3, 8):
        if cnt >= g:
            lists.append((f"real{g}", list(range(g))))
    return lists


def single(ex, base, q, k, metric=0, **opts):
    e = ex.BruteForceEngine(metric=metric)
    for o, v in opts.items():
        e.set_option(o, v)
    e.store_many_vectors(base); e.build()
    out = e.query_k_batch(q, k)
    e.close()
    return out


@pytest.mark.parametrize("kind", ["u8", "real", "pad100"])
def test_brute_force_identical_to_one_gpu(ex, kind):
    rng = np.random.default_rng(41)
    if kind == "u8":
        base = rng.integers(0, 256, (50_003, 128)).astype(np.float32); q = rng.integers(0, 256, (1300, 128)).astype(np.float32)
    elif kind == "real":
        base = rng.standard_normal((50_003, 96)).astype(np.float32); q = rng.standard_normal((1300, 96)).astype(np.float32)
    else:
        base = rng.integers(0, 256, (20_001, 100)).astype(np.float32); q = rng.integers(0, 256, (700, 100)).astype(np.float32)
    want = {m: single(ex, base, q[:m], 10) for m in (1, 4, 300, len(q))}
    for name, devs in device_lists():
        for rg in sorted({0, len(devs), 2 if len(devs) % 2 == 0 else 0}):   # automatic / one shard per GPU / 2 shards x G/2 query groups
            e = ex.BruteForceEngine(devices=devs, row_groups=rg)
            e.store_many_vectors(base[:30_000]); e.store_many_vectors(base[30_000:])   # two add() calls
            e.build()
            st = e.stats()
            assert st["n"] == len(base) and st["query_groups"] * st["row_groups"] <= len(devs)
            assert st["row_groups"] == (1 if rg == 0 else rg)        # automatic: a base below 2^20 rows is not cut, the GPUs split the queries
            for m, w in want.items():
                got = e.query_k_batch(q[:m], 10)
                assert np.array_equal(got[0], w[0]) and np.array_equal(got[1], w[1]) and np.array_equal(got[2], w[2]), (name, rg, m)
            st = e.stats()
            assert st["tc_launches"] >= 1 and st["scan_launches"] >= 1
            if name.startswith("real"):
                assert st["peer_bytes"] > 0
            e.close()


def test_ip_metric_k100_and_tiny_bases(ex):
    rng = np.random.default_rng(42)
    base = rng.integers(0, 256, (9000, 64)).astype(np.float32)
    q = rng.integers(0, 256, (520, 64)).astype(np.float32)
    for k, metric in ((100, 0), (10, 1), (1, 1)):
        w = single(ex, base, q, k, metric=metric)
        for name, devs in device_lists():
            e = ex.BruteForceEngine(metric=metric, devices=devs, row_groups=len(devs))
            e.store_many_vectors(base); e.build()
            got = e.query_k_batch(q, k)
            e.close()
            assert np.array_equal(got[0], w[0]) and np.array_equal(got[1], w[1]) and np.array_equal(got[2], w[2]), (name, k, metric)
    # fewer rows than GPUs / than k: no empty shard is created, short lists are padded exactly like the one-GPU result
    tiny = base[:5]
    w = single(ex, tiny, q[:7], 10)
    e = ex.BruteForceEngine(devices=[0] * 8, row_groups=8)
    e.store_many_vectors(tiny); e.build()
    got = e.query_k_batch(q[:7], 10)
    st = e.stats()
    e.close()
    assert st["row_groups"] == 5
    assert np.array_equal(got[0], w[0]) and np.array_equal(got[1], w[1]) and np.array_equal(got[2], w[2])


def test_device_pointer_entry_and_back_to_back_calls(ex):
    """expann_gpu_query_device on a multi-device handle: device buffers and the stream belong to the first GPU; calls
    queued back to back without a host synchronisation (the gather buffer is double-buffered by call parity) must each
    give the one-GPU keys"""
    import torch
    rng = np.random.default_rng(43)
    base = rng.integers(0, 256, (40_000, 128)).astype(np.float32)
    qs = [rng.integers(0, 256, (900, 128)).astype(np.float32) for _ in range(5)]
    one = ex.BruteForceEngine()
    one.store_many_vectors(base); one.build()
    dev = torch.device("cuda", 0)
    st = torch.cuda.current_stream(dev).cuda_stream
    dq = [torch.from_numpy(x).to(dev) for x in qs]
    want = [torch.empty((900, 10), dtype=torch.int64, device=dev) for _ in qs]
    for a, w in zip(dq, want):
        one.query_device(a.data_ptr(), 900, 10, w.data_ptr(), st)
    torch.cuda.synchronize()
    one.close()
    for name, devs in device_lists():
        e = ex.BruteForceEngine(devices=devs, row_groups=len(devs))
        e.store_many_vectors(base); e.build()
        got = [torch.empty((900, 10), dtype=torch.int64, device=dev) for _ in qs]
        for a, g in zip(dq, got):
            e.query_device(a.data_ptr(), 900, 10, g.data_ptr(), st)   # no synchronisation in between
        torch.cuda.synchronize()
        for i, (g, w) in enumerate(zip(got, want)):
            assert torch.equal(g, w), (name, i)
        e.close()


def test_calls_on_several_streams_overlap_safely(ex):
    """calls issued on different streams without a host synchronisation: the brute-force handle keeps two sets of work
    buffers (one per stream, csrc/api.cu scratch_for_stream), a third stream takes over the set used longest ago behind
    an event; a multi-device handle alternates two streams per GPU. Every call must give the keys of the same call made
    alone, whatever else is in flight; also for the engines that have one set (quantised scan)."""
    import torch
    rng = np.random.default_rng(47)
    base = rng.integers(0, 256, (60_000, 64)).astype(np.float32)
    sizes = (700, 300, 1000, 40, 3, 513)
    qs = [rng.integers(0, 256, (m, 64)).astype(np.float32) for m in sizes]
    dev = torch.device("cuda", 0)
    dq = [torch.from_numpy(x).to(dev) for x in qs]
    streams = [torch.cuda.Stream(device=dev) for _ in range(3)]
    makers = [("bf", lambda: ex.BruteForceEngine()), ("bf_real", lambda: ex.BruteForceEngine()),
              ("quant", lambda: ex.QuantScanEngine(quantizer=0, kprime=30)),
              ("multi", lambda: ex.BruteForceEngine(devices=[0, 0, 0], row_groups=0)),
              ("multi_rows", lambda: ex.BruteForceEngine(devices=[0, 0], row_groups=2))]
    for name, mk in makers:
        b = base + 0.25 if name == "bf_real" else base
        e = mk()
        e.store_many_vectors(b); e.build()
        main = torch.cuda.current_stream(dev).cuda_stream
        want = [torch.empty((m, 10), dtype=torch.int64, device=dev) for m in sizes]
        for a, w, m in zip(dq, want, sizes):
            e.query_device(a.data_ptr(), m, 10, w.data_ptr(), main)
            torch.cuda.synchronize()
        for nstreams in (2, 3):
            got = [[torch.zeros((m, 10), dtype=torch.int64, device=dev) for m in sizes] for _ in range(3)]
            for rep in range(3):
                for i, (a, m) in enumerate(zip(dq, sizes)):
                    stm = streams[(i + rep) % nstreams]
                    e.query_device(a.data_ptr(), m, 10, got[rep][i].data_ptr(), stm.cuda_stream)
            torch.cuda.synchronize()
            for rep in range(3):
                for i in range(len(sizes)):
                    assert torch.equal(got[rep][i], want[i]), (name, nstreams, rep, i)
        e.close()


def test_immutable_after_build_and_errors(ex):
    from expann_b200._lib import ExpannError, ERR_STATE, ERR_UNSUPPORTED, ERR_ARG
    rng = np.random.default_rng(44)
    base = rng.integers(0, 256, (3000, 32)).astype(np.float32)
    e = ex.BruteForceEngine(devices=[0, 0])
    with pytest.raises(ExpannError) as ei:
        e.build()                                  # empty engine
    assert ei.value.code == ERR_STATE
    e.store_many_vectors(base); e.build()
    with pytest.raises(ExpannError) as ei:
        e.store_many_vectors(base)
    assert ei.value.code == ERR_STATE
    with pytest.raises(ExpannError) as ei:
        e.set_option("no_such_knob", 1)
    assert ei.value.code == ERR_ARG
    e.close()
    with pytest.raises(ExpannError) as ei:
        ex.BruteForceEngine(devices=[0, 99])
    assert ei.value.code == ERR_ARG
    g = ex.GraphEngine(devices=[0, 0])
    g.store_many_vectors(base)
    g.build_graph(M=8, M0=16, K=32)
    with pytest.raises(ExpannError) as ei:
        g.save_index("/tmp/never_written.index")
    assert ei.value.code == ERR_UNSUPPORTED
    g.close()


def test_quant_scan_does_not_depend_on_the_partitioning(ex):
    """ranged q8: ONE value range is fitted over the whole base, so codes, survivors and fp32 results equal the one-GPU
    engine's"""
    rng = np.random.default_rng(45)
    base = rng.random((20_000, 192), dtype=np.float32)
    base[7777] *= 1.7          # the global maximum sits in one shard only
    q = rng.random((64, 192), dtype=np.float32)
    for quantizer in (0, 1):
        b = (base * 200).astype(np.float32) if quantizer == 0 else base
        qq = (q * 200).astype(np.float32) if quantizer == 0 else q
        one = ex.QuantScanEngine(quantizer=quantizer, kprime=50)
        one.store_many_vectors(b); one.build()
        w = one.query_k_batch(qq, 10)
        wp = one.quantizer_params()
        one.close()
        for name, devs in device_lists():
            e = ex.QuantScanEngine(quantizer=quantizer, kprime=50, devices=devs, row_groups=len(devs))
            e.store_many_vectors(b); e.build()
            got = e.query_k_batch(qq, 10)
            gp = e.quantizer_params()
            e.close()
            if quantizer == 1:
                assert gp == wp
            assert np.array_equal(got[0], w[0]) and np.array_equal(got[1], w[1]), (name, quantizer)


def test_graph_replicas_and_sharded_subgraphs(ex):
    """load_index on a multi-device handle replicates the graph (queries split, results identical to one GPU);
    build_graph builds one sub-graph per row shard (every query searches every shard: recall >= the single graph's)"""
    g = gold("graph_u8_2000")
    q = g["q"].astype(np.float32)
    path = os.path.join(GOLD, "graph_u8_2000.index")
    one = ex.GraphEngine()
    one.load_index(path); one.build(); one.set_ef_search(40)
    w = one.query_k_batch(q, 10)
    one.close()
    for name, devs in device_lists():
        e = ex.GraphEngine(devices=devs)
        e.load_index(path); e.build(); e.set_ef_search(40)
        got = e.query_k_batch(q, 10)
        st = e.stats()
        e.close()
        assert st["row_groups"] == 1 and st["query_groups"] == len(devs) and st["n"] == 2000
        assert np.array_equal(got[0], w[0]) and np.array_equal(got[1], w[1]), name
    from expann_b200 import datasets
    base, qq = datasets.make("sift", 24_000, 200, 128, seed=5, variant="lowrank")
    truth = single(ex, base, qq, 10)[0]

    def rec(ids):
        return float(np.mean([len(set(a.tolist()) & set(b.tolist())) / 10.0 for a, b in zip(ids, truth)]))
    one = ex.GraphEngine()
    one.store_many_vectors(base); one.build_graph(M=12, M0=24, K=48); one.build(); one.set_ef_search(30)
    r1 = rec(one.query_k_batch(qq, 10)[0])
    one.close()
    for name, devs in device_lists()[:2] + device_lists()[3:4]:
        e = ex.GraphEngine(devices=devs, row_groups=len(devs))     # one sub-graph per row shard
        e.store_many_vectors(base); e.build_graph(M=12, M0=24, K=48); e.build(); e.set_ef_search(30)
        ids, dists, counts = e.query_k_batch(qq, 10)
        e.close()
        assert np.all(counts == 10) and np.all(np.diff(dists, axis=1) >= 0)
        assert rec(ids) >= r1 - 1e-9, (name, rec(ids), r1)
