"""CPU, world_size 2 over gloo: the host-side logic of the sharded search (row ranges, id offsets, rank-major
all-gather layout, key encoding). The per-rank top-k and the merge are played by the oracle / numpy here (the
CUDA kernels cannot run without a GPU); the GPU test test_shard_merge_is_partition_independent covers K7."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, n, m, k, outdir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from expann_b200 import sharding
    from oracle.pyoracle import Oracle
    rng = np.random.default_rng(0)
    base = rng.integers(0, 256, (n, 32)).astype(np.float32)
    q = rng.integers(0, 256, (m, 32)).astype(np.float32)
    lo, hi = sharding.shard_range(n, world, rank)
    ids, d, _ = Oracle().bf_query(base[lo:hi], q, k)
    gids = np.where(ids == np.uint64(2**64 - 1), ids, ids + np.uint64(lo))      # id_offset
    local = torch.from_numpy(sharding.encode_keys(d, gids).view(np.int64))
    allk = sharding.gather_keys(local, world)                                   # [world][m][k]
    assert tuple(allk.shape) == (world, m, k)
    merged = np.sort(allk.numpy().view(np.uint64).transpose(1, 0, 2).reshape(m, world * k), axis=1)[:, :k]
    md, mi = sharding.decode_keys(merged)
    oi, od, _ = Oracle().bf_query(base, q, k)
    assert np.array_equal(mi, oi) and np.array_equal(md, od)
    open(os.path.join(outdir, f"ok{rank}"), "w").write("ok")
    dist.destroy_process_group()


def test_two_rank_sharded_search(tmp_path):
    port = _free_port()
    mp.spawn(_worker, args=(2, port, 1001, 7, 10, str(tmp_path)), nprocs=2, join=True)
    assert (tmp_path / "ok0").exists() and (tmp_path / "ok1").exists()


def test_shard_range_covers_everything():
    from expann_b200 import sharding
    for n in (1, 7, 1000, 1_000_000):
        for w in (1, 2, 3, 4, 8):
            r = [sharding.shard_range(n, w, g) for g in range(w)]
            assert r[0][0] == 0 and r[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(r, r[1:]))


def test_key_roundtrip():
    from expann_b200 import sharding
    d = np.array([[0.0, 1.5, -2.0, np.inf, 3e38, -0.0]], np.float32)
    ids = np.array([[5, 4, 3, 2**64 - 1, 1, 0]], np.uint64)
    k = sharding.encode_keys(d, ids)
    dd, ii = sharding.decode_keys(k)
    assert np.array_equal(ii, ids)
    assert np.array_equal(dd[ids != np.uint64(2**64 - 1)], d[ids != np.uint64(2**64 - 1)])
    # order of keys == order of (float, id)
    order = np.argsort(k[0][[0, 1, 2, 4, 5]])
    assert list(np.array([0.0, 1.5, -2.0, 3e38, -0.0], np.float32)[order]) == [-2.0, -0.0, 0.0, 1.5, np.float32(3e38)]
