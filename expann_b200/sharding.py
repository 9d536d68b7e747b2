"""Base-set sharding across GPUs: contiguous row ranges, one process per GPU, per-rank top-k,
`all_gather` of the [m][k] 64-bit keys over torch.distributed (NCCL on GPUs, gloo in the CPU tests) and a K7 merge.
The reference has no multi-device analogue; because keys carry (key, id) the merged result is bit-identical to the
single-GPU result for any number of shards."""
from __future__ import annotations

import numpy as np


def shard_range(n: int, world: int, rank: int) -> tuple[int, int]:
    """rows [lo, hi) of rank `rank`: GPU g holds rows [g*ceil(n/G), min(n, (g+1)*ceil(n/G)))"""
    per = (n + world - 1) // world
    lo = min(n, rank * per)
    return lo, min(n, lo + per)


def gather_keys(local_keys, world: int):
    """all-gather the per-rank [m][k] int64 key tensors -> [world][m][k] (rank-major, what merge_device expects
    with stride_q = k, stride_s = m*k)"""
    import torch
    import torch.distributed as dist
    if world == 1:
        return local_keys.unsqueeze(0)
    m, k = local_keys.shape
    out = torch.empty((world * m, k), dtype=local_keys.dtype, device=local_keys.device)
    dist.all_gather_into_tensor(out, local_keys.contiguous())
    return out.view(world, m, k)


def orderable_u32(x: np.ndarray) -> np.ndarray:
    """host mirror of f32_orderable (csrc/common.cuh): monotone float32 -> uint32"""
    u = np.ascontiguousarray(x, np.float32).view(np.uint32)
    return np.where(u >> 31 != 0, ~u, u | np.uint32(0x80000000)).astype(np.uint32)


def encode_keys(keys_f32: np.ndarray, ids: np.ndarray) -> np.ndarray:
    """(fp32 ranking key, id) -> the library's 64-bit key; padding (id == UINT64_MAX) -> KEY_EMPTY"""
    ids = np.asarray(ids, np.uint64)
    k = (orderable_u32(keys_f32).astype(np.uint64) << np.uint64(32)) | (ids & np.uint64(0xFFFFFFFF))
    return np.where(ids == np.uint64(2**64 - 1), np.uint64(2**64 - 1), k)


def decode_keys(keys: np.ndarray):
    """64-bit keys -> (fp32 ranking key, ids) like decode_keys_kernel"""
    keys = np.asarray(keys).view(np.uint64)
    o = (keys >> np.uint64(32)).astype(np.uint32)
    u = np.where(o >> 31 != 0, o & np.uint32(0x7FFFFFFF), ~o).astype(np.uint32)
    d = u.view(np.float32).copy()
    ids = (keys & np.uint64(0xFFFFFFFF)).astype(np.uint64)
    empty = keys == np.uint64(2**64 - 1)
    d[empty] = np.inf
    ids[empty] = np.uint64(2**64 - 1)
    return d, ids
