"""Python host mirror of the reference engine interface over the C ABI.

Reference interface mirrored (same names, argument meaning, result order):
  ann_engine<T,Derived>::name/param_list/store_vector/build/query_k      src/ann_engine.h:16-29
  brute_force_engine<T>                                                  src/brute_force_engine.h:10-46
  pybind AntitopoEngine.store_many_vectors/query_k_numpy/set_ef_search   src/pyrunner.cpp:55-91
plus the batched entry (`query_k_batch`) the GPU needs; `query_k` is the m == 1 case of the same call.
All compute happens in libexpann_gpu.so on the GPU; there is no CPU fallback.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib
from ._lib import ExpannError, METRIC_IP, METRIC_L2


def _f32c(a):
    return np.ascontiguousarray(a, dtype=np.float32)


class GpuEngine:
    """Owns one expann_gpu_t handle (one engine instance on one GPU)."""

    _engine_kind = _lib.ENGINE_BRUTE_FORCE
    _name = "GPU Engine"

    def __init__(self, metric: int = METRIC_L2, device: int = -1, qtile: int = 0, id_offset: int = 0, devices=None,
                 row_groups: int = 0, **kw):
        """devices: list of CUDA ordinals -> ONE engine over several GPUs of this process (rows sharded by contiguous
        ranges, per-GPU top-k merged on devices[0]; an ordinal may repeat); row_groups: R = R row shards each held by
        len(devices) / R GPUs that split the queries (R = len(devices): one shard per GPU), 0 = automatic (shards of up to
        2^20 rows). Results do not depend on either."""
        self._lib = _lib.load()
        cfg = _lib.Cfg()
        cfg.engine = self._engine_kind
        cfg.metric = metric
        cfg.device = device
        cfg.qtile = qtile
        cfg.id_offset = id_offset
        cfg.M = kw.get("M", 0)
        cfg.M0 = kw.get("M0", 0)
        cfg.ef_search_mult = kw.get("ef_search_mult", 1)
        cfg.use_compression = int(kw.get("use_compression", False))
        cfg.quantizer = int(kw.get("quantizer", 0))
        cfg.row_groups = int(row_groups)
        h = C.c_void_p()
        if devices is not None:
            devs = (C.c_int32 * len(devices))(*[int(x) for x in devices])
            rc = self._lib.expann_gpu_create_multi(C.byref(cfg), devs, len(devices), C.byref(h))
        else:
            rc = self._lib.expann_gpu_create(C.byref(cfg), C.byref(h))
        if rc != 0:
            raise ExpannError(rc, self._lib.expann_gpu_last_error(None).decode())
        self._h = h
        self.devices = list(devices) if devices is not None else None
        self.metric = metric
        self.qtile = qtile
        self.id_offset = id_offset
        self.ef_search = 0
        self.d = None

    # -- helpers
    def _check(self, rc):
        if rc != 0:
            raise ExpannError(rc, self._lib.expann_gpu_last_error(self._h).decode())

    def close(self):
        if getattr(self, "_h", None):
            self._lib.expann_gpu_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- ann_engine interface (src/ann_engine.h:16-29)
    def name(self) -> str:
        return self._name

    def param_list(self) -> dict:
        st = self.stats()
        return {"metric": "ip" if self.metric == METRIC_IP else "l2", "qtile": str(st["last_qtile"] or self.qtile),
                "batched": "1", "device": "B200"}

    def store_vector(self, v) -> None:
        v = _f32c(v).reshape(1, -1)
        self.store_many_vectors(v)

    def store_many_vectors(self, rows, take_norms: bool = False) -> None:
        """src/pyrunner.cpp:70-81 (take_norms: L2-normalise on ingest, :78-79)"""
        rows = _f32c(rows)
        if rows.ndim != 2:
            raise ValueError("rows must be [n][d]")
        if take_norms:
            rows = rows / np.linalg.norm(rows, axis=1, keepdims=True).astype(np.float32)
            rows = _f32c(rows)
        self._check(self._lib.expann_gpu_add(self._h, rows.ctypes.data, rows.shape[0], rows.shape[1]))
        self.d = rows.shape[1]

    def store_device_rows(self, ptr: int, n: int, d: int, stream: int = 0) -> None:
        """rows already resident in HBM (e.g. a torch tensor's data_ptr())"""
        self.d = d
        self._check(self._lib.expann_gpu_add_device(self._h, ptr, n, d, stream))

    def build(self) -> None:
        self._check(self._lib.expann_gpu_build(self._h))

    def set_ef_search(self, ef: int) -> None:
        """src/antitopo_engine.h:189-195"""
        self.ef_search = int(ef)

    def query_k(self, v, k: int):
        """one query -> list of ids, nearest first (src/ann_engine.h:26-28)"""
        ids, _, counts = self.query_k_batch(_f32c(v).reshape(1, -1), k)
        return [int(x) for x in ids[0, :int(counts[0])]]

    query_k_numpy = query_k  # src/pyrunner.cpp:85-90

    def query_k_batch(self, q, k: int, ef: int | None = None, out=None):
        """m queries -> (ids [m][k] uint64 padded with UINT64_MAX, keys [m][k] float32 padded with +inf, counts [m]).
        `out` = (ids, keys, counts) preallocated C-contiguous arrays to fill instead; page-locked buffers (for q too)
        are copied by the GPU's copy engines directly, pageable ones through the handle's pinned staging."""
        q = _f32c(q)
        if q.ndim != 2:
            raise ValueError("q must be [m][d]")
        m = q.shape[0]
        if out is not None:
            ids, dists, counts = out
            if (ids.shape != (m, k) or ids.dtype != np.uint64 or dists.shape != (m, k) or dists.dtype != np.float32 or
                    counts.shape != (m,) or counts.dtype != np.uint64 or
                    not (ids.flags.c_contiguous and dists.flags.c_contiguous and counts.flags.c_contiguous)):
                raise ValueError("out must be (uint64 [m][k], float32 [m][k], uint64 [m]) C-contiguous arrays")
        else:
            ids = np.empty((m, k), np.uint64)
            dists = np.empty((m, k), np.float32)
            counts = np.empty(m, np.uint64)
        if self.d is not None and q.shape[1] != self.d:
            raise ValueError("query dimension mismatch")
        self._check(self._lib.expann_gpu_query(self._h, q.ctypes.data, m, k, self.ef_search if ef is None else ef,
                                               ids.ctypes.data, dists.ctypes.data, counts.ctypes.data))
        return ids, dists, counts

    # -- HBM-resident entry points (device pointers as ints, CUDA stream handle as int)
    def query_device(self, q_ptr: int, m: int, k: int, keys_out_ptr: int, stream: int = 0, ef: int | None = None):
        self._check(self._lib.expann_gpu_query_device(self._h, q_ptr, m, k, self.ef_search if ef is None else ef,
                                                      keys_out_ptr, stream))

    def scan_device_noalloc(self, q_ptr: int, m: int, k: int, keys_out_ptr: int, stream: int = 0):
        """the exact fp32 scan with its scratch memory in the library's static device array: no allocation in the call
        (expann_gpu_scan_device_noalloc; limits in include/expann_gpu.h)"""
        self._check(self._lib.expann_gpu_scan_device_noalloc(self._h, q_ptr, m, k, keys_out_ptr, stream))

    def merge_device(self, keys_in_ptr: int, m: int, nlists: int, klist: int, stride_q: int, stride_s: int, k: int,
                     keys_out_ptr: int, stream: int = 0):
        self._check(self._lib.expann_gpu_merge_device(self._h, keys_in_ptr, m, nlists, klist, stride_q, stride_s, k,
                                                      keys_out_ptr, stream))

    def decode_device(self, keys_ptr: int, m: int, k: int, ids_ptr: int, dist_ptr: int = 0, counts_ptr: int = 0,
                      stream: int = 0):
        self._check(self._lib.expann_gpu_decode_device(self._h, keys_ptr, m, k, ids_ptr, dist_ptr or None,
                                                       counts_ptr or None, stream))

    def set_timing(self, on: bool) -> None:
        self._check(self._lib.expann_gpu_set_timing(self._h, int(on)))

    def set_option(self, name: str, value: int) -> None:
        self._check(self._lib.expann_gpu_set_option(self._h, name.encode(), int(value)))

    def stats(self) -> dict:
        st = _lib.Stats()
        self._check(self._lib.expann_gpu_stats(self._h, C.byref(st)))
        return {f: getattr(st, f) for f, _ in st._fields_}


class BruteForceEngine(GpuEngine):
    """brute_force_engine<float> (src/brute_force_engine.h:10-46) on the GPU: exact k-NN by exhaustive scan,
    results ordered by (distance asc, id asc)."""
    _engine_kind = _lib.ENGINE_BRUTE_FORCE
    _name = "Brute-Force Engine (B200)"


class GraphEngine(GpuEngine):
    """Query half of antitopo_engine<float> (src/antitopo_engine.h:495-928) on the GPU: greedy descent through
    the upper layers, beam search of width ef on the bottom layer (optionally over u8 codes with an fp32 rerank).
    The graph comes from the reference's CPU build: `load_index(path)` reads the file written by serialize
    (src/antitopo_engine.h:931-991); `set_graph(...)` takes the same graph as CSR arrays.
    Constructor arguments mirror the pybind AntitopoEngine (src/pyrunner.cpp:55-62) where they matter for queries."""
    _engine_kind = _lib.ENGINE_GRAPH
    _name = "Anti-Topo Engine (B200 query path)"

    def __init__(self, M: int = 0, M0: int = 0, ef_search_mult: int = 1, use_compression: bool = False, **kw):
        super().__init__(M=M, M0=M0, ef_search_mult=ef_search_mult, use_compression=use_compression, **kw)
        self.use_compression = use_compression
        self.ef_search_mult = ef_search_mult

    def load_index(self, path: str) -> None:
        self._check(self._lib.expann_gpu_load_index(self._h, path.encode()))
        self.d = None

    def build_graph(self, M: int = 16, M0: int = 32, K: int = 0, ortho_factor: float = 0.5, ortho_bias: float = 0.0,
                    prune_overflow: int = 0, seed: int = 0) -> None:
        """construct the graph on the GPU from the stored vectors (argument names follow antitopo_engine_config,
        src/antitopo_engine.h:73-101; K plays the role of ef_construction: candidates per vertex)"""
        self._check(self._lib.expann_gpu_build_graph(self._h, M, M0, K, ortho_factor, ortho_bias, prune_overflow, seed))

    def save_index(self, path: str) -> None:
        self._check(self._lib.expann_gpu_save_index(self._h, path.encode()))

    def set_graph(self, nlayers, offsets, edges, starting_vertex: int, max_layer: int) -> None:
        nl = np.ascontiguousarray(nlayers, np.uint32)
        off = np.ascontiguousarray(offsets, np.uint64)
        ed = np.ascontiguousarray(edges, np.uint32)
        self._check(self._lib.expann_gpu_set_graph(self._h, nl.size, nl.ctypes.data, off.ctypes.data, ed.ctypes.data,
                                                   int(starting_vertex), int(max_layer)))

    def param_list(self) -> dict:
        pl = super().param_list()
        pl.update({"ef_search_mult": str(self.ef_search_mult), "use_compression": str(int(self.use_compression)),
                   "ef_search": str(self.ef_search)})
        return pl


QUANT_SIMPLE_U8, QUANT_RANGED_Q8 = 0, 1


class QuantScanEngine(GpuEngine):
    """Exhaustive scan over 8-bit codes + fp32 rerank: quantizer_simple<uint8_t>
    (src/quantizer.h:125-150; lossless on u8-valued data) or quantizer_ranged_q8 (src/quantizer.h:186-238, scored
    with the sign-corrected int8 distance). The k' = `kprime` best codes are re-ranked with exact fp32 keys."""
    _engine_kind = _lib.ENGINE_QUANT_SCAN
    _name = "Quantized Scan Engine (B200)"

    def __init__(self, quantizer: int = QUANT_SIMPLE_U8, kprime: int = 0, **kw):
        super().__init__(quantizer=quantizer, **kw)
        self.quantizer = quantizer
        if kprime:
            self.set_option("kprime", kprime)

    def quantizer_params(self):
        s, o = C.c_float(), C.c_float()
        self._check(self._lib.expann_gpu_get_quantizer(self._h, C.byref(s), C.byref(o)))
        return np.float32(s.value), np.float32(o.value)


PAIR_F32_L2, PAIR_F32_DOT, PAIR_U8_COMPRESSED, PAIR_I8, PAIR_I16 = range(5)


def pair_distances(kind: int, a, b, device: int = -1):
    """distance primitives on explicit pairs (rows of a vs rows of b) computed on the GPU; see include/expann_gpu.h"""
    lib = _lib.load()
    ta = {0: np.float32, 1: np.float32, 2: np.float32, 3: np.int8, 4: np.int16}[kind]
    tb = {0: np.float32, 1: np.float32, 2: np.uint8, 3: np.int8, 4: np.int16}[kind]
    a, b = np.ascontiguousarray(a, ta), np.ascontiguousarray(b, tb)
    n, d = a.shape
    out = np.empty(n, np.float32 if kind <= 1 else np.int32)
    rc = lib.expann_gpu_pair_distances(device, kind, a.ctypes.data, b.ctypes.data, n, d,
                                       out.ctypes.data if kind <= 1 else None, out.ctypes.data if kind > 1 else None)
    if rc != 0:
        raise ExpannError(rc, "pair_distances failed")
    return out
