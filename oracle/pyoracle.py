"""TEST INFRASTRUCTURE — ctypes loaders for oracle/liboracle.so (C restatement) and oracle/_ref/*.so
(the unmodified reference engines compiled by oracle/Makefile).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this module.
Nothing under expann_b200/ does.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
L2, IP = 0, 1
_u64p = np.ctypeslib.ndpointer(np.uint64, flags="C_CONTIGUOUS")
_f32p = np.ctypeslib.ndpointer(np.float32, flags="C_CONTIGUOUS")
_u8p = np.ctypeslib.ndpointer(np.uint8, flags="C_CONTIGUOUS")
_i8p = np.ctypeslib.ndpointer(np.int8, flags="C_CONTIGUOUS")
_i16p = np.ctypeslib.ndpointer(np.int16, flags="C_CONTIGUOUS")
_u32p = np.ctypeslib.ndpointer(np.uint32, flags="C_CONTIGUOUS")
_i32p = np.ctypeslib.ndpointer(np.int32, flags="C_CONTIGUOUS")


def build(ref: bool = True) -> None:
    """compile liboracle.so and (when /root/reference exists) oracle/_ref/*.so"""
    subprocess.check_call(["make", "-C", HERE, "oracle"] + (["ref"] if ref else []),
                          stdout=subprocess.DEVNULL)


def host_has_avx512() -> bool:
    try:
        with open("/proc/cpuinfo") as f:
            txt = f.read()
        flags = set(txt.split("flags", 1)[1].split("\n", 1)[0].split())
        return {"avx512f", "avx512bw", "avx512dq", "avx512vl"} <= flags
    except Exception:
        return False


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


# ------------------------------------------------------------------------------------------------------------
class Oracle:
    """C restatement (oracle/oracle.c)."""

    def __init__(self, path: str | None = None):
        path = path or os.path.join(HERE, "liboracle.so")
        if not os.path.exists(path):
            build(ref=False)
        self.lib = lib = C.CDLL(path)
        lib.orc_dist2_f32.restype = C.c_float
        lib.orc_dist2_f32.argtypes = [_f32p, _f32p, C.c_size_t]
        lib.orc_dot_f32.restype = C.c_float
        lib.orc_dot_f32.argtypes = [_f32p, _f32p, C.c_size_t]
        lib.orc_dist2_u8.restype = C.c_int32
        lib.orc_dist2_u8.argtypes = [_f32p, _u8p, C.c_size_t]
        for nm, pt in (("orc_dist2_i8_asref", _i8p), ("orc_dist2_i8", _i8p),
                       ("orc_dist2_i16_asref", _i16p), ("orc_dist2_i16", _i16p)):
            f = getattr(lib, nm)
            f.restype = C.c_int32
            f.argtypes = [pt, pt, C.c_size_t]
        lib.orc_quant_simple_u8.argtypes = [_f32p, C.c_size_t, C.c_size_t, _u8p]
        lib.orc_quant_ranged_q8_fit.argtypes = [_f32p, C.c_size_t, C.c_size_t, C.POINTER(C.c_float),
                                                C.POINTER(C.c_float)]
        lib.orc_quant_ranged_q8_encode.argtypes = [_f32p, C.c_size_t, C.c_size_t, C.c_float, C.c_float, _i8p]
        lib.orc_topk_create.restype = C.c_void_p
        lib.orc_topk_create.argtypes = [C.c_size_t]
        lib.orc_topk_destroy.argtypes = [C.c_void_p]
        lib.orc_topk_consider.argtypes = [C.c_void_p, C.c_float, C.c_uint64]
        lib.orc_topk_consider.restype = C.c_int
        lib.orc_topk_discard_until_size.argtypes = [C.c_void_p, C.c_size_t]
        lib.orc_topk_size.argtypes = [C.c_void_p]
        lib.orc_topk_size.restype = C.c_size_t
        lib.orc_topk_worst.argtypes = [C.c_void_p]
        lib.orc_topk_worst.restype = C.c_uint64
        lib.orc_topk_worst_val.argtypes = [C.c_void_p]
        lib.orc_topk_worst_val.restype = C.c_float
        lib.orc_topk_at_capacity.argtypes = [C.c_void_p]
        lib.orc_topk_to_combined.argtypes = [C.c_void_p, _f32p, _u64p]
        lib.orc_topk_to_combined.restype = C.c_size_t
        lib.orc_bf_query.argtypes = [_f32p, C.c_size_t, C.c_size_t, _f32p, C.c_size_t, C.c_size_t, C.c_int,
                                     _u64p, _f32p, _u64p, C.c_int]
        lib.orc_scan_u8_query.argtypes = [_f32p, _u8p, C.c_size_t, C.c_size_t, _f32p, C.c_size_t, C.c_size_t,
                                          C.c_size_t, C.c_int, _u64p, _f32p, C.c_void_p, C.c_void_p, C.c_int]
        lib.orc_graph_load.restype = C.c_void_p
        lib.orc_graph_load.argtypes = [C.c_char_p]
        lib.orc_graph_from_arrays.restype = C.c_void_p
        lib.orc_graph_from_arrays.argtypes = [_f32p, C.c_size_t, C.c_size_t, _u32p, _u64p, _u32p, C.c_uint64,
                                              C.c_uint64, C.c_uint64, C.c_uint64]
        lib.orc_graph_destroy.argtypes = [C.c_void_p]
        lib.orc_graph_info.restype = C.c_uint64
        lib.orc_graph_info.argtypes = [C.c_void_p, C.c_int]
        lib.orc_graph_query.argtypes = [C.c_void_p, _f32p, C.c_size_t, C.c_size_t, C.c_size_t, C.c_int, _u64p,
                                        _f32p, _u64p, C.POINTER(C.c_uint64)]
        lib.orc_graph_query_mt.argtypes = [C.c_void_p, _f32p, C.c_size_t, C.c_size_t, C.c_size_t, C.c_int,
                                           _u64p, _u64p, C.c_int]
        lib.orc_graph_neighbours.restype = C.c_size_t
        lib.orc_graph_neighbours.argtypes = [C.c_void_p, C.c_uint64, C.c_uint32, _u32p, C.c_size_t]

    # primitives
    def dist2(self, a, b):
        a, b = _f32(a), _f32(b)
        return np.float32(self.lib.orc_dist2_f32(a, b, a.size))

    def dot(self, a, b):
        a, b = _f32(a), _f32(b)
        return np.float32(self.lib.orc_dot_f32(a, b, a.size))

    def dist2_u8(self, q, code):
        q = _f32(q)
        code = np.ascontiguousarray(code, np.uint8)
        return int(self.lib.orc_dist2_u8(q, code, q.size))

    def int_dist(self, name, a, b):
        return int(getattr(self.lib, name)(a, b, a.size))

    def quant_simple_u8(self, rows):
        rows = _f32(rows)
        out = np.empty(rows.shape, np.uint8)
        self.lib.orc_quant_simple_u8(rows, rows.shape[0], rows.shape[1], out)
        return out

    def quant_ranged_q8(self, rows, encode_rows=None):
        rows = _f32(rows)
        s, o = C.c_float(), C.c_float()
        self.lib.orc_quant_ranged_q8_fit(rows, rows.shape[0], rows.shape[1], C.byref(s), C.byref(o))
        enc = rows if encode_rows is None else _f32(encode_rows)
        out = np.empty(enc.shape, np.int8)
        self.lib.orc_quant_ranged_q8_encode(enc, enc.shape[0], enc.shape[1], s, o, out)
        return out, np.float32(s.value), np.float32(o.value)

    def bf_query(self, base, q, k, metric=L2, threads=1):
        base, q = _f32(base), _f32(q)
        n, d = base.shape
        m = q.shape[0]
        ids = np.empty((m, k), np.uint64)
        dists = np.empty((m, k), np.float32)
        counts = np.empty(m, np.uint64)
        self.lib.orc_bf_query(base, n, d, q, m, k, metric, ids, dists, counts, threads)
        return ids, dists, counts

    def scan_u8_query(self, base, codes, q, k, kprime, metric=L2, threads=1, want_cand=False):
        base, q = _f32(base), _f32(q)
        codes = np.ascontiguousarray(codes, np.uint8)
        n, d = base.shape
        m = q.shape[0]
        kprime = max(kprime, k)
        ids = np.empty((m, k), np.uint64)
        dists = np.empty((m, k), np.float32)
        scores = np.empty((m, kprime), np.int32)
        cand = np.empty((m, kprime), np.uint64)
        self.lib.orc_scan_u8_query(base, codes, n, d, q, m, k, kprime, metric, ids, dists,
                                   scores.ctypes.data, cand.ctypes.data, threads)
        if want_cand:
            return ids, dists, scores, cand
        return ids, dists

    def topk(self, k):
        return _TopK(self.lib, k)

    def graph_load(self, path):
        h = self.lib.orc_graph_load(path.encode())
        if not h:
            raise IOError(f"cannot load index {path}")
        return _Graph(self.lib, h)

    def graph_from_arrays(self, vectors, nlayers, offsets, edges, start, max_layer, M, M0):
        vectors = _f32(vectors)
        h = self.lib.orc_graph_from_arrays(vectors, vectors.shape[0], vectors.shape[1],
                                           np.ascontiguousarray(nlayers, np.uint32),
                                           np.ascontiguousarray(offsets, np.uint64),
                                           np.ascontiguousarray(edges, np.uint32), start, max_layer, M, M0)
        return _Graph(self.lib, h)


class _TopK:
    def __init__(self, lib, k):
        self.lib, self.k = lib, k
        self.h = lib.orc_topk_create(k)

    def consider(self, d, v):
        return bool(self.lib.orc_topk_consider(self.h, float(d), int(v)))

    def size(self):
        return self.lib.orc_topk_size(self.h)

    def worst(self):
        return self.lib.orc_topk_worst(self.h)

    def worst_val(self):
        return self.lib.orc_topk_worst_val(self.h)

    def at_capacity(self):
        return bool(self.lib.orc_topk_at_capacity(self.h))

    def discard_until_size(self, g):
        self.lib.orc_topk_discard_until_size(self.h, g)

    def to_combined(self):
        n = self.size()
        d = np.empty(max(n, 1), np.float32)
        i = np.empty(max(n, 1), np.uint64)
        self.lib.orc_topk_to_combined(self.h, d, i)
        return d[:n], i[:n]

    def __del__(self):
        try:
            self.lib.orc_topk_destroy(self.h)
        except Exception:
            pass


class _Graph:
    def __init__(self, lib, h):
        self.lib, self.h = lib, h
        self.start, self.max_layer, self.n, self.M, self.M0, self.d = (
            int(lib.orc_graph_info(h, i)) for i in range(6))

    def query(self, q, k, ef, compressed=False):
        q = _f32(q)
        m = q.shape[0]
        ids = np.empty((m, k), np.uint64)
        dists = np.empty((m, k), np.float32)
        counts = np.empty(m, np.uint64)
        dc = C.c_uint64(0)
        self.lib.orc_graph_query(self.h, q, m, k, ef, int(compressed), ids, dists, counts, C.byref(dc))
        return ids, dists, counts, dc.value

    def query_mt(self, q, k, ef, compressed=False, threads=1):
        q = _f32(q)
        m = q.shape[0]
        ids = np.empty((m, k), np.uint64)
        counts = np.empty(m, np.uint64)
        self.lib.orc_graph_query_mt(self.h, q, m, k, ef, int(compressed), ids, counts, threads)
        return ids, counts

    def neighbours(self, v, layer=0, cap=4096):
        out = np.empty(cap, np.uint32)
        deg = self.lib.orc_graph_neighbours(self.h, v, layer, out, cap)
        return out[:deg].copy()

    def __del__(self):
        try:
            self.lib.orc_graph_destroy(self.h)
        except Exception:
            pass


# ------------------------------------------------------------------------------------------------------------
class Reference:
    """the UNMODIFIED reference engines (oracle/_ref/libexpann_ref_{dyn,128}.so)."""

    def __init__(self, dim: int | None = None):
        name = "libexpann_ref_128.so" if dim == 128 else "libexpann_ref_dyn.so"
        path = os.path.join(HERE, "_ref", name)
        if not os.path.exists(path):
            raise FileNotFoundError(path)
        if not host_has_avx512():
            raise RuntimeError("reference build needs an AVX-512 host (src/distance.h uses it unconditionally)")
        self.lib = lib = C.CDLL(path)
        self.dim = dim
        for nm in ("ref_dist2", "ref_dot", "ref_dist2_batch128", "ref_dist2_f16", "ref_dot_f16"):
            f = getattr(lib, nm)
            f.restype = C.c_float
            f.argtypes = [_f32p, _f32p, C.c_size_t]
        lib.ref_dist2_i16.restype = C.c_int32
        lib.ref_dist2_i16.argtypes = [_i16p, _i16p, C.c_size_t]
        lib.ref_dist2_i8.restype = C.c_int32
        lib.ref_dist2_i8.argtypes = [_i8p, _i8p, C.c_size_t]
        lib.ref_dist2_compressed.restype = C.c_int32
        lib.ref_dist2_compressed.argtypes = [_f32p, _u8p, C.c_size_t]
        lib.ref_bf_create.restype = C.c_void_p
        lib.ref_bf_destroy.argtypes = [C.c_void_p]
        lib.ref_bf_store.argtypes = [C.c_void_p, _f32p, C.c_size_t, C.c_size_t]
        lib.ref_bf_build.argtypes = [C.c_void_p]
        lib.ref_bf_query.argtypes = [C.c_void_p, _f32p, C.c_size_t, C.c_size_t, _u64p, _f32p, _u64p, C.c_int]
        lib.ref_graph_create.restype = C.c_void_p
        lib.ref_graph_create.argtypes = [C.c_size_t] * 5 + [C.c_float, C.c_float, C.c_size_t, C.c_int]
        lib.ref_graph_destroy.argtypes = [C.c_void_p]
        lib.ref_graph_store.argtypes = [C.c_void_p, _f32p, C.c_size_t, C.c_size_t]
        lib.ref_graph_build.argtypes = [C.c_void_p]
        lib.ref_graph_set_ef.argtypes = [C.c_void_p, C.c_size_t]
        lib.ref_graph_set_compression.argtypes = [C.c_void_p, C.c_int]
        lib.ref_graph_save.argtypes = [C.c_void_p, C.c_char_p]
        lib.ref_graph_load.argtypes = [C.c_void_p, C.c_char_p, C.c_size_t]
        lib.ref_graph_distcomps.restype = C.c_uint64
        lib.ref_graph_distcomps.argtypes = [C.c_void_p]
        lib.ref_graph_reset_stats.argtypes = [C.c_void_p]
        lib.ref_graph_query.argtypes = [C.c_void_p, _f32p, C.c_size_t, C.c_size_t, _u64p, _u64p]
        lib.ref_graph_info.restype = C.c_uint64
        lib.ref_graph_info.argtypes = [C.c_void_p, C.c_int]

    def prim(self, name, a, b):
        a, b = _f32(a), _f32(b)
        return np.float32(getattr(self.lib, name)(a, b, a.size))

    def bf(self, base):
        return _RefBF(self.lib, _f32(base))

    def graph(self, M, M0, ef_search_mult=1, ef_construction=64, ortho_count=1, ortho_factor=0.5,
              ortho_bias=0.0, prune_overflow=0, use_compression=False):
        h = self.lib.ref_graph_create(M, M0, ef_search_mult, ef_construction, ortho_count, ortho_factor,
                                      ortho_bias, prune_overflow, int(use_compression))
        return _RefGraph(self.lib, h)


class _RefBF:
    def __init__(self, lib, base):
        self.lib = lib
        self.h = lib.ref_bf_create()
        lib.ref_bf_store(self.h, base, base.shape[0], base.shape[1])
        lib.ref_bf_build(self.h)

    def query(self, q, k, threads=1):
        q = _f32(q)
        m = q.shape[0]
        ids = np.empty((m, k), np.uint64)
        dists = np.empty((m, k), np.float32)
        counts = np.empty(m, np.uint64)
        self.lib.ref_bf_query(self.h, q, m, k, ids, dists, counts, threads)
        return ids, dists, counts

    def __del__(self):
        try:
            self.lib.ref_bf_destroy(self.h)
        except Exception:
            pass


class _RefGraph:
    def __init__(self, lib, h):
        self.lib, self.h = lib, h

    def store(self, rows):
        rows = _f32(rows)
        self.lib.ref_graph_store(self.h, rows, rows.shape[0], rows.shape[1])

    def build(self):
        self.lib.ref_graph_build(self.h)

    def set_ef(self, ef):
        self.lib.ref_graph_set_ef(self.h, ef)

    def set_compression(self, on):
        self.lib.ref_graph_set_compression(self.h, int(on))

    def save(self, path):
        if self.lib.ref_graph_save(self.h, path.encode()) != 0:
            raise IOError(path)

    def load(self, path, d):
        if self.lib.ref_graph_load(self.h, path.encode(), d) != 0:
            raise IOError(path)

    def query(self, q, k):
        q = _f32(q)
        m = q.shape[0]
        ids = np.empty((m, k), np.uint64)
        counts = np.empty(m, np.uint64)
        self.lib.ref_graph_query(self.h, q, m, k, ids, counts)
        return ids, counts

    def distcomps(self):
        return int(self.lib.ref_graph_distcomps(self.h))

    def reset_stats(self):
        self.lib.ref_graph_reset_stats(self.h)

    def info(self):
        return {k: int(self.lib.ref_graph_info(self.h, i)) for i, k in
                enumerate(("start", "max_layer", "n", "M", "M0"))}

    def __del__(self):
        try:
            self.lib.ref_graph_destroy(self.h)
        except Exception:
            pass
