"""ctypes binding of libexpann_gpu.so (C ABI: include/expann_gpu.h). No CPU fallback: if the library or a
B200 is missing, construction of an engine raises."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
# EXPANN_GPU_LIB: another build of the same library (kernel A/B experiments under tools/)
LIB_PATH = os.environ.get("EXPANN_GPU_LIB") or os.path.join(HERE, "libexpann_gpu.so")

OK, ERR_ARG, ERR_CUDA, ERR_STATE, ERR_UNSUPPORTED, ERR_IO = 0, -1, -2, -3, -4, -5
ENGINE_BRUTE_FORCE, ENGINE_GRAPH, ENGINE_QUANT_SCAN = 0, 1, 2
METRIC_L2, METRIC_IP = 0, 1


class Cfg(C.Structure):
    _fields_ = [("engine", C.c_int32), ("metric", C.c_int32), ("device", C.c_int32), ("qtile", C.c_int32),
                ("id_offset", C.c_uint64), ("M", C.c_uint64), ("M0", C.c_uint64), ("ef_search_mult", C.c_uint64),
                ("use_compression", C.c_int32), ("quantizer", C.c_int32), ("n_devices", C.c_int32), ("row_groups", C.c_int32),
                ("reserved", C.c_int32 * 4)]


class Stats(C.Structure):
    _fields_ = [("n", C.c_uint64), ("d", C.c_uint64), ("d_pad", C.c_uint64), ("scan_launches", C.c_uint64),
                ("merge_launches", C.c_uint64), ("graph_launches", C.c_uint64), ("aux_launches", C.c_uint64),
                ("last_nsplit", C.c_uint64), ("last_qtile", C.c_uint64), ("distcomps", C.c_uint64),
                ("hot_time_ns", C.c_uint64), ("hot_timed_launches", C.c_uint64), ("tc_launches", C.c_uint64),
                ("tc_fallbacks", C.c_uint64), ("tc_mode", C.c_uint64), ("peer_bytes", C.c_uint64), ("query_groups", C.c_uint64),
                ("row_groups", C.c_uint64)]


# every symbol include/expann_gpu.h declares: name -> (restype, argtypes)
_vp, _sz, _u64 = C.c_void_p, C.c_size_t, C.c_uint64
SYMBOLS = {
    "expann_gpu_create": (C.c_int, [C.POINTER(Cfg), C.POINTER(_vp)]),
    "expann_gpu_create_multi": (C.c_int, [C.POINTER(Cfg), C.POINTER(C.c_int32), _sz, C.POINTER(_vp)]),
    "expann_gpu_destroy": (None, [_vp]),
    "expann_gpu_last_error": (C.c_char_p, [_vp]),
    "expann_gpu_add": (C.c_int, [_vp, _vp, _sz, _sz]),
    "expann_gpu_add_device": (C.c_int, [_vp, _vp, _sz, _sz, _vp]),
    "expann_gpu_load_index": (C.c_int, [_vp, C.c_char_p]),
    "expann_gpu_set_graph": (C.c_int, [_vp, _sz, _vp, _vp, _vp, C.c_uint64, C.c_uint64]),
    "expann_gpu_build_graph": (C.c_int, [_vp, C.c_uint64, C.c_uint64, C.c_uint64, C.c_float, C.c_float, C.c_uint64, C.c_uint64]),
    "expann_gpu_save_index": (C.c_int, [_vp, C.c_char_p]),
    "expann_gpu_build": (C.c_int, [_vp]),
    "expann_gpu_query": (C.c_int, [_vp, _vp, _sz, _sz, _sz, _vp, _vp, _vp]),
    "expann_gpu_query_device": (C.c_int, [_vp, _vp, _sz, _sz, _sz, _vp, _vp]),
    "expann_gpu_scan_device_noalloc": (C.c_int, [_vp, _vp, _sz, _sz, _vp, _vp]),
    "expann_gpu_merge_device": (C.c_int, [_vp, _vp, _sz, _sz, _sz, _sz, _sz, _sz, _vp, _vp]),
    "expann_gpu_decode_device": (C.c_int, [_vp, _vp, _sz, _sz, _vp, _vp, _vp, _vp]),
    "expann_gpu_get_quantizer": (C.c_int, [_vp, C.POINTER(C.c_float), C.POINTER(C.c_float)]),
    "expann_gpu_pair_distances": (C.c_int, [C.c_int, C.c_int, _vp, _vp, _sz, _sz, _vp, _vp]),
    "expann_gpu_set_timing": (C.c_int, [_vp, C.c_int]),
    "expann_gpu_stats": (C.c_int, [_vp, C.POINTER(Stats)]),
    "expann_gpu_set_option": (C.c_int, [_vp, C.c_char_p, C.c_int64]),
    "expann_gpu_version": (C.c_char_p, []),
}

_lib = None


def build_library() -> str:
    """compile libexpann_gpu.so in-tree (nvcc -gencode arch=compute_100a,code=sm_100a)"""
    subprocess.check_call(["make", "-C", os.path.join(HERE, "csrc"), "-j8"], stdout=subprocess.DEVNULL)
    return LIB_PATH


def load():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                           "(there is no CPU fallback)")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SYMBOLS.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


class ExpannError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"expann_gpu error {code}: {msg}")
        self.code = code
