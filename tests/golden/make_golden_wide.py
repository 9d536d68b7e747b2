"""Wider pins of the distance primitives the reference ships (src/distance.h:14-190, src/antitopo_engine.h:38-61,
src/vec.h:109-155): 256 pairs per dimension, d = 832 from the reference's compiled dimension set (CMakeLists.txt:100-153)
added, and rows of extreme values (1e15-scale, denormal-scale, integer-valued, identical vectors, zeros). Minted by
RUNNING THE UNMODIFIED REFERENCE (oracle/_ref/libexpann_ref_dyn.so) like make_golden.py; kept in its own file so that
the first-round goldens stay byte-stable.  Run in the build container:  python tests/golden/make_golden_wide.py"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from oracle import pyoracle  # noqa: E402

pyoracle.build(ref=True)
ref = pyoracle.Reference(None)
from tests.util import WIDE_DIMS, wide_primitive_inputs  # noqa: E402

NP = 256
inp = wide_primitive_inputs()   # the tests regenerate the same arrays from the seed; only the reference's outputs are stored
out = {}
for d in WIDE_DIMS:
    a, b = inp[f"a{d}"], inp[f"b{d}"]
    for name, fn in (("eigen_dist2", "ref_dist2"), ("eigen_dot", "ref_dot"), ("avx_dist2", "ref_dist2_f16"), ("avx_dot", "ref_dot_f16")):
        out[f"{name}_{d}"] = np.array([ref.prim(fn, a[i], b[i]) for i in range(NP)], np.float32)
    if d % 128 == 0:
        out[f"avx128_dist2_{d}"] = np.array([ref.prim("ref_dist2_batch128", a[i], b[i]) for i in range(NP)], np.float32)
    qi, ci = inp[f"cq{d}"], inp[f"cc{d}"]
    out[f"compressed_{d}"] = np.array([ref.lib.ref_dist2_compressed(np.ascontiguousarray(qi[i]), np.ascontiguousarray(ci[i]), d)
                                       for i in range(NP)], np.int32)
    i8a, i8b = inp[f"i8a{d}"], inp[f"i8b{d}"]
    out[f"i8_asref_{d}"] = np.array([ref.lib.ref_dist2_i8(np.ascontiguousarray(i8a[i]), np.ascontiguousarray(i8b[i]), d) for i in range(NP)], np.int32)
    i16a, i16b = inp[f"i16a{d}"], inp[f"i16b{d}"]
    out[f"i16_asref_{d}"] = np.array([ref.lib.ref_dist2_i16(np.ascontiguousarray(i16a[i]), np.ascontiguousarray(i16b[i]), d) for i in range(NP)], np.int32)
np.savez_compressed(os.path.join(HERE, "primitives_wide.npz"), **out)
print("wrote primitives_wide", {k: v.shape for k, v in out.items()})
