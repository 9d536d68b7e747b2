"""Mint golden vectors by RUNNING THE UNMODIFIED REFERENCE (oracle/_ref/*.so, compiled from /root/reference by
oracle/Makefile) on seeded inputs. The reference ships no tests or fixtures of its own, so these
files are the pins. Run in the build container:  python tests/golden/make_golden.py
Outputs (committed): tests/golden/*.npz and tests/golden/graph_u8_2000.index (the reference's own index format,
src/antitopo_engine.h:931-991)."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from oracle import pyoracle  # noqa: E402

pyoracle.build(ref=True)
ref_dyn = pyoracle.Reference(None)
ref_128 = pyoracle.Reference(128)


def save(name, **kw):
    np.savez_compressed(os.path.join(HERE, name + ".npz"), **kw)
    print("wrote", name, {k: getattr(v, "shape", v) for k, v in kw.items()})


# ---- 1. primitives ---------------------------------------------------------------------------------------
rng = np.random.default_rng(7)
prim = {}
for d in (16, 64, 96, 128, 256, 960):
    a = rng.standard_normal((8, d)).astype(np.float32) * 3
    b = rng.standard_normal((8, d)).astype(np.float32) * 3
    prim[f"a{d}"], prim[f"b{d}"] = a, b
    prim[f"eigen_dist2_{d}"] = np.array([ref_dyn.prim("ref_dist2", a[i], b[i]) for i in range(8)], np.float32)
    prim[f"eigen_dot_{d}"] = np.array([ref_dyn.prim("ref_dot", a[i], b[i]) for i in range(8)], np.float32)
    prim[f"avx_dist2_{d}"] = np.array([ref_dyn.prim("ref_dist2_f16", a[i], b[i]) for i in range(8)], np.float32)
    prim[f"avx_dot_{d}"] = np.array([ref_dyn.prim("ref_dot_f16", a[i], b[i]) for i in range(8)], np.float32)
    if d % 128 == 0:
        prim[f"avx128_dist2_{d}"] = np.array([ref_dyn.prim("ref_dist2_batch128", a[i], b[i]) for i in range(8)],
                                              np.float32)
    if d % 64 == 0:
        qi = rng.integers(0, 256, (8, d)).astype(np.float32) + rng.random((8, d)).astype(np.float32) * 0.9
        ci = rng.integers(0, 256, (8, d)).astype(np.uint8)
        prim[f"cq{d}"], prim[f"cc{d}"] = qi, ci
        prim[f"compressed_{d}"] = np.array(
            [ref_dyn.lib.ref_dist2_compressed(np.ascontiguousarray(qi[i]), np.ascontiguousarray(ci[i]), d)
             for i in range(8)], np.int32)
        i8a = rng.integers(-128, 128, (8, d)).astype(np.int8)
        i8b = rng.integers(-128, 128, (8, d)).astype(np.int8)
        prim[f"i8a{d}"], prim[f"i8b{d}"] = i8a, i8b
        prim[f"i8_asref_{d}"] = np.array([ref_dyn.lib.ref_dist2_i8(np.ascontiguousarray(i8a[i]),
                                                                    np.ascontiguousarray(i8b[i]), d) for i in range(8)],
                                         np.int32)
        i16a = rng.integers(-300, 300, (8, d)).astype(np.int16)
        i16b = rng.integers(-300, 300, (8, d)).astype(np.int16)
        prim[f"i16a{d}"], prim[f"i16b{d}"] = i16a, i16b
        prim[f"i16_asref_{d}"] = np.array([ref_dyn.lib.ref_dist2_i16(np.ascontiguousarray(i16a[i]),
                                                                      np.ascontiguousarray(i16b[i]), d)
                                           for i in range(8)], np.int32)
save("primitives", **prim)


# ---- 2. brute force --------------------------------------------------------------------------------------
def bf_case(name, base, q, k, ref):
    ids, dists, counts = ref.bf(base).query(q, k)
    store_base = base.astype(np.uint8) if np.all(base == np.floor(base)) and base.min() >= 0 and base.max() <= 255 else base
    store_q = q.astype(np.uint8) if np.all(q == np.floor(q)) and q.min() >= 0 and q.max() <= 255 else q
    save(name, base=store_base, q=store_q, k=np.int64(k), ids=ids, dists=dists, counts=counts)


rng = np.random.default_rng(11)
# planted exact ties: 64 rows, each vector appears 4 times, integer coordinates
u = rng.integers(0, 8, (16, 16)).astype(np.float32)
base = np.concatenate([u, u, u, u])[rng.permutation(64)]
bf_case("bf_ties_64x16", base, rng.integers(0, 8, (8, 16)).astype(np.float32), 10, ref_dyn)
# k > n and n < 32
bf_case("bf_tiny_5x16", rng.standard_normal((5, 16)).astype(np.float32), rng.standard_normal((3, 16)).astype(np.float32),
        10, ref_dyn)
# u8-valued (exact fp32 arithmetic): ids AND distances must match bit for bit
b = rng.integers(0, 256, (3000, 128)).astype(np.float32)
q = rng.integers(0, 256, (16, 128)).astype(np.float32)
bf_case("bf_u8_3000x128", b, q, 10, ref_128)
# duplicates inside u8 data: many equal distances
b2 = b[rng.integers(0, 40, 1500)]
bf_case("bf_u8dup_1500x128", b2, q[:8], 10, ref_128)
# gaussian (reference's own generator shape), fixed-DIM build and dynamic build
g = rng.standard_normal((3000, 128)).astype(np.float32)
gq = rng.standard_normal((16, 128)).astype(np.float32)
bf_case("bf_gauss_3000x128_dim128", g, gq, 10, ref_128)
bf_case("bf_gauss_3000x128_dyn", g, gq, 10, ref_dyn)
g96 = rng.standard_normal((2000, 96)).astype(np.float32)
g96 /= np.linalg.norm(g96, axis=1, keepdims=True)
bf_case("bf_deep_2000x96", g96, g96[:8] + 0.01 * rng.standard_normal((8, 96)).astype(np.float32), 10, ref_dyn)
g960 = rng.random((600, 960)).astype(np.float32)
bf_case("bf_gist_600x960", g960, rng.random((8, 960)).astype(np.float32), 10, ref_dyn)
bf_case("bf_k100_3000x128", g, gq[:4], 100, ref_128)

# ---- 3. graph engine: reference-built index + its own query results ---------------------------------------
sys.path.insert(0, ROOT)
from expann_b200 import datasets  # noqa: E402

gb, gq = datasets.make("sift", 2000, 32, 128, seed=5, variant="lowrank")
G = ref_128.graph(M=16, M0=32, ef_search_mult=1, ef_construction=64, ortho_count=1, ortho_factor=0.5, ortho_bias=0.0,
                  prune_overflow=0, use_compression=False)
G.store(gb)
G.build()
idx = os.path.join(HERE, "graph_u8_2000.index")
G.save(idx)
truth, _, _ = ref_128.bf(gb).query(gq, 10)
out = dict(q=gq.astype(np.uint8), truth=truth, info=np.array(list(G.info().values()), np.int64))
for ef in (10, 20, 40, 80):
    G.set_ef(ef)
    G.set_compression(False)
    G.reset_stats()
    ids, counts = G.query(gq, 10)
    out[f"ids_ef{ef}"], out[f"distcomps_ef{ef}"] = ids, np.int64(G.distcomps())
    G.set_compression(True)
    G.reset_stats()
    idc, _ = G.query(gq, 10)
    out[f"ids_comp_ef{ef}"] = idc
save("graph_u8_2000", **out)
print("index bytes", os.path.getsize(idx))
