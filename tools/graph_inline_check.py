"""Latency form of the graph search on neighbourhood blocks (graph_inline) against the default latency form:
identical ids / distances / counts, and microseconds per one-query call. usage: graph_inline_check.py [n] [d]"""
import os, sys, time, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import expann_b200
from expann_b200 import datasets

n = int(sys.argv[1]) if len(sys.argv) > 1 else 20000
d = int(sys.argv[2]) if len(sys.argv) > 2 else 128
nq = 512
base, q = datasets.make("sift", n, nq, d, variant="lowrank")
eng = {}
for inl in (0, 1):
    e = expann_b200.GraphEngine(use_compression=bool(int(os.environ.get("COMPRESSED", "0"))))
    e.set_option("graph_inline", inl)
    e.store_many_vectors(base)
    t = time.time(); e.build_graph(M=16, M0=32, K=64); e.build(); tb = time.time() - t
    eng[inl] = e
    print(f"inline={inl} build {tb:.2f}s", flush=True)
out = {"n": n, "d": d}
for ef in (10, 40, 100, 200):
    res = {}
    for inl in (0, 1):
        e = eng[inl]; e.set_ef_search(ef)
        dc0 = e.stats()["distcomps"]
        one = [e.query_k_batch(q[i:i + 1], 10) for i in range(128)]
        dc1 = e.stats()["distcomps"] - dc0
        many = e.query_k_batch(q[:128], 10)
        res[inl] = (np.concatenate([o[0] for o in one]), np.concatenate([o[1] for o in one]), np.concatenate([o[2] for o in one]), many, dc1)
    a, b = res[0], res[1]
    ok = np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1]) and np.array_equal(a[2], b[2])
    okm = all(np.array_equal(x, y) for x, y in zip(a[3], b[3])) and np.array_equal(b[3][0], b[0])
    print(f"ef={ef} one-query identical={ok} batch identical={okm} distcomps {a[4]} vs {b[4]}", flush=True)
    out[f"identical_ef{ef}"] = bool(ok and okm and a[4] == b[4])
    for inl in (0, 1):
        e = eng[inl]
        for i in range(50): e.query_k_batch(q[i:i + 1], 10)
        t = time.perf_counter()
        R = 400
        for i in range(R): e.query_k_batch(q[i % nq:i % nq + 1], 10)
        us = (time.perf_counter() - t) / R * 1e6
        print(f"  ef={ef} inline={inl}: {us:.1f} us per one-query call", flush=True)
        out[f"us_ef{ef}_inline{inl}"] = round(us, 1)
print(json.dumps(out))
for e in eng.values(): e.close()
