"""GPU: one small invocation of every kernel family, for compute-sanitizer (memcheck / racecheck / synccheck):
  compute-sanitizer --tool memcheck python tools/sanitize_small.py
Families: K1 scan (1 / 4 / 8 queries per tile, L2 and IP, padded d), K3 single-CTA / pair / seeded / streamed / first form in
the exact, one-term and split formats, K6 SIMT and the kind::i8 code filter (k' <= 16 and > 16), K4, K7, K5 warp form and
CTA form (fp32 and u8 codes), graph build, the multi-device handle (two shards on one GPU). Shapes are tiny: the tools
slow kernels down 10-100x."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import expann_b200 as ex
rng = np.random.default_rng(0)
only = set(sys.argv[1:])


def want(name):
    return not only or name in only


def bf(base, q, k, metric=0, devices=None, **opts):
    e = ex.BruteForceEngine(metric=metric, devices=devices)
    for o, v in opts.items():
        e.set_option(o, v)
    e.store_many_vectors(base); e.build()
    out = e.query_k_batch(q, k); e.close()
    return out


u8 = rng.integers(0, 256, (3000, 128)).astype(np.float32)
qu8 = rng.integers(0, 256, (300, 128)).astype(np.float32)
real = rng.standard_normal((3000, 100)).astype(np.float32)
qreal = rng.standard_normal((300, 100)).astype(np.float32)
if want("k1"):
    for m in (1, 4, 9):
        for metric in (0, 1):
            bf(u8, qu8[:m], 10, metric=metric, tc=0); bf(real, qreal[:m], 10, metric=metric, tc=0)
    bf(u8, qu8[:20], 100, tc=0)
    print("k1 ok", flush=True)
if want("k3"):
    ref = bf(u8, qu8, 10, tc=0)
    for opts in (dict(), dict(tc_pair=0), dict(tc_boot=2), dict(tc_astream=2), dict(tc_v1=1)):
        out = bf(u8, qu8, 10, **opts)
        assert np.array_equal(out[0], ref[0]), opts
    bf(u8, qu8[:100], 10); bf(u8, qu8[:6], 10)
    rref = bf(real, qreal, 10, tc=0)
    for opts in (dict(tc_real=1), dict(tc_real=0), dict(tc_real=0, tc_astream=2)):
        out = bf(real, qreal, 10, **opts)
        assert np.array_equal(out[0], rref[0]), opts
    bf(u8[:, :64], qu8[:, :64], 100)      # first form, k > 16
    long = rng.random((1500, 960), dtype=np.float32)
    bf(long, long[:40] + 0.01, 10, metric=1, tc_real=0)
    print("k3 ok", flush=True)
if want("k6"):
    for quantizer, kp in ((0, 40), (0, 10), (1, 100)):
        for tc in (0, 1):
            e = ex.QuantScanEngine(quantizer=quantizer, kprime=kp, metric=quantizer)
            e.set_option("quant_tc", tc); e.set_option("quant_tc_min_batch", 8)
            b = u8 if quantizer == 0 else rng.random((6000, 192), dtype=np.float32)
            qq = qu8[:64] if quantizer == 0 else rng.random((64, 192), dtype=np.float32)
            e.store_many_vectors(b); e.build(); e.query_k_batch(qq, 10); e.query_k_batch(qq[:3], 10); e.close()
    print("k6 ok", flush=True)
if want("k5"):
    from expann_b200 import datasets
    gb, gq = datasets.make("sift", 3000, 64, 128, seed=5, variant="lowrank")
    for comp in (False, True):
        g = ex.GraphEngine(use_compression=comp)
        g.store_many_vectors(gb); g.build_graph(M=8, M0=16, K=32); g.build(); g.set_ef_search(30)
        g.set_option("graph_cta_batch", 0); g.query_k_batch(gq, 10)          # warp form
        g.set_option("graph_cta_batch", -1); g.query_k_batch(gq[:8], 10); g.query_k_batch(gq[:1], 10)   # CTA form
        g.set_option("graph_cta_width", 2); g.query_k_batch(gq[:8], 10)
        g.close()
    print("k5 + graph build ok", flush=True)
if want("multi"):
    ref = bf(u8, qu8, 10, tc=0)
    for rg in (0, 1):
        e = ex.BruteForceEngine(devices=[0, 0], row_groups=rg)
        e.store_many_vectors(u8); e.build()
        out = e.query_k_batch(qu8, 10); o1 = e.query_k_batch(qu8[:2], 10); e.close()
        assert np.array_equal(out[0], ref[0]) and np.array_equal(o1[0], ref[0][:2])
    print("multi ok", flush=True)
print("sanitize_small: done")
