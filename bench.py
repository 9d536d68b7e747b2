#!/usr/bin/env python
"""bench.py — headline benchmark of the expANN query hot path on B200 (contract: see DESIGN.md §Measurement).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload c2|c1|c4|c5]

A "step" is one pass of the hot path over one batch of m synthetic queries against the HBM-resident base set.
  value : QPS, device-resident queries in, merged 64-bit (dist,id) keys out (CUDA events on the launching stream)
  e2e   : QPS through the reference-facing host-buffer call (expann_gpu_query), H2D of the queries and D2H of the
          results inside the timed region; page-locked caller buffers (headline) and pageable ones (`pageable_value`)
  roofline : the dominant kernel of the step (algorithmic flops or bytes / CUDA-event time recorded inside the library on
          the launching stream) plus, as sub-objects the driver keeps whole: `hbm_scan` (the HBM-bound one-query-tile
          passes of the fp32 scan), `graph` (BASELINE config 3), `c4` / `c5` (the other named configs, bounded legs)
  cpu_baseline : the UNMODIFIED reference brute_force_engine (oracle/_ref) — or the C restatement when that
          build is not usable on this host — timed on a bounded query sample with all host threads
N > 1: ONE process drives all N GPUs through ONE multi-device handle (expann_gpu_create_multi, csrc/multi.cu): the base
set is sharded by contiguous row ranges inside the library, the queries fan out over NVLink, every GPU's last kernel
stores its [m][k] keys into GPU 0's gather buffer and K7 merges there — the path a reference-side caller gets. Under the
driver's torchrun launch rank 0 owns the handle; the other ranks initialise the N-rank NCCL communicator with it and then
only join the CPU-side barriers that bracket the timed region (they hold no data and launch nothing: an NCCL barrier
kernel spinning on GPU r would be time-sliced against the handle's own kernels there), so `value` is the whole job's
throughput.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: (generator, variant, n, m, d, k, metric, description)
    "c1": ("gauss", None, 56_000, 400, 128, 10, "l2", "config_synthetic.json shape 56000x128 gaussian, m=400 (L2-resident)"),
    "c2": ("sift", "iid", 1_000_000, 10_000, 128, 10, "l2", "SIFT-shape synthetic 1Mx128 fp32 (u8-valued iid), 10k queries, brute-force L2 exact top-10"),
    "c4": ("deep", None, 10_000_000, 10_000, 96, 10, "l2", "Deep-shape synthetic 10Mx96 fp32, 10k queries, sharded brute-force L2 top-10"),
    "c5": ("gist", None, 1_000_000, 1_000, 960, 10, "ip", "GIST-shape synthetic 1Mx960 fp32, 1k queries, brute-force inner-product top-10"),
}
T_START = time.perf_counter()


def ncu_traffic(key):
    """DRAM bytes (read + write) per launch of a kernel from the committed ncu --set full capture of the same kernel on
    the C2 shape (profiles/*_ncu_summary.json, newest round first); (None, None) when there is no capture"""
    for name in ("r2_ncu_summary.json", "r1_ncu_summary.json"):
        try:
            j = json.load(open(os.path.join(ROOT, "profiles", name)))
            for kname, d in j.items():
                if kname.startswith(key):
                    unit = {"Mbyte": 1e6, "Gbyte": 1e9, "Kbyte": 1e3, "byte": 1.0}
                    r, w = d["dram__bytes_read.sum"], d["dram__bytes_write.sum"]
                    return int(float(r[0]) * unit[r[1]] + float(w[0]) * unit[w[1]]), "committed ncu --set full capture, profiles/" + name
        except Exception:  # noqa: BLE001
            pass
    return None, None


def tensor_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            j = json.load(open(p))
            return float(j["bf16_tflops"]), float(j.get("bf16_tflops_sustained", j["bf16_tflops"])), "measured (MEASURED_PEAKS.json, cuBLAS bf16)"
        except Exception:  # noqa: BLE001
            pass
    return 1590.0, 1400.0, "fallback (B200_PROFILING.md)"


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
        except Exception:  # noqa: BLE001
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)"""

    def __init__(self, index):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={q}", "--format=csv,noheader,nounits", "-lms", "50",
                                          "-i", str(self.index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
            time.sleep(0.3)   # let the first samples arrive before the timed region starts
        except Exception:  # noqa: BLE001
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.perf_counter(), [x.strip() for x in line.split(",")]))

    def window(self, t0, t1):
        return [r for t, r in self.rows if t0 <= t <= t1 + 0.06]

    def stop(self, t0=None, t1=None):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:  # noqa: BLE001
            self.proc.kill()
        rows = self.window(t0, t1) if t0 is not None else [r for _, r in self.rows]
        sm = [float(r[0]) for r in rows if r and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        pw = [float(r[2]) for r in rows if len(r) > 2 and r[2].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in rows if len(r) >= 7 for i in range(4) if r[3 + i].lower().startswith("active")})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm), "power_w_max": max(pw) if pw else None}


def cpu_reference_qps(base, q, k, metric, budget_s=15.0):
    """time the reference's own CPU brute force (oracle/_ref) on a bounded query sample, all host threads.
    Returns (qps, cores, kind, sample description)."""
    from oracle import pyoracle
    threads = os.cpu_count() or 1
    d = base.shape[1]
    kind, eng = "port", None
    if metric == "l2":
        try:
            ref = pyoracle.Reference(128 if d == 128 else None)
            eng = ref.bf(base)  # store_vector x n + build (not timed as query time)
            kind = "reference"
        except Exception:  # noqa: BLE001
            eng = None
    orc = pyoracle.Oracle() if eng is None else None

    def run(qs):
        t0 = time.perf_counter()
        if eng is not None:
            eng.query(qs, k, threads=threads)
        else:
            orc.bf_query(base, qs, k, metric=1 if metric == "ip" else 0, threads=threads)
        return time.perf_counter() - t0

    probe = q[:min(len(q), threads)]
    t = run(probe)
    per_q = t / len(probe)
    msample = int(max(len(probe), min(len(q), budget_s / max(per_q, 1e-9))))
    msample = max(threads, msample // threads * threads) if msample >= threads else msample
    msample = min(msample, len(q))
    t = run(q[:msample])
    return msample / t, threads, kind, f"{msample} of {len(q)} queries, full {base.shape[0]}x{d} base, {threads} threads, {t:.1f}s"


def scan_hbm_leg(eng, torch, dq, k, n_rows, d, stream, peak, peak_src, steps):
    """HBM-bound regime of the path: ONE pass of the streaming fp32 scan (K1) over the base for a tile of 1, 4 and 8
    queries (brute_force_engine::_query_k, src/brute_force_engine.h:29-46, is one such pass per query). L2 is
    NOT flushed: the input of a pass (512 MB at C2) is 4x the L2 capacity, so back-to-back passes stream it from DRAM
    again. The headline numbers of this object are the one-query pass (the reference's own call shape)."""
    keys = torch.empty((8, k), dtype=torch.int64, device=dq.device)
    eng.set_option("tc", 0)   # this leg measures K1 itself (by default batches >= 5 take the tensor-core path)
    alg = n_rows * d * 4
    reps = max(20, min(steps, 200))
    passes = []
    for mq in (1, 4, 8):
        for _ in range(3):
            eng.query_device(dq.data_ptr(), mq, k, keys.data_ptr(), stream)
        torch.cuda.synchronize()
        s0 = eng.stats()
        eng.set_timing(True)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):   # back-to-back passes: each streams the whole 512 MB base again (4x the L2 capacity)
            eng.query_device(dq.data_ptr(), mq, k, keys.data_ptr(), stream)
        e1.record()
        torch.cuda.synchronize()
        s1 = eng.stats()
        eng.set_timing(False)
        hot_ms = (s1["hot_time_ns"] - s0["hot_time_ns"]) / 1e6 / max(1, s1["hot_timed_launches"] - s0["hot_timed_launches"])
        call_ms = e0.elapsed_time(e1) / reps
        passes.append({"queries": mq, "qtile": int(s1["last_qtile"]), "nsplit": int(s1["last_nsplit"]), "kernel_ms": hot_ms,
                       "call_ms": call_ms, "qps_call": mq / (call_ms * 1e-3), "achieved_GBs": alg / (hot_ms * 1e-3) / 1e9,
                       "frac": alg / (hot_ms * 1e-3) / 1e9 / peak, "frac_whole_call": alg / (call_ms * 1e-3) / 1e9 / peak})
    eng.set_option("tc", 1)
    p1 = passes[0]
    traffic, tsrc = ncu_traffic("scan_topk_f32_kernel<1") if (n_rows == 1_000_000 and d == 128) else (None, None)
    return {"bound": "hbm", "kernel": "scan_topk_f32_kernel (one pass of the fp32 scan + fused split merge, 1 query per pass)",
            "workload": "one pass over the %dx%d base for 1 / 4 / 8 queries; base is 4x L2, passes back to back, no flush" % (n_rows, d),
            "achieved": p1["achieved_GBs"], "peak": peak, "unit": "GB/s", "frac": p1["frac"], "peak_source": peak_src,
            "algorithmic_bytes_per_launch": int(alg), "kernel_ms_per_launch": p1["kernel_ms"],
            "traffic": traffic, "traffic_source": tsrc, "passes": passes}


_REAL_STDOUT = None


def protect_stdout():
    """the reference's engines print to std::cout (e.g. "Total query time", src/antitopo_engine.h destructor path):
    keep fd 1 clean for the ONE JSON line by pointing it at stderr for the whole run"""
    global _REAL_STDOUT
    if _REAL_STDOUT is None:
        sys.stdout.flush()
        _REAL_STDOUT = os.dup(1)
        os.dup2(2, 1)


def emit(obj):
    line = (json.dumps(obj) + "\n").encode()
    if _REAL_STDOUT is None:
        sys.stdout.write(line.decode())
        sys.stdout.flush()
    else:
        os.write(_REAL_STDOUT, line)


def tensor_roofline(st0, st1, m, n_rows_per_gpu, d, G, gq, workload, hot_ms):
    """roofline object of the K3 filter: issued flops 2*m*n*kd per GPU against the bf16 tensor peak; `useful` counts
    2*m*n*d only (the two-term split issues 3x that)"""
    seg = (d + 63) // 64 * 64
    mode = int(st1["tc_mode"])
    kd = 3 * seg if mode == 2 else ((d + 31) // 32 * 32 if mode == 4 else seg)
    m_gpu = (m + gq - 1) // gq
    flops = 2.0 * m_gpu * n_rows_per_gpu * kd
    useful = 2.0 * m_gpu * n_rows_per_gpu * d
    achieved = flops / (hot_ms * 1e-3) / 1e12
    tpeak, tpeak_sus, tsrc = tensor_peaks()
    if mode == 4:   # kind::tf32 runs at half the bf16 rate (K = 8 per instruction instead of 16): half the measured bf16 peaks
        tpeak, tpeak_sus, tsrc = tpeak / 2, tpeak_sus / 2, tsrc + " / 2 for tf32"
    qt = int(st1["last_qtile"])
    kname = "tc_filter_pair_kernel" if qt == 256 else "tc_filter2_kernel"
    traffic, trsrc = ncu_traffic(kname) if (workload == "c2" and G == 1 and m == 10000) else (None, None)
    return {"bound": "tensor", "kernel": kname + ((" (tcgen05.mma.cta_group::2 %s, fp32 accumulate in TMEM)" % ("tf32" if mode == 4 else "bf16")) if qt == 256 else
                                                  " (tcgen05.mma bf16, fp32 accumulate in TMEM)"),
            "achieved": achieved, "peak": tpeak, "unit": "TFLOP/s", "frac": achieved / tpeak, "frac_of_sustained_peak": achieved / tpeak_sus,
            "useful_flops_per_launch": useful, "achieved_useful": useful / (hot_ms * 1e-3) / 1e12, "frac_useful": useful / (hot_ms * 1e-3) / 1e12 / tpeak,
            "traffic": traffic, "traffic_source": trsrc, "traffic_unit": "DRAM bytes per launch (read + write)",
            "peak_source": tsrc, "algorithmic_flops_per_launch": flops, "kernel_ms_per_launch": hot_ms,
            "tc_mode": {1: "bf16-exact (integer rows)", 2: "two-term bf16 split (3x the flops)", 3: "one bf16 term per value (real-valued rows)",
                        4: "tf32 on the fp32 rows (kind::tf32)"}.get(mode, str(mode)),
            "kd": kd, "per_gpu": G > 1,
            "note": "flops = 2*m*n*kd of the candidate filter on ONE GPU's shard (the slowest GPU's kernel time when several run side by "
                    "side); useful = 2*m*n*d; survivors (about k per query) are re-ranked in fp32 by K4. peak = cuBLAS bf16 burst figure; "
                    "frac_of_sustained_peak uses the seconds-long figure"}


def timed_steps(torch, sync, fn, steps, join=None):
    """K steps bracketed by barrier + synchronize on both sides, CUDA events on the launching stream, max over ranks.
    join(): makes the launching stream wait for the side streams the steps were issued on (N > 1: two batches in flight)"""
    sync.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    e0.record()
    for _ in range(steps):
        fn()
    if join:
        join()
    e1.record()
    sync.barrier()
    t1 = time.perf_counter()
    return sync.max(e0.elapsed_time(e1)), t0, t1


class RankSync:
    """Barrier / max-over-ranks for the torchrun launch. NCCL is initialised and used ONCE (an all-reduce that proves the
    N-rank communicator), after that the ranks meet on a gloo group: an NCCL barrier is a kernel that spins on its GPU
    until every rank has arrived, and ranks > 0 reach each barrier long before rank 0 does -- their spinning kernels would
    sit on GPUs 1..N-1 (another process's context: time-sliced against rank 0's kernels there) through the whole timed
    region. A gloo barrier blocks on the CPU and leaves those GPUs to the multi-device handle."""

    def __init__(self, torch, dist, world, dev):
        self.torch, self.dist, self.world, self.dev, self.cpu = torch, dist, world, dev, None
        if world > 1:
            t = torch.ones(1, device=dev)
            dist.all_reduce(t)                      # NCCL communicator over all N ranks
            torch.cuda.synchronize()
            assert int(t.item()) == world
            self.cpu = dist.new_group(backend="gloo")

    def barrier(self):
        self.torch.cuda.synchronize()
        if self.world > 1:
            self.dist.barrier(group=self.cpu)

    def max(self, v):
        if self.world == 1:
            return float(v)
        t = self.torch.tensor([float(v)], dtype=self.torch.float64)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX, group=self.cpu)
        return float(t.item())


def extra_leg(expann_b200, torch, name, G, budget_left, steps=5):
    """bounded, non-headline legs of the other named configurations through the same handle type (one handle over G GPUs):
    c4 = Deep-shape 10M x 96 exact brute force; c5 = GIST-shape 1M x 960 inner product, ranged-q8 code scan + fp32
    rerank of 100 and the exact scan beside it. Returns a roofline-style dict with QPS."""
    from expann_b200 import datasets
    gen, variant, n, m, d, k, metric, desc = WORKLOADS[name]
    t0 = time.perf_counter()
    base, q = datasets.make(gen, n, m, d, variant=variant or "lowrank")
    gen_s = time.perf_counter() - t0
    devs = list(range(G)) if G > 1 else None
    out = {"workload": desc, "n": n, "m": m, "d": d, "k": k, "metric": metric, "gpus": G, "datagen_s": gen_s}
    dev0 = torch.device("cuda", 0)
    dq = torch.from_numpy(q).to(dev0)
    keys = torch.empty((m, k), dtype=torch.int64, device=dev0)
    stream = torch.cuda.current_stream(dev0).cuda_stream

    def run(eng, label):
        for _ in range(3):
            eng.query_device(dq.data_ptr(), m, k, keys.data_ptr(), stream)
        torch.cuda.synchronize()
        s0 = eng.stats()
        eng.set_timing(True)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            eng.query_device(dq.data_ptr(), m, k, keys.data_ptr(), stream)
        e1.record()
        torch.cuda.synchronize()
        s1 = eng.stats()
        eng.set_timing(False)
        ms = e0.elapsed_time(e1) / steps
        hot = (s1["hot_time_ns"] - s0["hot_time_ns"]) / 1e6 / max(1, s1["hot_timed_launches"] - s0["hot_timed_launches"])
        t1 = time.perf_counter()
        ids = eng.query_k_batch(q, k)[0]
        e2e_ms = (time.perf_counter() - t1) * 1e3
        row = {"qps": m / ms * 1e3, "ms_per_step": ms, "kernel_ms": hot, "e2e_qps_pageable": m / e2e_ms * 1e3,
               "tc_fallbacks": int(s1["tc_fallbacks"]), "row_groups": int(s1["row_groups"]) or 1, "query_groups": int(s1["query_groups"]) or 1}
        out[label] = row
        return ids, s0, s1, hot

    mt = 1 if metric == "ip" else 0
    if name == "c4":
        eng = expann_b200.BruteForceEngine(metric=mt, devices=devs)
        tb = time.perf_counter()
        eng.store_many_vectors(base)
        eng.build()
        out["build_s"] = time.perf_counter() - tb
        ids, s0, s1, hot = run(eng, "exact_scan")
        gr = out["exact_scan"]["row_groups"]
        out["exact_scan"]["roofline"] = tensor_roofline(s0, s1, m, (n + gr - 1) // gr, d, G, out["exact_scan"]["query_groups"], name, hot)
        eng.close()
    else:
        ex = expann_b200.BruteForceEngine(metric=mt, devices=devs)
        ex.store_many_vectors(base)
        ex.build()
        eids, s0, s1, hot = run(ex, "exact_scan")
        gr = out["exact_scan"]["row_groups"]
        out["exact_scan"]["roofline"] = tensor_roofline(s0, s1, m, (n + gr - 1) // gr, d, G, out["exact_scan"]["query_groups"], name, hot)
        ex.close()
        if time.perf_counter() - T_START < budget_left:
            qe = expann_b200.QuantScanEngine(quantizer=1, kprime=100, metric=mt, devices=devs)
            qe.store_many_vectors(base)
            qe.build()
            qids, s0, s1, hot = run(qe, "code_scan")
            dcp = (d + 127) // 128 * 128
            gr_q, gq_q = out["code_scan"]["row_groups"], out["code_scan"]["query_groups"]
            ops = 2.0 * ((m + gq_q - 1) // gq_q) * ((n + gr_q - 1) // gr_q) * dcp
            tpk = tensor_peaks()[0]
            out["code_scan"]["roofline"] = {"bound": "tensor", "kernel": "tc_filter_pair_kernel KIND = 1 (tcgen05.mma.kind::i8)", "achieved": ops / (hot * 1e-3) / 1e12,
                                            "peak": 2.0 * tpk, "unit": "Tera-op/s", "frac": ops / (hot * 1e-3) / 1e12 / (2.0 * tpk),
                                            "frac_of_nominal": ops / (hot * 1e-3) / 1e12 / 4500.0, "algorithmic_ops_per_launch": ops, "kernel_ms_per_launch": hot,
                                            "peak_source": "2 x measured bf16 burst peak; nominal dense int8 4500"}
            hit = sum(len(set(a.tolist()) & set(b.tolist())) for a, b in zip(qids, eids))
            out["code_scan"]["recall_at_k_vs_exact"] = hit / float(m * k)
            out["code_scan"]["path"] = "quantizer_ranged_q8 codes (src/quantizer.h:186-238), k' = 100 survivors re-ranked in fp32"
            qe.close()
    return out


def main():
    protect_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c2", choices=sorted(WORKLOADS))
    ap.add_argument("--ef", type=int, default=40, help="beam width of --engine graph")
    ap.add_argument("--engine", default="auto", choices=["auto", "bf", "quant", "graph"],
                    help="bf: exact brute force (K1/K3+K4); quant: quantizer compressed scan + fp32 rerank (K6+K4); auto: quant for c5")
    ap.add_argument("--qtile", type=int, default=0)
    ap.add_argument("--m", type=int, default=0, help="override the query count")
    ap.add_argument("--row-groups", type=int, default=0, help="N > 1: row shards (0 = automatic: shards of up to 2^20 rows; N = one per GPU); N / row_groups GPUs split the queries")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-extras", action="store_true", help="skip the extra legs (hbm scan, graph engine sweep, c4 / c5)")
    ap.add_argument("--extras-budget-s", type=float, default=300.0, help="extra legs start only while the run is younger than this")
    args = ap.parse_args()

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    G = max(args.gpus, world)
    gen, variant, n, m, d, k, metric, desc = WORKLOADS[args.workload]
    if args.engine == "graph":
        # graph engines need data with neighbourhood structure: the low-rank SIFT-shape variant of SURVEY 8(d) (C3), not iid
        if gen == "sift":
            variant = "lowrank"
        desc = desc.replace("(u8-valued iid)", "(u8-valued, low-rank)").replace("brute-force", "graph engine (one sub-graph per row shard),").replace("exact ", "")
    if args.m:
        m = args.m
    from expann_b200 import datasets

    if args.impl == "reference":
        if rank != 0:
            return 0
        base, q = datasets.make(gen, n, m, d, variant=variant or "lowrank")
        # each step = a bounded sample of the workload's queries on the full base
        budget = max(2.0, 120.0 / max(1, args.steps + args.warmup))
        vals = []
        info = None
        for i in range(args.warmup + args.steps):
            qps, cores, kind, sample = cpu_reference_qps(base, q, k, metric, budget_s=budget)
            if i >= args.warmup:
                vals.append(qps)
            info = (cores, kind, sample)
            if time.perf_counter() - T_START > 240 and len(vals) >= 3:   # bounded: the arm must end within minutes whatever K
                break
        v = float(np.mean(vals))
        out = {"impl": "reference", "metric": "QPS (exact brute-force top-%d)" % k, "value": v, "unit": "queries/s",
               "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1000.0 * m / v,
               "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
               "config": {"workload": desc, "n": n, "m": m, "d": d, "k": k, "metric": metric, "steps_measured": len(vals)},
               "cpu_baseline": {"value": v, "unit": "queries/s", "cores": info[0], "kind": info[1], "sample": info[2]},
               "e2e": {"value": v, "unit": "queries/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
        emit(out)
        return 0

    import torch
    import torch.distributed as dist
    import expann_b200

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a B200: no CUDA device visible (there is no CPU fallback)")
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    dev = torch.device("cuda", local_rank)

    sync = RankSync(torch, dist, world, dev)
    if rank != 0:
        # ONE process (rank 0) drives all N GPUs through one multi-device handle; this rank only joins the barriers and the
        # max-over-ranks reductions that bracket rank 0's timed regions, in the same order (headline, then e2e), on the CPU
        for _ in range(2):
            sync.barrier()
            sync.barrier()
            sync.max(0.0)
        sync.barrier()
        dist.destroy_process_group()
        return 0

    if G > torch.cuda.device_count():
        raise SystemExit(f"--gpus {G} but only {torch.cuda.device_count()} visible")
    devs = list(range(G)) if G > 1 else None
    base, q = datasets.make(gen, n, m, d, variant=variant or "lowrank")
    use_quant = args.engine == "quant" or (args.engine == "auto" and args.workload == "c5")
    use_graph = args.engine == "graph"
    mt = 1 if metric == "ip" else 0
    truth_ids = None
    graph_build_s = None
    if use_graph:
        # the graph engines shard like the scans: one GPU-built sub-graph per row shard, every query searches every shard
        # with the same beam width, K7 merges. Ground truth for recall@k: the exact scan of the same handle type (not timed).
        bf = expann_b200.BruteForceEngine(metric=mt, devices=devs)
        bf.store_many_vectors(base); bf.build()
        truth_ids = bf.query_k_batch(q, k)[0]
        bf.close()
        eng = expann_b200.GraphEngine(metric=mt, devices=devs)
        eng.store_many_vectors(base)
        t_b = time.perf_counter()
        eng.build_graph(M=16, M0=32, K=64)
        graph_build_s = time.perf_counter() - t_b
        eng.set_ef_search(args.ef)
    elif use_quant:   # BASELINE config 5: quantizer_ranged_q8 codes (src/quantizer.h:186-238) scanned, k' = 100 survivors re-ranked in fp32
        eng = expann_b200.QuantScanEngine(quantizer=1, kprime=100, metric=mt, qtile=args.qtile, devices=devs)
    else:
        eng = expann_b200.BruteForceEngine(metric=mt, qtile=args.qtile, devices=devs, row_groups=args.row_groups)
    if not use_graph:
        eng.store_many_vectors(base)
    eng.build()

    stream = torch.cuda.current_stream().cuda_stream
    dq = torch.from_numpy(q).to(dev)
    hq = torch.from_numpy(q).pin_memory()
    final_keys = torch.empty((m, k), dtype=torch.int64, device=dev)
    ids_dev = torch.empty((m, k), dtype=torch.int64, device=dev)
    hq_np = hq.numpy()   # a view of the page-locked tensor
    out_pinned = (torch.empty((m, k), dtype=torch.int64).pin_memory().numpy().view(np.uint64),
                  torch.empty((m, k), dtype=torch.float32).pin_memory().numpy(),
                  torch.empty((m,), dtype=torch.int64).pin_memory().numpy().view(np.uint64))
    out_pageable = (np.empty((m, k), np.uint64), np.empty((m, k), np.float32), np.empty((m,), np.uint64))
    q_pageable = np.array(q, copy=True)

    # Consecutive steps are issued on two alternating streams with their own result buffers -- two batches in flight, the way
    # a serving loop double-buffers: a call's results are ordered on ITS stream, so on one stream step s+1 could not start on
    # any GPU before the slowest GPU had finished step s, and one batch's conversion / select / re-rank could not run beside
    # the other's filter. The brute-force handle keeps two sets of work buffers for this (csrc/api.cu scratch_for_stream; a
    # multi-device handle alternates two streams per GPU); the graph and quantised engines have one set and take one stream.
    alt = None
    if G > 1 or (not use_graph and not use_quant):
        alt = {"streams": [torch.cuda.Stream(device=dev), torch.cuda.Stream(device=dev)],
               "keys": [final_keys, torch.empty_like(final_keys)], "i": 0}
        torch.cuda.synchronize()

    def step_device():
        if alt is None:
            eng.query_device(dq.data_ptr(), m, k, final_keys.data_ptr(), stream)
        else:
            i = alt["i"] = alt["i"] ^ 1
            eng.query_device(dq.data_ptr(), m, k, alt["keys"][i].data_ptr(), alt["streams"][i].cuda_stream)

    def join_streams():
        if alt is not None:
            for s_ in alt["streams"]:
                torch.cuda.current_stream().wait_stream(s_)

    def step_e2e():
        # the C-ABI host call: H2D of the pinned queries, scan (all GPUs), merge, decode, D2H into pinned result buffers
        return eng.query_k_batch(hq_np, k, out=out_pinned)[0]

    def step_e2e_pageable():
        return eng.query_k_batch(q_pageable, k, out=out_pageable)[0]

    for _ in range(args.warmup):
        step_device()
    torch.cuda.synchronize()
    st0 = eng.stats()
    eng.set_timing(True)
    sampler = ClockSampler(local_rank)
    sampler.start()
    ms, tw0, tw1 = timed_steps(torch, sync, step_device, args.steps, join=join_streams)
    st1 = eng.stats()
    eng.set_timing(False)
    # sustained region: the same step for >= 0.5 s, so that the clock sampler sees the loaded clock and the seconds-long tensor
    # peak is the right denominator; the headline value stays the K-step region above
    sus = None
    if ms < 400.0 and not args.no_extras:
        reps = int(min(2000, max(args.steps, 500.0 / max(ms / args.steps, 1e-3))))
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(reps):
            step_device()
        join_streams()
        e1.record()
        torch.cuda.synchronize()
        tw1 = time.perf_counter()
        sus = {"steps": reps, "ms_per_step": e0.elapsed_time(e1) / reps, "qps": m * reps / (e0.elapsed_time(e1) * 1e-3)}
    clocks = sampler.stop(tw0, tw1)
    # N > 1: with two batches in flight a GPU works on two calls side by side (the handle's second stream and second set of
    # work buffers), so the event pairs around consecutive filter launches overlap and over-count. The kernel time of the
    # roofline is therefore taken over a region of the same step with ONE batch in flight, right behind the timed one.
    hot_region = None
    if alt is not None:
        torch.cuda.synchronize()
        hs0 = eng.stats()
        eng.set_timing(True)
        nrep_hot = max(10, min(args.steps, 50))
        for _ in range(nrep_hot):
            eng.query_device(dq.data_ptr(), m, k, final_keys.data_ptr(), stream)
        torch.cuda.synchronize()
        hs1 = eng.stats()
        eng.set_timing(False)
        hot_region = (hs0, hs1, nrep_hot)
    # e2e leg (host buffers): its own warm-up, then K timed steps by wall clock bracketed by barriers
    for _ in range(max(1, min(args.warmup, 2))):
        step_e2e()
    sync.barrier()
    p0 = eng.stats()["peer_bytes"]
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step_e2e()
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    peer_per_step = (eng.stats()["peer_bytes"] - p0) / max(1, args.steps)
    sync.barrier()
    e2e_s = sync.max(e2e_s)
    step_e2e_pageable()
    nrep = max(3, min(args.steps, 20))
    t0 = time.perf_counter()
    for _ in range(nrep):
        step_e2e_pageable()
    e2e_pageable_s = (time.perf_counter() - t0) / nrep

    # correctness (not timed): keys decode to the ids of the host call; N > 1: ids of 256 queries equal a ONE-GPU scan of the whole base
    torch.cuda.synchronize()
    eng.query_device(dq.data_ptr(), m, k, final_keys.data_ptr(), stream)
    eng.decode_device(final_keys.data_ptr(), m, k, ids_dev.data_ptr(), 0, 0, stream)
    torch.cuda.synchronize()
    got = ids_dev.cpu().numpy().view(np.uint64)
    check = {"first_ids": [int(x) for x in got[0]], "device_path_equals_host_call": bool(np.array_equal(got, step_e2e()))}
    if G > 1 and not use_graph:
        msub = min(m, 256)
        one = (expann_b200.QuantScanEngine(quantizer=1, kprime=100, metric=mt, device=0) if use_quant
               else expann_b200.BruteForceEngine(metric=mt, device=0))
        one.store_many_vectors(base); one.build()
        w = one.query_k_batch(q[:msub], k)
        one.close()
        g = eng.query_k_batch(q[:msub], k)
        same = bool(np.array_equal(w[0], g[0]) and np.array_equal(w[1], g[1]))
        check["ids_equal_one_gpu_scan"] = {"queries": msub, "equal": same}
        assert same, "multi-GPU result differs from the one-GPU scan of the same rows"

    peak, peak_src = peaks()
    launches = sum(st1[x] - st0[x] for x in ("scan_launches", "merge_launches", "graph_launches", "aux_launches", "tc_launches"))
    hs0, hs1 = (hot_region[0], hot_region[1]) if hot_region else (st0, st1)
    hot_ms = (hs1["hot_time_ns"] - hs0["hot_time_ns"]) / 1e6 / max(1, hs1["hot_timed_launches"] - hs0["hot_timed_launches"])
    qt = st1["last_qtile"]
    gr, gq = int(st1["row_groups"]) or 1, int(st1["query_groups"]) or 1
    rows_gpu = (n + gr - 1) // gr
    used_tc = st1["tc_launches"] > st0["tc_launches"] and st1["tc_fallbacks"] == st0["tc_fallbacks"]
    if use_graph:
        dcq = (st1["distcomps"] - st0["distcomps"]) / float(args.steps * m)   # all shards together
        alg_bytes = dcq * d * 4 * m / max(1, gr * gq)
        achieved = alg_bytes / (hot_ms * 1e-3) / 1e9
        roof = {"bound": "hbm", "kernel": "graph_search_kernel", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": None, "traffic_source": None, "peak_source": peak_src, "algorithmic_bytes_per_launch": int(alg_bytes), "kernel_ms_per_launch": hot_ms,
                "distcomps_per_query_all_gpus": dcq,
                "note": "algorithmic bytes = distance evaluations x d x 4 per GPU (random row gathers); latency-bound"}
    elif used_tc and use_quant:
        # K6-TC: tcgen05.mma.kind::i8 over the 8-bit codes; ops = 2 * m * n_shard * dc_pad (code row stride: d rounded up to 128)
        dcp = (d + 127) // 128 * 128
        m_gpu = (m + gq - 1) // gq
        ops = 2.0 * m_gpu * rows_gpu * dcp
        tpeak, tpeak_sus, tsrc = tensor_peaks()
        achieved = ops / (hot_ms * 1e-3) / 1e12
        code_gbs = rows_gpu * dcp / (hot_ms * 1e-3) / 1e9
        roof = {"bound": "tensor", "kernel": "tc_filter_pair_kernel<.., KIND = 1> (tcgen05.mma.cta_group::2.kind::i8, u8 x u8 -> s32 in TMEM)",
                "achieved": achieved, "peak": 2.0 * tpeak, "unit": "Tera-op/s", "frac": achieved / (2.0 * tpeak),
                "peak_source": "2 x the measured bf16 burst peak (MEASURED_PEAKS.json has no int8 figure; an int8 instruction moves twice the K of a bf16 one "
                               "in the same time); nominal dense int8 = 4500 Tera-op/s (B200_PROFILING.md): frac_of_nominal beside it",
                "frac_of_nominal": achieved / 4500.0, "algorithmic_ops_per_launch": ops, "useful_ops_per_launch": 2.0 * m_gpu * rows_gpu * d,
                "kernel_ms_per_launch": hot_ms, "code_bytes_per_launch": int(rows_gpu * dcp), "code_GBs": code_gbs, "code_frac_of_hbm": code_gbs / peak,
                "traffic": None, "traffic_source": None, "per_gpu": G > 1,
                "note": "the quantizer's compressed scan as a tensor-core candidate filter over the code matrix; the k' = 100 best codes per query "
                        "(bit-identical to the SIMT scan's) are re-ranked in fp32 by K4"}
    elif used_tc:
        roof = tensor_roofline(st0, st1, m, rows_gpu, d, G, gq, args.workload, hot_ms)
    else:
        passes = (m + qt - 1) // max(1, qt)
        alg_bytes = passes * rows_gpu * d * (1 if use_quant else 4)
        achieved = alg_bytes / (hot_ms * 1e-3) / 1e9
        roof = {"bound": "hbm", "kernel": "scan_topk_u8_kernel" if use_quant else "scan_topk_f32_kernel", "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": achieved / peak, "traffic": None, "traffic_source": None, "peak_source": peak_src,
                "algorithmic_bytes_per_launch": int(alg_bytes), "kernel_ms_per_launch": hot_ms,
                "note": "algorithmic bytes = ceil(m/qtile) passes x shard bytes (SURVEY 8(d)); CTAs of one "
                        "wave share rows through L2, so DRAM traffic is lower than this figure"}
    if sus:
        roof["sustained"] = sus
    if hot_region:
        roof["kernel_time_region"] = ("%d steps of the same call with ONE batch in flight, right behind the timed region (with two in flight "
                                      "a GPU runs two calls side by side and the event pairs of consecutive filter launches overlap)" % hot_region[2])
    ms_step = ms / args.steps
    out = {"metric": ("QPS (graph engine, ef=%d, top-%d)" % (args.ef, k)) if use_graph else "QPS (exact brute-force top-%d)" % k,
           "value": m * args.steps / (ms * 1e-3), "unit": "queries/s",
           "n_gpus": G, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step,
           "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
           "config": {"workload": desc, "n": n, "m": m, "d": d, "k": k, "metric": metric, "qtile": int(qt),
                      "decomposition": ("automatic: %d row shard(s) of up to 2^20 rows x %d query group(s)" % (gr, gq)) if (G > 1 and not args.row_groups) else None,
                      "path": "K5 graph beam search (one GPU-built graph per shard, M0=32)" if use_graph else
                              ("K6-TC 8-bit code filter (tcgen05 kind::i8) + K4 fp32 rerank of 100 survivors" if used_tc else "K6 8-bit code scan (dp4a) + K4 fp32 rerank of 100 survivors") if use_quant else
                              ("K3 tcgen05 filter + K4 fp32 rerank" if used_tc else "K1 fp32 SIMT scan"),
                      "nsplit": int(st1["last_nsplit"]), "rows_per_gpu": rows_gpu, "row_groups": gr, "query_groups": gq,
                      "l2": "base shard (%.0f MB) %s L2 (126 MB); no explicit flush" % (rows_gpu * d * 4 / 1e6, "larger than" if rows_gpu * d * 4 > 126e6 else "fits in"),
                      "step_minus_kernel_ms": ms_step - hot_ms, "batches_in_flight": 2 if alt is not None else 1,
                      "parallelism": ("one process, one multi-device handle over %d GPUs: %d row shards x %d query groups, queries fanned out over "
                                      "NVLink, per-GPU [m][k] keys stored by each GPU's last kernel into GPU 0's gather buffer, K7 merge "
                                      "(no NCCL on the data path; torchrun ranks > 0 only join the CPU-side barriers)" % (G, gr, gq)) if G > 1 else "single GPU"},
           "e2e": {"value": m * args.steps / e2e_s, "unit": "queries/s", "h2d_bytes_per_step": int(m * d * 4),
                   "d2h_bytes_per_step": int(m * k * 8 + m * k * 4 + m * 8),
                   "api": "expann_gpu_query (page-locked caller buffers)", "pageable_value": m / e2e_pageable_s,
                   "pageable_api": "expann_gpu_query (pageable numpy arrays, staged through the handle's page-locked buffers)",
                   "peer_bytes_per_step": int(peer_per_step)},
           "gpu_launches": int(launches),
           "roofline": roof,
           "clocks": clocks}
    if not args.no_cpu and G == 1 and not use_graph:
        qps, cores, kind, sample = cpu_reference_qps(base, q, k, metric)
        out["cpu_baseline"] = {"value": qps, "unit": "queries/s", "cores": cores, "kind": kind, "sample": sample}
    out["check"] = check
    if use_graph:
        hit = sum(len(set(a.tolist()) & set(b.tolist())) for a, b in zip(got, truth_ids))
        out["check"]["recall_at_k_vs_exact"] = hit / float(m * k)
        out["config"]["ef"] = args.ef
        out["config"]["graph_build_s"] = graph_build_s
    if use_quant and G == 1:
        # quality of the compressed scan: recall@10 against the exact fp32 scan of the same library (not timed)
        try:
            ex = expann_b200.BruteForceEngine(metric=mt, device=local_rank)
            ex.store_many_vectors(base)
            ex.build()
            msub = min(m, 1000)
            eids = ex.query_k_batch(q[:msub], k)[0]
            ex.close()
            qids = eng.query_k_batch(q[:msub], k)[0]
            hit = sum(len(set(a.tolist()) & set(b.tolist())) for a, b in zip(qids, eids))
            out["check"]["recall_at_k_vs_exact"] = hit / float(msub * k)
        except Exception as exn:  # noqa: BLE001
            out["check"]["recall_error"] = repr(exn)[:200]
    extras = not args.no_extras and not use_quant and not use_graph and args.workload == "c2"

    def young():
        return time.perf_counter() - T_START < args.extras_budget_s

    if extras and G > 1 and gr != G and young():
        # the decomposition the north star names beside the automatic one of the headline: ONE row shard per GPU, all queries
        # on every GPU, K7 merge of G lists (not the headline: the same numbers for the other decomposition)
        try:
            e2 = expann_b200.BruteForceEngine(metric=mt, qtile=args.qtile, devices=devs, row_groups=G)
            e2.store_many_vectors(base)
            e2.build()
            for _ in range(args.warmup):
                e2.query_device(dq.data_ptr(), m, k, final_keys.data_ptr(), stream)
            torch.cuda.synchronize()
            s0 = e2.stats()
            e2.set_timing(True)
            ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            ev0.record()
            for i in range(args.steps):   # two batches in flight, like the headline
                e2.query_device(dq.data_ptr(), m, k, alt["keys"][i & 1].data_ptr(), alt["streams"][i & 1].cuda_stream)
            join_streams()
            ev1.record()
            torch.cuda.synchronize()
            ms2 = ev0.elapsed_time(ev1) / args.steps
            s0 = e2.stats()              # kernel time: one batch in flight (see the headline's kernel_time_region)
            for _ in range(max(10, min(args.steps, 50))):
                e2.query_device(dq.data_ptr(), m, k, final_keys.data_ptr(), stream)
            torch.cuda.synchronize()
            s1 = e2.stats()
            e2.set_timing(False)
            hot2 = (s1["hot_time_ns"] - s0["hot_time_ns"]) / 1e6 / max(1, s1["hot_timed_launches"] - s0["hot_timed_launches"])
            e2.query_k_batch(hq_np, k, out=out_pinned)
            t0 = time.perf_counter()
            for _ in range(args.steps):
                e2.query_k_batch(hq_np, k, out=out_pinned)
            e2e2 = (time.perf_counter() - t0) / args.steps
            rs = tensor_roofline(s0, s1, m, (n + G - 1) // G, d, G, 1, args.workload, hot2)
            rs.update({"decomposition": "%d row shards x 1 query group (one shard per GPU, K7 merge of %d lists)" % (G, G),
                       "qps": m / ms2 * 1e3, "ms_per_step": ms2, "step_minus_kernel_ms": ms2 - hot2, "e2e_qps": m / e2e2,
                       "ids_equal_headline": bool(np.array_equal(e2.query_k_batch(q[:256], k)[0], got[:256]))})
            roof["row_sharded"] = rs
            e2.close()
            del e2
        except Exception as ex:  # noqa: BLE001
            roof["row_sharded"] = {"error": repr(ex)[:300]}
    if extras and G == 1:
        try:
            roof["hbm_scan"] = scan_hbm_leg(eng, torch, dq, k, n, d, stream, peak, peak_src, args.steps)
        except Exception as ex:  # noqa: BLE001
            roof["hbm_scan"] = {"error": repr(ex)[:300]}
    eng.close()
    del eng
    if extras and G == 1 and young():
        # BASELINE config 3 (graph engine, QPS at recall@10 >= 0.95) measured in the same run
        try:
            from tools import bench_extras
            torch.cuda.empty_cache()
            g3 = bench_extras.graph_c3(local_rank, max(2, min(args.steps, 20)), args.warmup, cpu=not args.no_cpu, peak_gbs=peak)
            if "roofline" in g3:   # DRAM bytes per launch of the ef = 40 search from the committed ncu capture
                g3["roofline"]["traffic"], g3["roofline"]["traffic_source"] = ncu_traffic("graph_search_blocks_kernel")
            roof["graph"] = g3
            if g3.get("at_recall_0.95"):
                out["graph_qps_at_recall_0_95"] = g3["at_recall_0.95"]["qps"]
        except Exception as ex:  # noqa: BLE001
            roof["graph"] = {"error": repr(ex)[:300]}
    # the other named configurations, bounded and not timed as the headline: C4 wherever it is sharded (N > 1), C5 at 1 and 8 GPUs
    del base, q
    for leg, want in (("c4", G > 1), ("c5", G in (1, 8))):
        if extras and want and young():
            try:
                torch.cuda.empty_cache()
                roof[leg] = extra_leg(expann_b200, torch, leg, G, args.extras_budget_s)
            except Exception as ex:  # noqa: BLE001
                roof[leg] = {"error": repr(ex)[:300]}
    out["run_s"] = time.perf_counter() - T_START
    emit(out)
    if world > 1:
        sync.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
