#!/usr/bin/env python
"""bench.py — headline benchmark of the expANN query hot path on B200 (contract: see DESIGN.md §Measurement).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload c2|c1|c4|c5]

A "step" is one pass of the hot path over one batch of m synthetic queries against the HBM-resident base set.
  value : QPS, device-resident queries in, merged 64-bit (dist,id) keys out (CUDA events, max over ranks)
  e2e   : QPS through the reference-facing host-buffer call (expann_gpu_query at N=1), H2D of the queries and
          D2H of ids inside the timed region
  roofline : algorithmic bytes of the dominant kernel / its CUDA-event time (events recorded inside the library
          around every scan launch on the launching stream)
  cpu_baseline : the UNMODIFIED reference brute_force_engine (oracle/_ref) — or the C restatement when that
          build is not usable on this host — timed on a bounded query sample with all host threads
N > 1 (torchrun): the base set is sharded by contiguous row ranges (strong scaling: the same 1M-row base over
N GPUs), every rank scans its shard for all queries, the per-rank [m][k] key lists are all-gathered over NCCL
and merged by K7 on every rank.
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


def ncu_traffic(key):
    """DRAM bytes (read + write) per launch of a kernel from the committed ncu --set full capture of the same kernel on
    the C2 shape (profiles/r1_ncu_summary.json); None when there is no capture"""
    try:
        j = json.load(open(os.path.join(ROOT, "profiles", "r1_ncu_summary.json")))
        for name, d in j.items():
            if name.startswith(key):
                unit = {"Mbyte": 1e6, "Gbyte": 1e9, "Kbyte": 1e3, "byte": 1.0}
                r, w = d["dram__bytes_read.sum"], d["dram__bytes_write.sum"]
                return int(float(r[0]) * unit[r[1]] + float(w[0]) * unit[w[1]])
    except Exception:  # noqa: BLE001
        pass
    return None


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
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={q}", "--format=csv,noheader,nounits", "-lms", "100",
                                          "-i", str(self.index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:  # noqa: BLE001
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:  # noqa: BLE001
            self.proc.kill()
        sm = [float(r[0]) for r in self.rows if r and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in self.rows if len(r) >= 7 for i in range(4) if r[3 + i].lower().startswith("active")})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


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
            t0 = time.perf_counter()
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
    again. The headline roofline object is the one-query pass (the reference's own call shape); `passes` lists all three."""
    keys = torch.empty((8, k), dtype=torch.int64, device=dq.device)
    eng.set_option("tc", 0)   # this leg measures K1 itself (by default batches >= 5 take the tensor-core path)
    alg = n_rows * d * 4
    reps = max(20, steps)
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
    return {"workload": "one pass of the fp32 scan (+ fused split merge) over the %dx%d base for 1 / 4 / 8 queries; "
                        "base is 4x L2, passes back to back, no flush" % (n_rows, d),
            "passes": passes,
            "roofline": {"bound": "hbm", "kernel": "scan_topk_f32_kernel (1 query per pass)", "achieved": p1["achieved_GBs"],
                         "peak": peak, "unit": "GB/s", "frac": p1["frac"],
                         "traffic": ncu_traffic("scan_topk_f32_kernel<1") if (n_rows == 1_000_000 and d == 128) else None,
                         "peak_source": peak_src,
                         "algorithmic_bytes_per_launch": int(alg), "kernel_ms_per_launch": p1["kernel_ms"]}}


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


def main():
    protect_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c2", choices=sorted(WORKLOADS))
    ap.add_argument("--ef", type=int, default=40, help="beam width of --engine graph")
    ap.add_argument("--engine", default="auto", choices=["auto", "bf", "quant", "graph"],
                    help="bf: exact brute force (K1/K3+K4); quant: quantizer compressed scan + fp32 rerank (K6+K4); auto: quant for c5")
    ap.add_argument("--qtile", type=int, default=0)
    ap.add_argument("--m", type=int, default=0, help="override the query count")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-extras", action="store_true", help="skip the extra legs (graph engine sweep, BASELINE config 3)")
    args = ap.parse_args()

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    gen, variant, n, m, d, k, metric, desc = WORKLOADS[args.workload]
    if args.engine == "graph":
        # graph engines need data with neighbourhood structure: the low-rank SIFT-shape variant of SURVEY 8(d) (C3), not iid
        if gen == "sift":
            variant = "lowrank"
        desc = desc.replace("(u8-valued iid)", "(u8-valued, low-rank)").replace("brute-force", "graph engine (sharded),").replace("exact ", "")
    if args.m:
        m = args.m
    from expann_b200 import datasets

    if args.impl == "reference":
        if rank != 0:
            return 0
        base, q = datasets.make(gen, n, m, d, variant=variant or "lowrank")
        # each step = a bounded sample of the workload's queries on the full base
        budget = max(4.0, 120.0 / max(1, args.steps + args.warmup))
        vals = []
        info = None
        for i in range(args.warmup + args.steps):
            qps, cores, kind, sample = cpu_reference_qps(base, q, k, metric, budget_s=budget)
            if i >= args.warmup:
                vals.append(qps)
            info = (cores, kind, sample)
        v = float(np.mean(vals))
        out = {"impl": "reference", "metric": "QPS (exact brute-force top-%d)" % k, "value": v, "unit": "queries/s",
               "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1000.0 * m / v,
               "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
               "config": {"workload": desc, "n": n, "m": m, "d": d, "k": k, "metric": metric},
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

    base, q = datasets.make(gen, n, m, d, variant=variant or "lowrank")
    from expann_b200 import sharding
    lo, hi = sharding.shard_range(n, world, rank)
    use_quant = args.engine == "quant" or (args.engine == "auto" and args.workload == "c5")
    use_graph = args.engine == "graph"
    truth_ids = None
    if use_graph:
        # BASELINE config 4's graph leg: every rank builds a graph over ITS rows (GPU build, reference prune rule) and
        # searches it for all queries; per-GPU top-k lists meet through the same NCCL gather + K7 merge as the scans.
        # Ground truth for recall@k: the sharded exact scan of the same rows (not timed).
        bf = expann_b200.BruteForceEngine(metric=1 if metric == "ip" else 0, device=local_rank, id_offset=lo)
        bf.store_many_vectors(base[lo:hi]); bf.build()
        tq = torch.from_numpy(q).to(dev)
        tk_local = torch.empty((m, k), dtype=torch.int64, device=dev)
        bf.query_device(tq.data_ptr(), m, k, tk_local.data_ptr(), torch.cuda.current_stream().cuda_stream)
        if world > 1:
            g = sharding.gather_keys(tk_local, world)
            tk_final = torch.empty_like(tk_local)
            bf.merge_device(g.data_ptr(), m, world, k, k, m * k, k, tk_final.data_ptr(), torch.cuda.current_stream().cuda_stream)
        else:
            tk_final = tk_local
        tids = torch.empty((m, k), dtype=torch.int64, device=dev)
        bf.decode_device(tk_final.data_ptr(), m, k, tids.data_ptr(), 0, 0, torch.cuda.current_stream().cuda_stream)
        torch.cuda.synchronize()
        truth_ids = tids.cpu().numpy()
        bf.close()
        del bf, tq, tk_local, tk_final, tids
        eng = expann_b200.GraphEngine(metric=1 if metric == "ip" else 0, device=local_rank, id_offset=lo)
        eng.store_many_vectors(base[lo:hi])
        t_b = time.perf_counter()
        eng.build_graph(M=16, M0=32, K=64)
        graph_build_s = time.perf_counter() - t_b
        eng.set_ef_search(args.ef)
    elif use_quant:   # BASELINE config 5: quantizer_ranged_q8 codes (src/quantizer.h:186-238) scanned with K6, k' = 100 survivors re-ranked in fp32
        eng = expann_b200.QuantScanEngine(quantizer=1, kprime=100, metric=1 if metric == "ip" else 0, device=local_rank,
                                          qtile=args.qtile, id_offset=lo)
    else:
        eng = expann_b200.BruteForceEngine(metric=1 if metric == "ip" else 0, device=local_rank, qtile=args.qtile, id_offset=lo)
    if not use_graph:
        eng.store_many_vectors(base[lo:hi])
    eng.build()

    stream = torch.cuda.current_stream().cuda_stream
    dq = torch.from_numpy(q).to(dev)
    hq = torch.from_numpy(q).pin_memory()
    local_keys = torch.empty((m, k), dtype=torch.int64, device=dev)
    final_keys = torch.empty((m, k), dtype=torch.int64, device=dev)
    ids_dev = torch.empty((m, k), dtype=torch.int64, device=dev)
    ids_host = torch.empty((m, k), dtype=torch.int64).pin_memory()
    dq_e2e = torch.empty_like(dq)
    hq_np = hq.numpy()   # a view of the page-locked tensor
    out_pinned = (torch.empty((m, k), dtype=torch.int64).pin_memory().numpy().view(np.uint64),
                  torch.empty((m, k), dtype=torch.float32).pin_memory().numpy(),
                  torch.empty((m,), dtype=torch.int64).pin_memory().numpy().view(np.uint64))

    def step_device(qptr):
        if world == 1:
            eng.query_device(qptr, m, k, final_keys.data_ptr(), stream)
        else:
            eng.query_device(qptr, m, k, local_keys.data_ptr(), stream)
            gathered = sharding.gather_keys(local_keys, world)   # NCCL all-gather -> [world][m][k]
            eng.merge_device(gathered.data_ptr(), m, world, k, k, m * k, k, final_keys.data_ptr(), stream)

    def step_e2e():
        if world == 1:
            # the C-ABI host call: H2D of the pinned queries, scan, decode, D2H into pinned result buffers
            return eng.query_k_batch(hq_np, k, out=out_pinned)[0]
        dq_e2e.copy_(hq, non_blocking=True)
        step_device(dq_e2e.data_ptr())
        eng.decode_device(final_keys.data_ptr(), m, k, ids_dev.data_ptr(), 0, 0, stream)
        ids_host.copy_(ids_dev, non_blocking=True)
        torch.cuda.current_stream().synchronize()
        return ids_host

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    for _ in range(args.warmup):
        step_device(dq.data_ptr())
    barrier()
    st0 = eng.stats()
    eng.set_timing(True)
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    ms = timed(lambda: step_device(dq.data_ptr()), args.steps)
    clocks = sampler.stop() if rank == 0 else None
    st1 = eng.stats()
    eng.set_timing(False)
    # e2e leg (host buffers): its own warm-up, then K timed steps by wall clock bracketed by barriers
    for _ in range(max(1, min(args.warmup, 2))):
        step_e2e()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step_e2e()
    barrier()
    e2e_s = torch.tensor([time.perf_counter() - t0], device=dev)
    if world > 1:
        dist.all_reduce(e2e_s, op=dist.ReduceOp.MAX)
    e2e_s = float(e2e_s.item())

    # correctness spot check on rank 0 (not timed): the keys decode to sorted unique ids
    eng.decode_device(final_keys.data_ptr(), m, k, ids_dev.data_ptr(), 0, 0, stream)
    torch.cuda.synchronize()
    got = ids_dev[:4].cpu().numpy()

    if rank == 0:
        peak, peak_src = peaks()
        launches = sum(st1[x] - st0[x] for x in ("scan_launches", "merge_launches", "graph_launches", "aux_launches", "tc_launches"))
        hot_ms = (st1["hot_time_ns"] - st0["hot_time_ns"]) / 1e6 / max(1, st1["hot_timed_launches"] - st0["hot_timed_launches"])
        qt = st1["last_qtile"]
        used_tc = st1["tc_launches"] > st0["tc_launches"] and st1["tc_fallbacks"] == st0["tc_fallbacks"]
        if use_graph:
            dcq = (st1["distcomps"] - st0["distcomps"]) / float(args.steps * m)   # this rank's shard
            alg_bytes = dcq * d * 4 * m
            achieved = alg_bytes / (hot_ms * 1e-3) / 1e9
            roof = {"bound": "hbm", "kernel": "graph_search_kernel", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                    "traffic": None, "peak_source": peak_src, "algorithmic_bytes_per_launch": int(alg_bytes), "kernel_ms_per_launch": hot_ms,
                    "distcomps_per_query_per_gpu": dcq,
                    "note": "algorithmic bytes = distance evaluations x d x 4 on rank 0's shard (random row gathers); latency-bound"}
        elif used_tc:
            # K3: bf16 tcgen05 GEMM of a'(q,x) over the whole shard; flops = 2 * m * n_shard * kd (kd = operand row length:
            # d rounded up to 64 in the bf16-exact mode, 3x that in the two-term split mode)
            seg = (d + 63) // 64 * 64
            kd = seg if st1["tc_mode"] == 1 else 3 * seg
            flops = 2.0 * m * (hi - lo) * kd
            achieved = flops / (hot_ms * 1e-3) / 1e12
            tpeak, tpeak_sus, tsrc = tensor_peaks()
            kname = "tc_filter_pair_kernel" if qt == 256 else "tc_filter2_kernel"
            roof = {"bound": "tensor", "kernel": kname + (" (tcgen05.mma.cta_group::2 bf16, fp32 accumulate in TMEM)" if qt == 256 else
                                                          " (tcgen05.mma bf16, fp32 accumulate in TMEM)"), "achieved": achieved,
                    "peak": tpeak, "unit": "TFLOP/s", "frac": achieved / tpeak, "frac_of_sustained_peak": achieved / tpeak_sus,
                    "traffic": ncu_traffic(kname) if (args.workload == "c2" and world == 1 and m == 10000) else None,
                    "traffic_unit": "DRAM bytes per launch (ncu --set full, profiles/r1_ncu_summary.json)",
                    "peak_source": tsrc, "algorithmic_flops_per_launch": flops, "kernel_ms_per_launch": hot_ms,
                    "tc_mode": "bf16-exact" if st1["tc_mode"] == 1 else "two-term bf16 split", "kd": kd,
                    "note": "flops = 2*m*n*kd of the candidate filter; survivors (about k per query) are re-ranked in fp32 by K4. "
                            "peak = cuBLAS bf16 burst figure; frac_of_sustained_peak uses the seconds-long figure"}
        else:
            passes = (m + qt - 1) // qt
            alg_bytes = passes * (hi - lo) * d * (1 if use_quant else 4)
            achieved = alg_bytes / (hot_ms * 1e-3) / 1e9
            roof = {"bound": "hbm", "kernel": "scan_topk_u8_kernel" if use_quant else "scan_topk_f32_kernel", "achieved": achieved, "peak": peak, "unit": "GB/s",
                    "frac": achieved / peak, "traffic": None, "peak_source": peak_src,
                    "algorithmic_bytes_per_launch": int(alg_bytes), "kernel_ms_per_launch": hot_ms,
                    "note": "algorithmic bytes = ceil(m/qtile) passes x shard bytes (SURVEY 8(d)); CTAs of one "
                            "wave share rows through L2, so DRAM traffic is lower than this figure"}
        out = {"metric": ("QPS (graph engine, ef=%d, top-%d)" % (args.ef, k)) if use_graph else "QPS (exact brute-force top-%d)" % k, "value": m * args.steps / (ms * 1e-3), "unit": "queries/s",
               "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps,
               "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
               "config": {"workload": desc, "n": n, "m": m, "d": d, "k": k, "metric": metric, "qtile": int(qt),
                          "path": "K5 graph beam search (one GPU-built graph per shard, M0=32)" if use_graph else
                                  "K6 8-bit code scan (dp4a) + K4 fp32 rerank of 100 survivors" if use_quant else
                                  ("K3 tcgen05 filter + K4 fp32 rerank" if used_tc else "K1 fp32 SIMT scan"),
                          "nsplit": int(st1["last_nsplit"]), "rows_per_gpu": hi - lo,
                          "l2": "base shard (%.0f MB) %s L2 (126 MB); no explicit flush" % ((hi - lo) * d * 4 / 1e6, "larger than" if (hi - lo) * d * 4 > 126e6 else "fits in"),
                          "parallelism": "base rows sharded x%d, NCCL all-gather of [m][k] keys + K7 merge" % world if world > 1 else "single GPU"},
               "e2e": {"value": m * args.steps / e2e_s, "unit": "queries/s", "h2d_bytes_per_step": int(m * d * 4),
                       "d2h_bytes_per_step": int(m * k * 8 + (m * k * 4 + m * 8 if world == 1 else 0)),
                       "api": "expann_gpu_query (page-locked host buffers)" if world == 1 else "pinned H2D + query_device + all_gather + merge + decode + D2H"},
               "gpu_launches": int(launches),
               "roofline": roof,
               "clocks": clocks}
        if not args.no_cpu and world == 1 and not use_graph:
            qps, cores, kind, sample = cpu_reference_qps(base, q, k, metric)
            out["cpu_baseline"] = {"value": qps, "unit": "queries/s", "cores": cores, "kind": kind, "sample": sample}
        out["check"] = {"first_ids": got[0].tolist()}
        if use_graph:
            allids = ids_dev.cpu().numpy()
            hit = sum(len(set(a.tolist()) & set(b.tolist())) for a, b in zip(allids, truth_ids))
            out["check"]["recall_at_k_vs_exact"] = hit / float(m * k)
            out["config"]["ef"] = args.ef
            out["config"]["graph_build_s_per_shard"] = graph_build_s
        if use_quant and world == 1:
            # quality of the compressed scan: recall@10 against the exact fp32 scan of the same library (not timed)
            try:
                ex = expann_b200.BruteForceEngine(metric=1 if metric == "ip" else 0, device=local_rank)
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
        if world == 1 and not args.no_extras and not use_quant and not use_graph:
            try:
                out["scan_hbm"] = scan_hbm_leg(eng, torch, dq, k, hi - lo, d, stream, peak, peak_src, args.steps)
            except Exception as ex:  # noqa: BLE001
                out["scan_hbm"] = {"error": repr(ex)[:300]}
    eng.close()
    del eng
    if rank == 0:
        if world == 1 and args.workload == "c2" and not args.no_extras and not use_graph:
            # BASELINE config 3 (graph engine, QPS at recall@10 >= 0.95) measured in the same run
            try:
                from tools import bench_extras
                torch.cuda.empty_cache()
                out["graph_c3"] = bench_extras.graph_c3(local_rank, max(2, args.steps), args.warmup, cpu=not args.no_cpu, peak_gbs=peak)
                if "roofline" in out["graph_c3"]:   # DRAM bytes per launch of the ef = 40 search from the committed ncu capture
                    out["graph_c3"]["roofline"]["traffic"] = ncu_traffic("graph_search_kernel")
            except Exception as ex:  # noqa: BLE001
                out["graph_c3"] = {"error": repr(ex)[:300]}
        emit(out)
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
