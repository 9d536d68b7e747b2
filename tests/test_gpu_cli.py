"""GPU: the C++ host CLI (expann_b200/host/expann_gpu_cli.cpp) runs the reference's config.json workflow
(src/main.cpp:52-109) through the C ABI and writes bench_data JSON (src/bench_data.h:8-43)."""
import json
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CLI = os.path.join(ROOT, "expann_b200", "bin", "expann_gpu")


def test_cli_builds_and_prints_usage():
    if not os.path.exists(CLI):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "expann_b200", "csrc"), "-j8"], stdout=subprocess.DEVNULL)
    out = subprocess.run([CLI, "--help"], capture_output=True, text=True)
    assert out.returncode == 0 and "--config" in out.stdout


@pytest.mark.gpu
def test_cli_config_workflow(tmp_path):
    cfg = tmp_path / "config_synthetic.json"
    cfg.write_text(json.dumps({"dataset": "Synthetic", "ds_name": "t_synth", "num_threads": 4, "n": 56000, "m": 400,
                               "d": 128, "k": 10}))
    prefix = str(tmp_path / "out") + "/"
    r = subprocess.run([CLI, "--config", str(cfg), "--n", "4000", "--m", "60", "--engines", "bf,graph,quant",
                        "--out_prefix", prefix], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    latest = json.load(open(prefix + "data/latest.json"))
    assert len(latest) == 1 + 6 + 1
    keys = {"time_per_query_ns", "time_to_build_ns", "average_distance", "average_squared_distance", "recall",
            "engine_name", "param_list"}
    assert all(set(b) == keys for b in latest)
    assert latest[0]["engine_name"].startswith("Brute-Force") and latest[0]["recall"] == 1.0
    graph = [b for b in latest if b["engine_name"].startswith("Anti-Topo")]
    rec = [b["recall"] for b in graph]
    assert [int(b["param_list"]["ef_search_mult"]) for b in graph] == [1, 2, 3, 4, 5, 6]
    assert rec[-1] >= rec[0] and rec[-1] >= 0.9
    assert latest[-1]["recall"] >= 0.9
    # second run appends to all.json, overwrites latest.json (src/bench_data_manager.h:70-73)
    r = subprocess.run([CLI, "--config", str(cfg), "--n", "4000", "--m", "60", "--engines", "bf", "--per_query", "1",
                        "--out_prefix", prefix], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    assert len(json.load(open(prefix + "data/latest.json"))) == 1
    assert len(json.load(open(prefix + "data/all.json"))) == 9


@pytest.mark.gpu
def test_cli_per_query_graph(tmp_path):
    """the reference harness's own loop (one query per call, src/basic_bench.h:82-126) through the C++ shim: the graph
    engine answers it with the one-CTA-per-query form of the search; search_width = 2 must not lower the recall"""
    cfg = tmp_path / "config_synthetic.json"
    cfg.write_text(json.dumps({"dataset": "Synthetic", "ds_name": "t_pq", "n": 4000, "m": 60, "d": 128, "k": 10}))
    rec = {}
    for width in ("1", "2"):
        prefix = str(tmp_path / ("out" + width)) + "/"
        r = subprocess.run([CLI, "--config", str(cfg), "--engines", "graph", "--per_query", "1", "--search_width", width,
                            "--out_prefix", prefix], capture_output=True, text=True, timeout=300)
        assert r.returncode == 0, r.stderr
        latest = json.load(open(prefix + "data/latest.json"))
        assert len(latest) == 6 and all(b["param_list"]["search_width"] == width for b in latest)
        rec[width] = [b["recall"] for b in latest]
        assert rec[width][-1] >= 0.9
    assert all(b >= a - 0.01 for a, b in zip(rec["1"], rec["2"])), rec
