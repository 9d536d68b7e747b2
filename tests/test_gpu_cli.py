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


def _write_tvecs(path, arr):
    """.fvecs / .ivecs: per record int32 d, then d 4-byte values (src/dataset_loader.h:96-125)"""
    import numpy as np
    arr = np.ascontiguousarray(arr)
    n, d = arr.shape
    rec = np.empty((n, d + 1), dtype=np.int32)
    rec[:, 0] = d
    rec[:, 1:] = arr.view(np.int32)
    rec.tofile(path)


@pytest.mark.gpu
def test_cli_reads_fvecs_ivecs(tmp_path):
    """--dataset Sift1M (load_sift1m, src/dataset_loader.h:127-168): base / query vectors from .fvecs files; with fewer than
    1M base rows the ground truth is recomputed by exhaustive search (the shipped .ivecs belongs to the full base), so the
    brute-force engine must come out at recall 1.0 and the graph engine high"""
    import numpy as np
    rng = np.random.default_rng(3)
    z = rng.standard_normal((3064, 12)).astype(np.float32) @ rng.standard_normal((12, 128)).astype(np.float32)
    x = np.clip(np.rint(40.0 + 12.0 * z), 0, 255).astype(np.float32)
    sd = tmp_path / "sift"
    sd.mkdir()
    _write_tvecs(str(sd / "sift_base.fvecs"), x[:3000])
    _write_tvecs(str(sd / "sift_query.fvecs"), x[3000:])
    _write_tvecs(str(sd / "sift_groundtruth.ivecs"), np.zeros((64, 100), np.int32))   # present but not used for a 3000-row base
    prefix = str(tmp_path / "out") + "/"
    r = subprocess.run([CLI, "--dataset", "Sift1M", "--sift_dir", str(sd) + "/", "--ds_name", "t_sift", "--n", "1000000", "--m", "10000",
                        "--k", "10", "--engines", "bf,graph", "--out_prefix", prefix], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    latest = json.load(open(prefix + "data/latest.json"))
    assert len(latest) == 7 and latest[0]["recall"] == 1.0
    assert latest[-1]["recall"] >= 0.9
    # the distances the harness reports are those of the vectors in the files (mean distance to the nearest neighbour)
    d2 = ((x[3000:, None, :] - x[None, :3000, :]) ** 2).sum(-1).min(1)
    assert abs(latest[0]["average_squared_distance"] - float(d2.mean())) <= 1e-3 * float(d2.mean())
    # a truncated file is reported, not crashed on
    (sd / "sift_base.fvecs").write_bytes(b"\\x80\\x00\\x00")
    r = subprocess.run([CLI, "--dataset", "Sift1M", "--sift_dir", str(sd) + "/", "--engines", "bf", "--out_prefix", prefix],
                       capture_output=True, text=True, timeout=60)
    assert r.returncode == 2 and "not found" in r.stderr


@pytest.mark.gpu
def test_cli_binary_dataset_cache(tmp_path):
    """binary form of the reference's dataset cache (src/dataset_loader.h:42-95): the second run loads vectors, queries and
    ground truth from the cache file and reports the same results; a corrupt cache is ignored and rebuilt"""
    cache = str(tmp_path / "data" / "t.dataset.bin")
    args = [CLI, "--n", "3000", "--m", "40", "--d", "64", "--k", "10", "--engines", "bf", "--cache", "1", "--cache_file", cache]
    runs = []
    for i in range(2):
        prefix = str(tmp_path / f"out{i}") + "/"
        r = subprocess.run(args + ["--out_prefix", prefix], capture_output=True, text=True, timeout=300)
        assert r.returncode == 0, r.stderr
        runs.append((r.stdout, json.load(open(prefix + "data/latest.json"))))
    assert "dataset cache: saved" in runs[0][0] and "dataset cache: loaded" in runs[1][0]
    assert os.path.getsize(cache) == 8 + 40 + (3000 + 40) * 64 * 4 + 40 * (8 + 10 * 8)
    a, b = runs[0][1][0], runs[1][1][0]
    assert a["recall"] == b["recall"] == 1.0 and a["average_squared_distance"] == b["average_squared_distance"]
    with open(cache, "r+b") as f:          # claim 2^40 vectors: the sizes are checked against the file size
        f.seek(16)
        f.write((2 ** 40).to_bytes(8, "little"))
    prefix = str(tmp_path / "out2") + "/"
    r = subprocess.run(args + ["--out_prefix", prefix], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "dataset cache: saved" in r.stdout
    assert json.load(open(prefix + "data/latest.json"))[0]["recall"] == 1.0
