"""The literal drop-in: the reference's OWN harness (bench_runner.h job machinery -> basic_bench::get_benchmark_data,
src/basic_bench.h:58-149; results written by bench_data_manager, src/bench_data_manager.h:70-73) driving the GPU engines
of expann_b200/host/reference_dropin.hpp, which derive from the real ann_engine<T, Derived> (src/ann_engine.h:16-29) and
take the real vec<float>. oracle/dropin_harness.cpp is compiled against /root/reference by oracle/Makefile (target
dropin) where that tree exists; the binary travels to the GPU box prebuilt (oracle/_ref/)."""
import json
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "oracle", "_ref", "expann_dropin")


def test_dropin_compiles_against_the_reference():
    if not os.path.isdir("/root/reference/src"):
        pytest.skip("no /root/reference here (GPU box): the prebuilt binary is used")
    from expann_b200 import _lib
    if not os.path.exists(_lib.LIB_PATH):
        _lib.build_library()
    subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle"), "dropin"], stdout=subprocess.DEVNULL)
    assert os.path.exists(BIN)
    # the header the reference-side maintainer adds derives from the reference's ann_engine, not from a look-alike
    hdr = open(os.path.join(ROOT, "expann_b200", "host", "reference_dropin.hpp")).read()
    assert '#include "ann_engine.h"' in hdr and "public ann_engine<T, gpu_brute_force_engine<T>>" in hdr


@pytest.mark.gpu
def test_reference_harness_over_gpu_engines(tmp_path):
    if not os.path.exists(BIN):
        pytest.skip("oracle/_ref/expann_dropin was not built (needs /root/reference at build time)")
    prefix = str(tmp_path / "dropin") + "/"
    r = subprocess.run([BIN, "4000", "2000", "10", prefix], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    latest = json.load(open(prefix + "data/latest.json"))        # written by the reference's bench_data_manager
    assert len(latest) == 7
    keys = {"time_per_query_ns", "time_to_build_ns", "average_distance", "average_squared_distance", "recall", "engine_name", "param_list"}
    assert all(set(b) == keys for b in latest)
    cpu = {b["param_list"]["ef_search_mult"]: b for b in latest if b["engine_name"].startswith("Anti-Topo Engine+")}
    gpu = {b["param_list"]["conf.ef_search_mult"]: b for b in latest if b["engine_name"].startswith("Anti-Topo Engine (B200")}
    bf = [b for b in latest if b["engine_name"].startswith("Brute-Force Engine (B200)")]
    assert len(cpu) == 3 and len(gpu) == 3 and len(bf) == 1
    assert bf[0]["recall"] == 1.0                                  # exact search == the reference's CPU ground truth
    for mult in ("1", "2", "4"):
        # same index file, same search: equal-or-better recall at each beam width (real-valued data: summation order can flip a
        # near-tie either way, hence the small allowance)
        assert gpu[mult]["recall"] >= cpu[mult]["recall"] - 0.005, (mult, gpu[mult]["recall"], cpu[mult]["recall"])
        assert abs(gpu[mult]["average_squared_distance"] - cpu[mult]["average_squared_distance"]) <= 1e-3 * cpu[mult]["average_squared_distance"]
    out = os.path.join(ROOT, "gpurun_out")
    if os.path.isdir(out):
        json.dump(latest, open(os.path.join(out, "dropin_latest.json"), "w"), indent=1)
