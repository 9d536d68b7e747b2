"""CPU: the C-ABI library loads and exports exactly what include/expann_gpu.h declares; no compute calls."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    txt = open(os.path.join(ROOT, "include", "expann_gpu.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(expann_gpu_[a-z_]+)\s*\(", txt)))


@pytest.fixture(scope="module")
def lib():
    from expann_b200 import _lib
    if not os.path.exists(_lib.LIB_PATH):
        _lib.build_library()
    return _lib.load()


def test_header_and_binding_agree(lib):
    from expann_b200 import _lib
    assert header_symbols() == sorted(_lib.SYMBOLS), "include/expann_gpu.h and expann_b200/_lib.py disagree"


def test_all_symbols_exported(lib):
    for name in header_symbols():
        assert hasattr(lib, name), name
    assert b"sm_100a" in lib.expann_gpu_version()


def test_struct_layout_matches_header(lib):
    from expann_b200 import _lib
    assert C.sizeof(_lib.Cfg) == 80
    assert C.sizeof(_lib.Stats) == 18 * 8


def test_no_cpu_fallback(lib):
    """without a GPU, creating an engine must fail loudly (the product path never computes on the CPU)"""
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    import expann_b200
    with pytest.raises(expann_b200.ExpannError):
        expann_b200.BruteForceEngine()


def test_product_does_not_touch_oracle():
    """nothing under expann_b200/ may import, link or execute oracle/"""
    pkg = os.path.join(ROOT, "expann_b200")
    for dp, _, fs in os.walk(pkg):
        for f in fs:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp", ".hpp")) or f == "Makefile":
                txt = open(os.path.join(dp, f), errors="ignore").read()
                assert "oracle" not in txt.lower() or f.endswith(".md"), os.path.join(dp, f)


def test_documented_options_exist():
    """every tuning knob the header documents for expann_gpu_set_option is accepted by the implementation, and every
    knob the implementation accepts is documented (string-level check; no GPU needed)"""
    import re
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    hdr = open(os.path.join(root, "include", "expann_gpu.h")).read()
    doc = hdr[hdr.index("tuning knobs by name"):hdr.index("int expann_gpu_set_option")]
    documented = set(re.findall(r'"([a-z_0-9]+)"', doc))
    src = open(os.path.join(root, "expann_b200", "csrc", "api.cu")).read()
    accepted = set(re.findall(r'nm == "([a-z_0-9]+)"', src))
    assert documented == accepted, (sorted(documented - accepted), sorted(accepted - documented))
