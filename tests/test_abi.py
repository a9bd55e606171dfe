"""CPU tests (no GPU): the C-ABI shared library loads and exports every symbol include/pebble_cuda.h declares, the
ctypes binding covers exactly that set, the host-only entry points work, and the product fails LOUDLY without CUDA
(there is no CPU fallback anywhere in libpebblecuda.so)."""
import ctypes
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    txt = open(os.path.join(ROOT, "include", "pebble_cuda.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(pebble_[a-z0-9_]+)\s*\(", txt)))


def test_header_symbols_exported_and_bound():
    import pebblesdr_b200 as pb
    from pebblesdr_b200 import _lib
    names = _declared()
    assert len(names) > 70
    raw = ctypes.CDLL(pb.LIB_PATH)
    for n in names:
        assert hasattr(raw, n), "libpebblecuda.so does not export " + n
    assert sorted(_lib.SIGNATURES) == names, set(names) ^ set(_lib.SIGNATURES)
    assert pb.lib.pebble_abi_version() >= 1


def test_host_only_entry_points():
    import pebblesdr_b200 as pb
    # WindowFunction tables are built on the host (windowfunction.cpp:50-241): BLACKMANHARRIS coherent gain
    w = pb.WindowFunction(pb.WindowFunction.BLACKMANHARRIS, 8192)
    assert 0.35 < w.coherentGain < 0.37 and abs(w.window.max() - 1.0) < 1e-3
    # chain planner (downconvert.cpp:139-213) without a device
    lens = (ctypes.c_int * 16)()
    out = ctypes.c_double(0)
    n = pb.lib.pebble_downconvert_plan(10e6, 15000.0, 0, lens, 16, ctypes.byref(out))
    assert list(lens[:n]) == [3, 11, 11, 11, 11, 15, 27] and out.value == 78125.0
    n = pb.lib.pebble_downconvert_plan(61.44e6, 15000.0, 0, lens, 16, ctypes.byref(out))
    assert list(lens[:n]) == [3, 3, 3, 11, 11, 11, 11, 15, 19, 35] and out.value == 60000.0
    n = pb.lib.pebble_downconvert_plan(2e6, 15000.0, 0, lens, 16, ctypes.byref(out))
    assert list(lens[:n]) == [11, 11, 15, 19, 31] and out.value == 62500.0
    n = pb.lib.pebble_downconvert_plan(2e6, 15000.0, 1, lens, 16, ctypes.byref(out))   # SetDataRateSimple
    assert list(lens[:n]) == [51, 51, 51] and out.value == 250000.0


def test_no_cpu_fallback_without_cuda():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    import pebblesdr_b200 as pb
    assert pb.lib.pebble_device_count() <= 0
    with pytest.raises(pb.PebbleError):
        pb.CFastFIR()
    assert pb.lib.pebble_last_error()        # the failure is explained
    with pytest.raises(pb.PebbleError):
        pb.CDownConvert()
    with pytest.raises(pb.PebbleError):
        pb.Receiver(10e6, 2048)


def test_product_never_imports_oracle():
    # the oracle is test infrastructure: nothing under pebblesdr_b200/ may reference it
    for dp, _, fs in os.walk(os.path.join(ROOT, "pebblesdr_b200")):
        for f in fs:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                txt = open(os.path.join(dp, f), errors="ignore").read()
                assert "oracle" not in txt.replace("build_oracle", "").replace("oracle/", "ORACLE_PATH/") or f == "build.py", f
