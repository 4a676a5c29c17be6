"""CPU-only: the C-ABI library builds, loads and exports every symbol include/sepb200.h declares."""
import ctypes
import os
import re

from multi_robot_slam_separators_b200 import build, capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "sepb200.h")).read()
    return sorted(set(re.findall(r"SEPB_API[^;(]*?\b(sepb_\w+)\s*\(", src)))


def test_library_builds_and_exports_all_symbols():
    path = build.build()
    L = ctypes.CDLL(path)
    names = declared_symbols()
    assert len(names) >= 10
    for n in names:
        assert hasattr(L, n), f"missing export {n}"


def test_no_cpu_fallback_without_device():
    import torch
    if torch.cuda.is_available():
        return
    try:
        capi.Context(0)
    except capi.SepbError:
        return
    raise AssertionError("Context creation must fail loudly without a CUDA device")
