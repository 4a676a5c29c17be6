"""Builds/loads the TEST-ONLY host compilation of the product's math headers (see harness.cpp)."""
import ctypes as C
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
SO = os.path.join(HERE, "_build", "libharness.so")


def load():
    src = os.path.join(HERE, "harness.cpp")
    csrc = os.path.join(HERE, "..", "..", "multi_robot_slam_separators_b200", "csrc")
    deps = [src] + [os.path.join(csrc, f) for f in os.listdir(csrc) if f.endswith(".cuh")]
    os.makedirs(os.path.dirname(SO), exist_ok=True)
    if not os.path.exists(SO) or any(os.path.getmtime(d) > os.path.getmtime(SO) for d in deps):
        # -ffp-contract=off + baseline x86-64 (no FMA): same arithmetic as the cv2 wheel and as nvcc -fmad=false
        subprocess.check_call(["g++", "-O2", "-std=c++17", "-ffp-contract=off", "-fPIC", "-shared", "-x", "c++", src,
                               "-o", SO])
    return C.CDLL(SO)
