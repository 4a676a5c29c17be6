"""In-tree build of libsepb200.so (nvcc, sm_100a only).  `python -m multi_robot_slam_separators_b200.build`."""
from __future__ import annotations

import hashlib
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
ROOT = os.path.dirname(HERE)
LIB = os.path.join(HERE, "libsepb200.so")
OBJ = os.path.join(HERE, "build")

ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
COMMON = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=hidden",
          "-I", os.path.join(ROOT, "include")]
# SEPB_DEVELOPER=1: the library reads its A/B switches (SEPB_K1_SPLITS, SEPB_K3_WPC, ...) from the environment;
# a normal build compiles every getenv out of the launch paths
if os.environ.get("SEPB_DEVELOPER"):
    COMMON.append("-DSEPB_DEVELOPER")
# per-file extra flags.  pnp_ransac.cu replays OpenCV's double-precision operation order bit for
# bit (x86-64 baseline build of cv2 has no FMA contraction), hence -fmad=false there.
# SEPB_K3_DEFS: developer A/B switches for the K3 translation unit (e.g. "-DSEPB_JACOBI_ILP=0 -DSEPB_SCORE_ILP=2")
EXTRA = {"guess_match.cu": ["-fmad=false"],
         "pnp_ransac.cu": ["-fmad=false"] + (["-DSEPB_K3_PROFILE"] if os.environ.get("SEPB_K3_PROFILE") else [])
                          + os.environ.get("SEPB_K3_DEFS", "").split(),
         # SEPB_K2T_DEFS: same for the tensor matcher (e.g. "-DSEPB_K2T_TRACE=1000": clock64 timeline of that pair's CTA)
         "hamming_mma.cu": os.environ.get("SEPB_K2T_DEFS", "").split()}


def nvcc() -> str:
    for c in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if c and os.path.exists(c):
            return c
    raise RuntimeError("nvcc not found")


def _sources():
    return sorted(f for f in os.listdir(CSRC) if f.endswith(".cu"))


def _stamp():
    h = hashlib.sha256()
    for f in sorted(os.listdir(CSRC)) + ["../../include/sepb200.h", "../build.py"]:
        with open(os.path.join(CSRC, f), "rb") as fh:
            h.update(fh.read())
    h.update((os.environ.get("SEPB_K3_PROFILE", "") + "|" + os.environ.get("SEPB_K3_DEFS", "") + "|"
              + os.environ.get("SEPB_K2T_DEFS", "") + "|" + os.environ.get("SEPB_DEVELOPER", "")).encode())
    return h.hexdigest()


def build(force: bool = False, verbose: bool = False) -> str:
    os.makedirs(OBJ, exist_ok=True)
    stamp_file = os.path.join(OBJ, "stamp")
    stamp = _stamp()
    if not force and os.path.exists(LIB) and os.path.exists(stamp_file) and open(stamp_file).read() == stamp:
        return LIB
    objs = []
    procs = []
    for src in _sources():
        o = os.path.join(OBJ, src[:-3] + ".o")
        cmd = [nvcc()] + ARCH + COMMON + EXTRA.get(src, []) + (["-Xptxas", "-v"] if verbose else []) + \
              ["-c", os.path.join(CSRC, src), "-o", o]
        procs.append((src, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
        objs.append(o)
    for src, p in procs:
        out, _ = p.communicate()
        if verbose or p.returncode:
            sys.stderr.write(out)
        if p.returncode:
            raise RuntimeError(f"nvcc failed on {src}")
    cmd = [nvcc()] + ARCH + ["-shared", "-o", LIB] + objs
    subprocess.check_call(cmd)
    with open(stamp_file, "w") as fh:
        fh.write(stamp)
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
