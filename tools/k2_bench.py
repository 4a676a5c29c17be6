"""Scratch timing of K2 (Hamming match) on resident frames + pipe peaks."""
import ctypes as C
import sys, os, time
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from multi_robot_slam_separators_b200 import capi, synth

if os.environ.get("SEPB_LIB"):          # developer: time another build of the library on the same data
    capi.LIB_PATH = os.path.abspath(os.environ["SEPB_LIB"])
ctx = capi.Context(0)
print("popc/s %.3e  dfma/s %.3e" % (ctx.measure_peak("popc"), ctx.measure_peak("dfma")))
nF = int(sys.argv[1]) if len(sys.argv) > 1 else 256
P = int(sys.argv[2]) if len(sys.argv) > 2 else 20000
A, B = synth.make_robots(2, nF, n_feat=1000, dim=128, seed=1002)
if os.environ.get("SEPB_K2_COUNT"):      # developer: shorter frames -> fixed cost per pair of the matcher kernels
    A.count[:] = int(os.environ["SEPB_K2_COUNT"]); B.count[:] = int(os.environ["SEPB_K2_COUNT"])
fa, fb = capi.frames_from_robot(ctx, A), capi.frames_from_robot(ctx, B)
rng = np.random.default_rng(0)
pairs = torch.from_numpy(rng.integers(0, nF, (P, 2)).astype(np.int32)).cuda()
nm = torch.zeros(P, dtype=torch.int32, device="cuda")
nu = torch.zeros(P, dtype=torch.int32, device="cuda")
mf = torch.zeros((P, 1000), dtype=torch.int16, device="cuda")
mt = torch.zeros((P, 1000), dtype=torch.int16, device="cuda")
prm = capi.MatchParams(0.8, 0)
L = capi.lib()
stream = torch.cuda.ExternalStream(ctx.stream)
which = sys.argv[3] if len(sys.argv) > 3 else "auto"
ctx.set_match_kernel(which)
ctx.set_match_tile(int(os.environ.get("SEPB_K2T_TN", "0")) % 256)   # tool-side switch -> SEPB_OPT_MATCH_TILE
def run():
    rc = L.sepb_match_pairs_dev(ctx._h, fa._h, fb._h, C.c_void_p(pairs.data_ptr()), P, C.byref(prm), None, None,
                                C.c_void_p(nm.data_ptr()), C.c_void_p(mf.data_ptr()), C.c_void_p(mt.data_ptr()),
                                C.c_void_p(nu.data_ptr()))
    ctx.check(rc)
torch.cuda.synchronize()
with torch.cuda.stream(stream):
    for _ in range(3): run()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream); 
    for _ in range(5): run()
    e1.record(stream); e1.synchronize()
ms = e0.elapsed_time(e1) / 5
print(which, "K2: %d pairs in %.3f ms -> %.3e pairs/s, %.3e popc/s" % (P, ms, P / ms * 1e3, P * 8e6 / ms * 1e3))
print("mean matches", nm.float().mean().item())
