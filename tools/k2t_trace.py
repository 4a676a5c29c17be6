"""Timeline of one K2T CTA (developer build: SEPB_K2T_DEFS="-DSEPB_K2T_TRACE=<pair>" python -c 'build(force=True)').

Prints, relative to the CTA's start and in SM clocks, when the MMA warp, the epilogue group and the producers
passed each mbarrier hand-off of every accumulator tile (slot layout: csrc/hamming_mma.cu)."""
import ctypes as C
import json
import os
import sys
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from multi_robot_slam_separators_b200 import capi, synth

ctx = capi.Context(0)
nF, P = 256, int(sys.argv[1]) if len(sys.argv) > 1 else 20000
A, B = synth.make_robots(2, nF, n_feat=1000, dim=128, seed=1002)
fa, fb = capi.frames_from_robot(ctx, A), capi.frames_from_robot(ctx, B)
rng = np.random.default_rng(0)
pairs = torch.from_numpy(rng.integers(0, nF, (P, 2)).astype(np.int32)).cuda()
nm = torch.zeros(P, dtype=torch.int32, device="cuda")
nu = torch.zeros(P, dtype=torch.int32, device="cuda")
mf = torch.zeros((P, 1000), dtype=torch.int16, device="cuda")
mt = torch.zeros((P, 1000), dtype=torch.int16, device="cuda")
prm = capi.MatchParams(0.8, 0)
L = capi.lib()
ctx.set_match_kernel("tensor")
for _ in range(3):
    ctx.check(L.sepb_match_pairs_dev(ctx._h, fa._h, fb._h, C.c_void_p(pairs.data_ptr()), P, C.byref(prm), None, None,
                                     C.c_void_p(nm.data_ptr()), C.c_void_p(mf.data_ptr()), C.c_void_p(mt.data_ptr()),
                                     C.c_void_p(nu.data_ptr())))
torch.cuda.synchronize()
tr = (C.c_longlong * 512)()
L.sepb_debug_k2t_trace.argtypes = [C.POINTER(C.c_longlong), C.c_int]
assert L.sepb_debug_k2t_trace(tr, 512) == 0
t = np.array(tr[:], dtype=np.int64)
t0 = t[0]
rel = lambda a: (np.asarray(a) - t0).astype(np.int64)
print("setup done %d, main loop done %d, tail done %d (clocks from CTA start)" % tuple(rel(t[1:4])))
wide = os.environ.get("SEPB_K2T_TN", "256") == "256"      # geometry the launch picks for 1000-feature frames
G = 32 if wide else 64
mw, mi, ew, er, ed = (rel(t[16:16 + G]), rel(t[80:80 + G]), rel(t[144:144 + G]), rel(t[208:208 + G]),
                      rel(t[272:272 + G]))
print("  g  mma_wait mma_issued  epi_wait acc_ready released | issue-wait ready-issued drain")
for g in range(G):
    print("%3d %9d %10d %9d %9d %8d | %10d %12d %5d" % (g, mw[g], mi[g], ew[g], er[g], ed[g], mi[g] - mw[g],
                                                        er[g] - mi[g], ed[g] - er[g]))
n_it = 16 if wide else 24
pw, pf, pd, aw = rel(t[336:336 + n_it]), rel(t[368:368 + n_it]), rel(t[400:400 + n_it]), rel(t[464:464 + n_it])
print(" it prod_wait buf_free expanded mma_waits_A | wait expand  mma_stall_on_A")
for i in range(n_it):
    print("%3d %9d %8d %8d %11d | %4d %6d %6d" % (i, pw[i], pf[i], pd[i], aw[i], pf[i] - pw[i], pd[i] - pf[i],
                                                 max(0, pd[i] - aw[i])))
print("From stages expanded (chunk, j):", {(c, j): int(rel(t[432 + 4 * c + j])) for c in range(3) for j in range(3)
                                           if t[432 + 4 * c + j]})
steady = slice(4, G - 4)
summary = {
    "main_loop_clocks": int(rel(t[2]) - rel(t[1])), "setup_clocks": int(rel(t[1])), "tail_clocks": int(t[3] - t[2]),
    "per_tile_clocks": float((rel(t[2]) - rel(t[1])) / G),
    "issue_interval_mean": float(np.diff(mi[steady]).mean()),
    "mma_stall_on_tmem_mean": float((mi - mw)[steady].mean()),
    "mma_latency_issue_to_ready_mean": float((er - mi)[steady].mean()),
    "epilogue_stall_mean": float((er - ew)[steady].mean()),
    "drain_mean": float((ed - er)[steady].mean()),
}
print(json.dumps(summary))
