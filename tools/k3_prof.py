"""Developer tool: per-phase cycle split of K3 (needs a library built with SEPB_K3_PROFILE=1)."""
import ctypes as C, os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from multi_robot_slam_separators_b200 import capi, synth
from multi_robot_slam_separators_b200.engine import Camera, SeparatorEngine

frames = int(sys.argv[1]) if len(sys.argv) > 1 else 2000
A, B = synth.make_robots(2, frames, n_feat=1000, dim=128, seed=2000)
eng = SeparatorEngine(Camera(synth.FX, synth.FY, synth.CX, synth.CY), 1000, 128, frames, frames)
eng.add_peer_keyframes(A.desc, A.kp2d, A.kp3d, A.count, A.octave)
eng.add_local_keyframes(B.desc, B.kp2d, B.kp3d, B.count, B.octave)
eng.receive_descriptors(A.netvlad)
d_q = torch.from_numpy(B.netvlad).cuda()
L = capi.lib()
L.sepb_debug_k3_prof.argtypes = [C.c_void_p, C.c_int]
out = (C.c_ulonglong * 8)()
res = eng.step_dev(d_q, 5, capi.NN_TENSOR); eng.sync()
L.sepb_debug_k3_prof(out, 1)
res = eng.step_dev(d_q, 5, capi.NN_TENSOR); eng.sync()
built = L.sepb_debug_k3_prof(out, 1)
v = np.array(list(out), dtype=np.float64)
it = res["iters_run"][res["n_matches"] >= 5].cpu().numpy()
print("profile build:", built, "pairs with RANSAC:", len(it), "iters mean/max:", it.mean(), it.max(),
      "hist chunks:", np.bincount((it + 31) // 32))
ch = max(v[4], 1)
print("chunk-solves:", int(v[4]))
nm = res["n_matches"][res["n_matches"] >= 5].cpu().numpy(); ni = res["n_inliers"][res["n_matches"] >= 5].cpu().numpy()
print("matches per RANSAC pair: mean %.1f min %d max %d; inliers mean %.1f" % (nm.mean(), nm.min(), nm.max(), ni.mean()))
for name, i in (("epnp5", 0), ("  of which jacobi 12x12", 5), ("rodrigues", 1), ("scoring (32 hyp)", 2)):
    print(f"{name:26s} {v[i] / ch / 1e3:10.1f} kcycles per chunk")
print(f"{'refit (per pair)':26s} {v[3] / max(len(it), 1) / 1e3:10.1f} kcycles")
print("jacobi 12x12 sweeps per solve (lane 0):", v[6] / ch)
