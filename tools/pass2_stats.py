"""RANSAC statistics of the two calls of estimate_transformation on a synthetic world: matches, inliers, iterations run
(why the second call's K3 costs as much as the first: the window matcher's lists have a lower inlier ratio)."""
import sys, os, numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from multi_robot_slam_separators_b200 import capi, synth
from multi_robot_slam_separators_b200.engine import Camera, SeparatorEngine
n = 4000
A, B = synth.make_robots(2, n, n_feat=1000, dim=128, seed=2000)
eng = SeparatorEngine(Camera(synth.FX, synth.FY, synth.CX, synth.CY), 1000, 128, n, n, device=0)
eng.add_peer_keyframes(A.desc, A.kp2d, A.kp3d, A.count, A.octave)
eng.add_local_keyframes(B.desc, B.kp2d, B.kp3d, B.count, B.octave)
eng.receive_descriptors(A.netvlad)
d_q = torch.from_numpy(B.netvlad).cuda()
idx, _ = eng.candidates_dev(d_q, 5, capi.NN_TENSOR)
pairs = eng.pairs_from_candidates(idx)
r1 = {k: v.clone() for k, v in eng.verify_dev(pairs).items()}
eng.sync()
r2 = eng.verify_dev(pairs, two_pass=True)
eng.sync()
ok1 = (r1["status"] == 1)
ok2 = (r2["status"] == 1)
f = lambda t, m: float(t[m].float().mean().item())
print("pass1: ok", int(ok1.sum()), "matches", f(r1["n_matches"], ok1), "inliers", f(r1["n_inliers"], ok1), "iters", f(r1["iters_run"], ok1))
print("pass2: ok", int(ok2.sum()), "matches", f(r2["n_matches"], ok1), "inliers", f(r2["n_inliers"], ok1), "iters", f(r2["iters_run"], ok1))
it = r2["iters_run"][ok1].cpu().numpy()
print("pass2 iters histogram (bins of 32):", np.bincount((it - 1) // 32)[:10])
it = r1["iters_run"][ok1].cpu().numpy()
print("pass1 iters histogram (bins of 32):", np.bincount((it - 1) // 32)[:10])
