"""Latency of ONE estimate_transformation-sized call through the host-buffer C ABI (what a ROS handler pays per
request): upload of the two keyframes, K2 + K3 (+ the second call), results back.  Prints medians over the overlapping
pairs of a small synthetic world, next to the same pairs on the CPU oracle (cv2 calls of the reference)."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from multi_robot_slam_separators_b200 import capi, synth
from oracle import matching, pnp

A, B = synth.make_robots(2, 30, n_feat=1000, dim=128, seed=1003)
same = [(a, b) for a in range(30) for b in range(30) if A.place[a] == B.place[b]][:20]
ctx = capi.Context(0)
fr, to = capi.Frames(ctx, 1, 1000), capi.Frames(ctx, 1, 1000)
prm = capi.pnp_params(synth.FX, synth.FY, synth.CX, synth.CY)
gprm = capi.guess_params(1241, 376, 20.0)
K = pnp.K_matrix(synth.FX, synth.FY, synth.CX, synth.CY)
rv0, tv0 = np.array(prm.guess_rvec), np.array(prm.guess_tvec)
t1, t2, tc = [], [], []
for rep in range(2):
    for (a, b) in same:
        t0 = time.perf_counter()
        fr.upload(0, A.desc[a:a + 1], A.kp2d[a:a + 1], A.kp3d[a:a + 1], A.count[a:a + 1], A.octave[a:a + 1])
        to.upload(0, B.desc[b:b + 1], B.kp2d[b:b + 1], B.kp3d[b:b + 1], B.count[b:b + 1], B.octave[b:b + 1])
        r1 = capi.verify_pairs(ctx, fr, to, [[0, 0]], prm, want_lists=False)
        t_first = time.perf_counter() - t0
        t0 = time.perf_counter()
        r2 = capi.estimate_transformations(ctx, fr, to, [[0, 0]], prm, gprm, want_lists=False)
        t_both = time.perf_counter() - t0
        if rep:
            t1.append(t_first); t2.append(t_both)
            t0 = time.perf_counter()
            pairs = matching.global_match(B.desc[b], A.desc[a], 0.8)
            obj, img, _, _ = matching.pnp_correspondences(pairs, A.kp3d[a], B.kp2d[b])
            if len(obj) >= 5:
                pnp.cv_solve_pnp_ransac(obj, img, K, rv0, tv0)
            tc.append(time.perf_counter() - t0)
print("one pair, host buffers in and out: first call (K2 + K3) median %.2f ms, whole service call (both calls) %.2f ms; "
      "CPU oracle first call (cv2 BFMatcher + solvePnPRansac, %d threads) %.2f ms" %
      (1e3 * np.median(t1), 1e3 * np.median(t2), os.cpu_count(), 1e3 * np.median(tc)))
ctx.close()
