import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from multi_robot_slam_separators_b200 import capi, synth
ctx = capi.Context(0)
A, B = synth.make_robots(2, 48, n_feat=1000, dim=128, seed=1207)
A.count[3] = 700
B.count[5] = 300
pin = lambda a: torch.from_numpy(np.ascontiguousarray(a)).pin_memory().numpy()
hA = {k: pin(getattr(A, k)) for k in ("desc", "kp2d", "kp3d", "count", "octave")}
hB = {k: pin(getattr(B, k)) for k in ("desc", "kp2d", "kp3d", "count", "octave")}
rng = np.random.default_rng(3)
same = [(a, b) for a in range(48) for b in range(48) if A.place[a] == B.place[b]]
pairs = same + [(int(rng.integers(48)), int(rng.integers(48))) for _ in range(60)]
prm = capi.pnp_params(synth.FX, synth.FY, synth.CX, synth.CY)
gprm = capi.guess_params(synth.WIDTH, synth.HEIGHT, 20.0)
fa, fb = capi.frames_from_robot(ctx, A), capi.frames_from_robot(ctx, B)
la, lb = capi.Frames(ctx, 48, 1000), capi.Frames(ctx, 48, 1000)
want1 = capi.verify_pairs(ctx, fa, fb, pairs, prm)
la.upload(0, hA["desc"], hA["kp2d"], hA["kp3d"], hA["count"], hA["octave"], lazy=True)
lb.upload(0, hB["desc"], hB["kp2d"], hB["kp3d"], hB["count"], hB["octave"], lazy=True)
got1 = capi.verify_pairs(ctx, la, lb, pairs, prm)
print("pulled", la.fetched_bytes, lb.fetched_bytes)
for k in want1:
    eq = np.array_equal(got1[k], want1[k], equal_nan=True)
    print(k, eq)
    if not eq:
        bad = [i for i in range(len(pairs)) if not np.array_equal(got1[k][i], want1[k][i], equal_nan=True)]
        print("  bad pairs", bad[:10], "of", len(bad), "n_matches", want1["n_matches"][bad[:10]], "status", want1["status"][bad[:10]], got1["status"][bad[:10]])
        if k in ("pose", "covariance"):
            print(got1[k][bad[0]], want1[k][bad[0]])
ctx.sync()
for f in (fa, fb, la, lb): f.close()
ctx.close()
