"""Times the 2-view bundle adjustment kernel alone (sepb_bundle_adjust_dev) on verified pairs of a synthetic world:
    python tools/ba_time.py [--frames 1500] [--lib path/to/libsepb200.so]
The world is made once per process; pass --lib to time a differently built library on the same data (A/B builds)."""
import argparse
import ctypes as C
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--frames", type=int, default=1500)
    ap.add_argument("--feat", type=int, default=1000)
    ap.add_argument("--lib", default=None)
    ap.add_argument("--reps", type=int, default=5)
    a = ap.parse_args()
    from multi_robot_slam_separators_b200 import capi, synth
    if a.lib:
        capi.LIB_PATH = os.path.abspath(a.lib)
    import torch
    t0 = time.time()
    A, B = synth.make_robots(2, a.frames, n_feat=a.feat, dim=8, seed=4321, overlap=0.5)
    pairs = np.array([(o, l) for l in range(a.frames) for o in np.nonzero(A.place == B.place[l])[0][:2]], np.int32)
    ctx = capi.Context(0)
    fa, fb = capi.frames_from_robot(ctx, A), capi.frames_from_robot(ctx, B)
    prm = capi.pnp_params(synth.FX, synth.FY, synth.CX, synth.CY)
    res = capi.verify_pairs(ctx, fa, fb, pairs, prm)
    ok = np.nonzero(res["status"] == 1)[0]
    pairs = np.ascontiguousarray(pairs[ok])
    P, F = len(ok), a.feat
    print("world %.1f s, %d verified pairs, mean inliers %.0f" % (time.time() - t0, P, res["n_inliers"][ok].mean()), flush=True)
    dev = lambda x: torch.from_numpy(np.ascontiguousarray(x)).cuda()  # noqa: E731
    d_pairs, d_nm = dev(pairs), dev(res["n_matches"][ok])
    d_mf, d_mt = dev(res["match_from"][ok].view(np.int16)), dev(res["match_to"][ok].view(np.int16))
    d_cov = dev(res["covariance"][ok])
    src = {k: dev(res[k][ok] if k != "inlier_mask" else res[k][ok].view(np.int32)) for k in ("status", "n_inliers", "pose", "transform", "inlier_mask")}
    d_stats = torch.zeros((P, 4), dtype=torch.float64, device="cuda")
    b = capi.ba_params(synth.BASELINE * synth.FY / synth.FX)
    p = lambda t: C.c_void_p(t.data_ptr())  # noqa: E731
    stream = torch.cuda.ExternalStream(ctx.stream)
    ms = []
    for rep in range(a.reps + 1):
        w = {k: v.clone() for k, v in src.items()}
        torch.cuda.synchronize()
        with torch.cuda.stream(stream):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            ctx.check(capi.lib().sepb_bundle_adjust_dev(ctx._h, fa._h, fb._h, p(d_pairs), P, p(d_nm), p(d_mf), p(d_mt), p(d_cov),
                                                        C.byref(prm), C.byref(b), p(w["status"]), p(w["n_inliers"]), p(w["pose"]),
                                                        p(w["transform"]), p(w["inlier_mask"]), (F + 31) // 32, p(d_stats)))
            e1.record()
        torch.cuda.synchronize()
        if rep:
            ms.append(e0.elapsed_time(e1))
    st = d_stats.cpu().numpy()
    print("lib %s: BA %.3f ms (min %.3f) for %d pairs = %.2f us/pair; LM iterations mean %.1f; still ok %d; dropped inliers mean %.1f"
          % (os.path.basename(capi.LIB_PATH), np.median(ms), min(ms), P, 1e3 * np.median(ms) / P, st[:, 0].mean(),
             int((w["status"] == 1).sum().item()), st[:, 2].mean()), flush=True)
    fa.close(); fb.close(); ctx.close()


if __name__ == "__main__":
    main()
