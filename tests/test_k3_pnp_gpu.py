"""K3 parity (GPU, through the C ABI): PnP-RANSAC inlier sets exactly equal to cv2.solvePnPRansac,
poses within 1e-4 rad / 1e-4 m (BASELINE.json north_star), on golden vectors and seeded keyframes."""
import numpy as np
import pytest

from multi_robot_slam_separators_b200 import capi, synth
from oracle import matching, pnp

cv2 = pytest.importorskip("cv2")
pytestmark = pytest.mark.gpu

ROT_TOL = 1e-4   # rad   (north_star)
TRANS_TOL = 1e-4  # m


@pytest.fixture(scope="module")
def ctx():
    c = capi.Context(0)
    yield c
    c.close()


def rot_angle(rv_a, rv_b):
    Ra, _ = cv2.Rodrigues(np.asarray(rv_a, float))
    Rb, _ = cv2.Rodrigues(np.asarray(rv_b, float))
    c = (np.trace(Ra.T @ Rb) - 1) / 2
    return float(np.arccos(np.clip(c, -1, 1)))


def oracle_pair(A, a, B, b, K, rv0, tv0, iters=300, min_inl=5):
    pairs = matching.global_match(B.desc[b][:B.count[b]], A.desc[a][:A.count[a]], 0.8)
    obj, img, fi, ti = matching.pnp_correspondences(pairs, A.kp3d[a], B.kp2d[b])
    if len(obj) < min_inl:
        return dict(status=0, M=len(obj), inliers=np.zeros(0, np.int32))
    ok, rv, tv, inl = pnp.cv_solve_pnp_ransac(obj, img, K, rv0, tv0, iterations=iters)
    st = 1 if (ok and len(inl) >= min_inl) else (2 if not ok else 3)
    return dict(status=st, M=len(obj), inliers=inl, rvec=rv, tvec=tv)


def check(res, p, ref):
    assert res["n_matches"][p] == ref["M"]
    assert res["status"][p] == ref["status"], (p, res["status"][p], ref["status"])
    if ref["status"] in (1, 3):
        inl = capi.inlier_indices(res["inlier_mask"][p], res["n_matches"][p])
        assert np.array_equal(inl, ref["inliers"]), p
        assert res["n_inliers"][p] == len(ref["inliers"])
        assert rot_angle(res["pose"][p, :3], ref["rvec"]) <= ROT_TOL
        assert np.abs(res["pose"][p, 3:] - ref["tvec"]).max() <= TRANS_TOL


def test_golden_pairs(ctx, golden):
    g = golden("pnp_ransac")
    K = g["K"]
    n = int(g["n"])
    F = 1000
    fr = capi.Frames(ctx, n, F)
    to = capi.Frames(ctx, n, F)
    z2 = np.zeros((n, F, 2), np.float32)
    z3 = np.ones((n, F, 3), np.float32)
    cnt = np.full(n, F, np.int32)
    fr.upload(0, np.stack([g[f"desc_from{i}"] for i in range(n)]), z2, np.stack([g[f"kp3d_from{i}"] for i in range(n)]), cnt)
    to.upload(0, np.stack([g[f"desc_to{i}"] for i in range(n)]), np.stack([g[f"kp2d_to{i}"] for i in range(n)]), z3, cnt)
    for iters in (300, 100, 1000):
        sel = [i for i in range(n) if int(g[f"iters{i}"]) == iters]
        prm = capi.pnp_params(K[0, 0], K[1, 1], K[0, 2], K[1, 2], iterations=iters)
        assert np.abs(np.array(prm.guess_rvec) - g["rv0"]).max() < 1e-12
        res = capi.verify_pairs(ctx, fr, to, [[i, i] for i in sel], prm)
        # and with the reference's own minimal-sample indices injected
        cs = np.stack([g[f"samples{i}"] for i in sel])
        cs_pad = np.zeros((len(sel), iters, 5), np.int32)
        cs_pad[:, :cs.shape[1]] = cs
        res2 = capi.verify_pairs(ctx, fr, to, [[i, i] for i in sel], prm, custom_samples=cs_pad)
        for p, i in enumerate(sel):
            ref = dict(status=1 if g[f"ok{i}"] else 2, M=len(g[f"obj{i}"]), inliers=g[f"inliers{i}"],
                       rvec=g[f"rvec{i}"], tvec=g[f"tvec{i}"])
            check(res, p, ref)
            check(res2, p, ref)


def test_synthetic_pairs_vs_cv2(ctx):
    A, B = synth.make_robots(2, 40, n_feat=1000, dim=128, seed=1003)
    # ragged counts and a degenerate frame
    A.count[3] = 700
    B.count[5] = 300
    A.count[7] = 4
    fa, fb = capi.frames_from_robot(ctx, A), capi.frames_from_robot(ctx, B)
    K = pnp.K_matrix(synth.FX, synth.FY, synth.CX, synth.CY)
    prm = capi.pnp_params(synth.FX, synth.FY, synth.CX, synth.CY)
    rv0, tv0 = np.array(prm.guess_rvec), np.array(prm.guess_tvec)
    same = [(a, b) for a in range(40) for b in range(40) if A.place[a] == B.place[b]]
    rng = np.random.default_rng(5)
    pairs = same + [(int(rng.integers(40)), int(rng.integers(40))) for _ in range(20)] + [(7, 1), (3, 5)]
    res = capi.verify_pairs(ctx, fa, fb, pairs, prm)
    n_ok = 0
    for p, (a, b) in enumerate(pairs):
        ref = oracle_pair(A, a, B, b, K, rv0, tv0)
        check(res, p, ref)
        n_ok += ref["status"] == 1
        if ref["status"] == 1:
            # the transform recovers the ground-truth relative pose of the synthetic world
            T = np.eye(4)
            T[:3] = res["transform"][p].reshape(3, 4)
            G = synth.relative_pose_gt(A, a, B, b)
            assert np.abs(T - G).max() < 0.05
    assert n_ok >= len(same) - 2


def test_persistent_scheduler_identical(ctx, golden):
    """The persistent work-queue kernel (idle warps compute chunks of other warps' pairs; the owner scans their inlier
    counts in hypothesis order) must equal the one-warp-per-pair kernel on every output, including pairs that finish in
    chunk 0 and degenerate pairs.  Few pairs against 740 workers: almost every chunk here comes from a helper."""
    A, B = synth.make_robots(2, 40, n_feat=1000, dim=128, seed=1003)
    A.count[3] = 700
    B.count[5] = 300
    A.count[7] = 4
    fa, fb = capi.frames_from_robot(ctx, A), capi.frames_from_robot(ctx, B)
    rng = np.random.default_rng(6)
    same = [(a, b) for a in range(40) for b in range(40) if A.place[a] == B.place[b]]
    pairs = same + [(int(rng.integers(40)), int(rng.integers(40))) for _ in range(30)] + [(7, 1), (3, 5)]
    for iters in (300, 40, 1000):
        prm = capi.pnp_params(synth.FX, synth.FY, synth.CX, synth.CY, iterations=iters)
        ctx.set_pnp_scheduler("per_pair")
        try:
            one = capi.verify_pairs(ctx, fa, fb, pairs, prm)
        finally:
            ctx.set_pnp_scheduler("persistent")
        two = capi.verify_pairs(ctx, fa, fb, pairs, prm)
        assert int((one["status"] == 1).sum()) >= len(same) - 2
        assert int((one["iters_run"] > 32).sum()) > 0 or iters <= 32
        for k in ("status", "n_matches", "n_inliers", "inlier_mask", "pose", "transform", "best_hyp", "iters_run", "covariance"):
            if k in one:
                assert np.array_equal(one[k], two[k], equal_nan=True), (iters, k)


def test_persistent_scheduler_identical_in_the_second_call(ctx):
    """The second call of estimate_transformation (guess = first result, only pairs whose first call succeeded) through
    the persistent kernel: every output equals the one-warp-per-pair run."""
    A, B = synth.make_robots(2, 40, n_feat=1000, dim=128, seed=1003)
    fa, fb = capi.frames_from_robot(ctx, A), capi.frames_from_robot(ctx, B)
    rng = np.random.default_rng(8)
    same = [(a, b) for a in range(40) for b in range(40) if A.place[a] == B.place[b]]
    pairs = same + [(int(rng.integers(40)), int(rng.integers(40))) for _ in range(30)]
    prm = capi.pnp_params(synth.FX, synth.FY, synth.CX, synth.CY)
    gprm = capi.guess_params(1241, 376, 20.0)
    ctx.set_pnp_scheduler("per_pair")
    try:
        one = capi.estimate_transformations(ctx, fa, fb, pairs, prm, gprm)
    finally:
        ctx.set_pnp_scheduler("persistent")
    two = capi.estimate_transformations(ctx, fa, fb, pairs, prm, gprm)
    assert int((one["status"] == 1).sum()) >= len(same) - 3
    for k in one:
        assert np.array_equal(one[k], two[k], equal_nan=True), k


def test_async_reupload_waits_for_the_covariance_kernel(ctx):
    """Double-buffer pattern of include/sepb200.h: an asynchronous re-upload of a store enqueued right after
    sepb_verify_pairs_async must wait for EVERY kernel of that call that reads the store -- including pose_cov, which runs
    after K3 -- so the results (covariance included) equal those of a blocking run on the original data."""
    import torch
    A, B = synth.make_robots(2, 40, n_feat=1000, dim=128, seed=1003)
    fa, fb = capi.frames_from_robot(ctx, A), capi.frames_from_robot(ctx, B)
    same = [(a, b) for a in range(40) for b in range(40) if A.place[a] == B.place[b]]
    pairs = np.array(same * 8, np.int32)
    prm = capi.pnp_params(synth.FX, synth.FY, synth.CX, synth.CY)
    want = capi.verify_pairs(ctx, fa, fb, pairs, prm)
    assert int((want["status"] == 1).sum()) >= len(same) - 2

    def pinned(shape, dtype):
        return torch.zeros(shape, dtype=getattr(torch, np.dtype(dtype).name)).pin_memory().numpy()

    pin = lambda a: torch.from_numpy(np.ascontiguousarray(a)).pin_memory().numpy()
    out = capi.verify_buffers(len(pairs), 1000, want_lists=True, alloc=pinned)
    pairs_pin = pin(pairs)
    junk = dict(desc=pin(np.zeros_like(A.desc)), kp2d=pin(np.full_like(A.kp2d, 7.0)), kp3d=pin(np.full_like(A.kp3d, np.nan)),
                count=pin(A.count), octave=pin(A.octave))
    for _ in range(3):
        capi.verify_pairs(ctx, fa, fb, pairs_pin, prm, out=out, async_=True)
        # overwrite BOTH stores behind the call, without waiting
        fa.upload(0, junk["desc"], junk["kp2d"], junk["kp3d"], junk["count"], junk["octave"], async_=True)
        fb.upload(0, junk["desc"], junk["kp2d"], junk["kp3d"], junk["count"], junk["octave"], async_=True)
        ctx.sync()
        fa.wait(); fb.wait()
        for k in ("status", "n_matches", "n_inliers", "pose", "transform", "inlier_mask", "covariance"):
            assert np.array_equal(out[k], want[k], equal_nan=True), k
        fa.upload(0, A.desc, A.kp2d, A.kp3d, A.count, A.octave)
        fb.upload(0, B.desc, B.kp2d, B.kp3d, B.count, B.octave)
    fa.close(); fb.close()
