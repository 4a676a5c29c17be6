"""Parity on the shapes of BASELINE.json configs[3] (KITTI-sized frames: 2000 features per frame, 4096-d NetVLAD,
RANSAC 1000 hypotheses) and the batched-sweep shape of configs[4], through the C ABI, against the oracle
(scipy cdist / cv2.BFMatcher / cv2.solvePnPRansac restatements)."""
import numpy as np
import pytest

from multi_robot_slam_separators_b200 import capi, synth
from multi_robot_slam_separators_b200.engine import Camera, SeparatorEngine
from oracle import find_matches as fm, matching, pnp

pytestmark = pytest.mark.gpu

ROT_TOL = 1e-4     # rad  (north_star)
TRANS_TOL = 1e-4   # m


def _check_pairs(res, matches, A, B, K, rv0, tv0, iterations):
    n_ok = 0
    for p, (l, o) in enumerate(matches):
        pairs = matching.global_match(B.desc[l][:B.count[l]], A.desc[o][:A.count[o]], 0.8)
        obj, img, fi, ti = matching.pnp_correspondences(pairs, A.kp3d[o], B.kp2d[l])
        m = int(res["n_matches"][p])
        assert m == len(fi), (p, m, len(fi))
        assert np.array_equal(res["match_from"][p, :m], fi.astype(np.uint16))
        assert np.array_equal(res["match_to"][p, :m], ti.astype(np.uint16))
        if len(obj) >= 5:
            ok, rv, tv, inl = pnp.cv_solve_pnp_ransac(obj, img, K, rv0, tv0, iterations=iterations)
            got = capi.inlier_indices(res["inlier_mask"][p], m)
            if ok:
                assert np.array_equal(got, inl), p
                assert np.abs(res["pose"][p, 3:] - tv).max() < TRANS_TOL and np.abs(res["pose"][p, :3] - rv).max() < ROT_TOL
                n_ok += int(res["status"][p] == 1)
    return n_ok


@pytest.mark.parametrize("mode", [capi.NN_EXACT, capi.NN_TENSOR])
@pytest.mark.parametrize("match_kernel", ["tensor", "popc"])
def test_config4_kitti_sized_frames(mode, match_kernel):
    n, F, D, H = 16, 2000, 4096, 1000
    A, B = synth.make_robots(2, n, n_feat=F, dim=D, seed=4000)
    A.count[2] = 1500        # ragged frames
    B.count[3] = 1999
    eng = SeparatorEngine(Camera(synth.FX, synth.FY, synth.CX, synth.CY), F, D, n, n, iterations=H)
    try:
        eng.ctx.set_match_kernel(match_kernel)
        eng.add_peer_keyframes(A.desc, A.kp2d, A.kp3d, A.count, A.octave)
        eng.add_local_keyframes(B.desc, B.kp2d, B.kp3d, B.count, B.octave)
        eng.receive_descriptors(A.netvlad)
        thr = float(np.median(fm.topk(B.netvlad, A.netvlad, 1)[1]))     # about half of the rows pass
        matches = eng.find_matches(B.netvlad, thr, 20, mode=mode)
        assert matches == fm.find_matches(B.netvlad, A.netvlad, thr, 20) and len(matches) >= 3
        idx, dist = eng.db.topk(B.netvlad, 5, mode)
        ridx, rdist = fm.topk(B.netvlad, A.netvlad, 5)
        assert np.array_equal(idx, ridx) and np.array_equal(dist, rdist)     # bit-exact ids and f64 distances
        res = eng.estimate_transformations(matches)
        K = pnp.K_matrix(synth.FX, synth.FY, synth.CX, synth.CY)
        rv0, tv0 = np.array(eng.pnp.guess_rvec), np.array(eng.pnp.guess_tvec)
        n_ok = _check_pairs(res, matches, A, B, K, rv0, tv0, H)
        assert n_ok >= 1
    finally:
        eng.close()


def test_config5_sweep_batch_is_order_independent():
    """configs[4]: one large batch of candidate pairs drawn from a resident frame store.  Size-independent property:
    every pair's result depends on the pair alone -- a 20 000-pair batch (persistent K3 work queue, multi-wave K2) gives,
    pair for pair, what the same pairs give in small one-warp-per-pair batches, and duplicates agree."""
    n, F = 64, 1000
    A, B = synth.make_robots(2, n, n_feat=F, dim=8, seed=5000)
    ctx = capi.Context(0)
    try:
        fa, fb = capi.frames_from_robot(ctx, A), capi.frames_from_robot(ctx, B)
        prm = capi.pnp_params(synth.FX, synth.FY, synth.CX, synth.CY)
        rng = np.random.default_rng(11)
        same = np.array([(a, b) for a in range(n) for b in range(n) if A.place[a] == B.place[b]], np.int32)
        P = 20000
        pairs = np.empty((P, 2), np.int32)
        hot = rng.random(P) < 0.2                   # 20 % overlapping pairs, the rest random (mostly rejected by matching)
        pairs[hot] = same[rng.integers(len(same), size=int(hot.sum()))]
        pairs[~hot] = rng.integers(n, size=(int((~hot).sum()), 2))
        big = capi.verify_pairs(ctx, fa, fb, pairs, prm, want_lists=False)
        assert int((big["status"] == 1).sum()) > 1000
        uniq, first = np.unique(pairs, axis=0, return_index=True)
        ctx.set_pnp_scheduler("per_pair")      # reference scheduling: one warp runs every chunk of its pair
        small = capi.verify_pairs(ctx, fa, fb, uniq[:500], prm, want_lists=False)
        key = {tuple(p): i for i, p in enumerate(uniq[:500])}
        rows = np.array([key.get(tuple(p), -1) for p in pairs])
        sel = rows >= 0
        assert sel.sum() > 1000
        for k in ("status", "n_matches", "n_inliers", "pose", "transform", "best_hyp", "iters_run"):
            assert np.array_equal(big[k][sel], small[k][rows[sel]], equal_nan=True), k
    finally:
        ctx.close()


def test_more_correspondences_than_the_staging_area_holds():
    """K3 stages a pair's correspondences in shared memory when they fit (1203 of them); beyond that it reads them from
    global memory.  Two 2000-feature frames with the same descriptors give 2000 correspondences: inliers and pose must
    still equal cv2.solvePnPRansac."""
    import cv2
    F = 2000
    rng = np.random.default_rng(99)
    desc = rng.integers(0, 256, (F, 32), dtype=np.uint8)
    K = pnp.K_matrix(synth.FX, synth.FY, synth.CX, synth.CY)
    pc = np.stack([rng.uniform(-8, 8, F), rng.uniform(-2, 2, F), rng.uniform(4, 40, F)], axis=1)     # camera frame
    kp3d = (pc @ synth.R_LOCAL.T).astype(np.float32)                                                   # base frame
    rv_true, tv_true = np.array([0.02, -0.05, 0.01]), np.array([0.3, -0.1, 0.5])
    Rt, _ = cv2.Rodrigues(rv_true)
    q = pc @ Rt.T + tv_true
    uv = np.stack([q[:, 0] / q[:, 2] * synth.FX + synth.CX, q[:, 1] / q[:, 2] * synth.FY + synth.CY], axis=1)
    uv += rng.normal(0, 0.4, uv.shape)
    out = rng.random(F) < 0.35
    uv[out] = np.stack([rng.uniform(0, 1241, out.sum()), rng.uniform(0, 376, out.sum())], axis=1)
    kp2d = uv.astype(np.float32)
    ctx = capi.Context(0)
    try:
        fr, to = capi.Frames(ctx, 1, F), capi.Frames(ctx, 1, F)
        cnt = np.array([F], np.int32)
        fr.upload(0, desc[None], np.zeros((1, F, 2), np.float32), kp3d[None], cnt)
        to.upload(0, desc[None], kp2d[None], np.ones((1, F, 3), np.float32), cnt)
        prm = capi.pnp_params(synth.FX, synth.FY, synth.CX, synth.CY)
        res = capi.verify_pairs(ctx, fr, to, [[0, 0]], prm)
        assert res["n_matches"][0] == F
        pairs = matching.global_match(desc, desc, 0.8)
        obj, img, fi, ti = matching.pnp_correspondences(pairs, kp3d, kp2d)
        assert len(obj) == F
        rv0, tv0 = np.array(prm.guess_rvec), np.array(prm.guess_tvec)
        ok, rv, tv, inl = pnp.cv_solve_pnp_ransac(obj, img, K, rv0, tv0)
        assert ok and res["status"][0] == 1
        assert np.array_equal(capi.inlier_indices(res["inlier_mask"][0], F), inl)
        assert np.abs(res["pose"][0, 3:] - tv).max() < TRANS_TOL and np.abs(res["pose"][0, :3] - rv).max() < ROT_TOL
    finally:
        ctx.close()
