"""K3 parity (GPU, through the C ABI) for the solver options beyond the launch default: Vis/PnPFlags = P3P / AP3P
(4-point minimal sets + EPnP refit), exactly four correspondences (the single P3P solve), the OpenCV 3.2 return
convention (minimal-set model) and RTAB-Map's Vis/PnPRefineIterations loop -- against cv2.solvePnPRansac /
cv2.solvePnP and the restatements in oracle/pnp.py.  R/src/myRegistrationVis.cpp:103-104,1084-1085."""
import numpy as np
import pytest

from multi_robot_slam_separators_b200 import capi
from oracle import pnp

cv2 = pytest.importorskip("cv2")
pytestmark = pytest.mark.gpu

ROT_TOL = 1e-4   # rad   (north_star)
TRANS_TOL = 1e-4  # m
FX, FY, CX, CY = 500.0, 510.0, 320.0, 240.0
F = 512


@pytest.fixture(scope="module")
def ctx():
    c = capi.Context(0)
    yield c
    c.close()


def rot_angle(rv_a, rv_b):
    Ra, _ = cv2.Rodrigues(np.asarray(rv_a, float))
    Rb, _ = cv2.Rodrigues(np.asarray(rv_b, float))
    return float(np.arccos(np.clip((np.trace(Ra.T @ Rb) - 1) / 2, -1, 1)))


def make_problem(rng, n, outlier_frac, noise=0.4):
    K = pnp.K_matrix(FX, FY, CX, CY)
    obj = rng.uniform(-2, 2, (n, 3)).astype(np.float32)
    obj[:, 2] += 6
    rv, tv = rng.normal(0, 0.3, 3), rng.normal(0, 0.5, 3)
    img, _ = cv2.projectPoints(obj.astype(np.float64), rv, tv, K, None)
    img = img.reshape(-1, 2).astype(np.float32)
    img += rng.normal(0, noise, (n, 2)).astype(np.float32)
    out = int(n * outlier_frac)
    img[:out] += rng.uniform(-60, 60, (out, 2)).astype(np.float32)
    return obj, img


def upload_problems(ctx, problems, seed):
    """One (From, To) frame pair per problem: identical random descriptors on both sides make the match list the
    identity, so correspondence m is (kp3d_from[m], kp2d_to[m]) -- the PnP input is exactly (obj, img)."""
    rng = np.random.default_rng(seed)
    n = len(problems)
    desc = rng.integers(0, 256, (n, F, 32), dtype=np.uint8)
    kp3 = np.ones((n, F, 3), np.float32)
    kp2 = np.zeros((n, F, 2), np.float32)
    cnt = np.zeros(n, np.int32)
    for i, (obj, img) in enumerate(problems):
        m = len(obj)
        kp3[i, :m], kp2[i, :m], cnt[i] = obj, img, m
    fr, to = capi.Frames(ctx, n, F), capi.Frames(ctx, n, F)
    fr.upload(0, desc, np.zeros_like(kp2), kp3, cnt)
    to.upload(0, desc, kp2, np.ones_like(kp3), cnt)
    return fr, to


def check_against(res, p, ok, rvec, tvec, inliers, M, min_inl):
    assert res["n_matches"][p] == M
    want = (1 if len(inliers) >= min_inl else 3) if ok else 2
    assert res["status"][p] == want, (p, res["status"][p], want)
    if ok:
        got = capi.inlier_indices(res["inlier_mask"][p], M)
        assert np.array_equal(got, inliers), (p, got, inliers)
        assert res["n_inliers"][p] == len(inliers)
        assert rot_angle(res["pose"][p, :3], rvec) <= ROT_TOL, p
        assert np.abs(res["pose"][p, 3:] - tvec).max() <= TRANS_TOL, p


@pytest.mark.parametrize("flags", [capi.PNP_P3P, capi.PNP_AP3P])
def test_p3p_flags_vs_cv2(ctx, flags):
    rng = np.random.default_rng(21)
    sizes = [5, 6, 7, 9, 12] + [int(rng.integers(20, 400)) for _ in range(27)]
    problems = [make_problem(rng, n, float(rng.uniform(0.1, 0.6)) if n > 8 else 0.0) for n in sizes]
    fr, to = upload_problems(ctx, problems, 1)
    K = pnp.K_matrix(FX, FY, CX, CY)
    pairs = [[i, i] for i in range(len(problems))]
    for iters in (300, 50):
        prm = capi.pnp_params(FX, FY, CX, CY, iterations=iters, flags=flags)
        rv0, tv0 = np.array(prm.guess_rvec), np.array(prm.guess_tvec)
        for sched in ("persistent", "per_pair"):
            ctx.set_pnp_scheduler(sched)
            res = capi.verify_pairs(ctx, fr, to, pairs, prm)
            ctx.set_pnp_scheduler("persistent")
            n_ok = 0
            for p, (obj, img) in enumerate(problems):
                ok, rv, tv, inl = pnp.cv_solve_pnp_ransac(obj, img, K, rv0, tv0, iterations=iters, flags=cv2.SOLVEPNP_P3P)
                check_against(res, p, ok, rv, tv, inl, len(obj), 5)
                n_ok += ok
            assert n_ok >= len(problems) - 3
    fr.close(); to.close()


def test_p3p_flags_with_the_references_own_samples(ctx):
    """north_star: fed the reference's own minimal-sample indices, inlier sets agree exactly."""
    rng = np.random.default_rng(22)
    problems = [make_problem(rng, int(rng.integers(30, 300)), float(rng.uniform(0.1, 0.5))) for _ in range(12)]
    fr, to = upload_problems(ctx, problems, 2)
    K = pnp.K_matrix(FX, FY, CX, CY)
    iters = 200
    prm = capi.pnp_params(FX, FY, CX, CY, iterations=iters, flags=capi.PNP_P3P)
    rv0, tv0 = np.array(prm.guess_rvec), np.array(prm.guess_tvec)
    cs = np.zeros((len(problems), iters, 4), np.int32)
    refs = []
    for i, (obj, img) in enumerate(problems):
        sam = np.stack([rng.permutation(len(obj))[:4] for _ in range(iters)]).astype(np.int32)   # NOT the cv::RNG stream
        cs[i] = sam
        refs.append(pnp.restated_solve_pnp_ransac(obj, img, K, rv0, tv0, iterations=iters, sample_indices=sam, flags=2))
    res = capi.verify_pairs(ctx, fr, to, [[i, i] for i in range(len(problems))], prm, custom_samples=cs)
    for p, (r, (obj, _)) in enumerate(zip(refs, problems)):
        check_against(res, p, r["ok"], r["rvec"], r["tvec"], r["inliers"], len(obj), 5)
        if r["ok"]:
            assert res["best_hyp"][p] == r["best_h"] and res["iters_run"][p] == r["n_iters_run"]
    fr.close(); to.close()


@pytest.mark.parametrize("flags", [capi.PNP_ITERATIVE, capi.PNP_P3P])
def test_exactly_four_correspondences(ctx, flags):
    """npoints == 4: cv::solvePnPRansac answers with ONE P3P solve over the four points, all of them inliers
    (separators_min_inliers = 4 is the only setting that lets such a pair through)."""
    rng = np.random.default_rng(23)
    problems = [make_problem(rng, 4, 0.0, noise=float(rng.uniform(0, 1))) for _ in range(40)]
    fr, to = upload_problems(ctx, problems, 3)
    K = pnp.K_matrix(FX, FY, CX, CY)
    prm = capi.pnp_params(FX, FY, CX, CY, min_inliers=4, flags=flags)
    rv0, tv0 = np.array(prm.guess_rvec), np.array(prm.guess_tvec)
    res = capi.verify_pairs(ctx, fr, to, [[i, i] for i in range(40)], prm)
    n_ok = 0
    for p, (obj, img) in enumerate(problems):
        ok, rv, tv, inl = pnp.cv_solve_pnp_ransac(obj, img, K, rv0, tv0)
        if ok and not np.all(np.isfinite(tv)):
            ok = False          # cv2 4.13 reports success with a NaN translation on a few degenerate sets
        check_against(res, p, ok, rv, tv, inl, 4, 4)
        n_ok += ok
    assert n_ok >= 36
    # the launch default (min_inliers = 5) never accepts such a pair: no RANSAC, null transform
    res5 = capi.verify_pairs(ctx, fr, to, [[i, i] for i in range(40)], capi.pnp_params(FX, FY, CX, CY, flags=flags))
    assert np.all(res5["status"] == 0) and np.all(res5["transform"] == 0)
    fr.close(); to.close()


@pytest.mark.parametrize("flags", [capi.PNP_ITERATIVE, capi.PNP_P3P])
def test_minimal_model_is_opencv_3_2_semantics(ctx, flags):
    """sepb_pnp_params.minimal_model = 1: the pose returned is the winning hypothesis itself (what the OpenCV 3.2 of
    the reference image hands back), same inlier set."""
    rng = np.random.default_rng(24)
    problems = [make_problem(rng, int(rng.integers(20, 300)), float(rng.uniform(0.1, 0.5))) for _ in range(16)]
    fr, to = upload_problems(ctx, problems, 4)
    K = pnp.K_matrix(FX, FY, CX, CY)
    prm = capi.pnp_params(FX, FY, CX, CY, flags=flags, minimal_model=True)
    rv0, tv0 = np.array(prm.guess_rvec), np.array(prm.guess_tvec)
    res = capi.verify_pairs(ctx, fr, to, [[i, i] for i in range(16)], prm)
    for p, (obj, img) in enumerate(problems):
        r = pnp.restated_solve_pnp_ransac(obj, img, K, rv0, tv0, flags=flags)
        if r["ok"]:
            r["rvec"], r["tvec"] = r["model"]
        check_against(res, p, r["ok"], r["rvec"], r["tvec"], r["inliers"], len(obj), 5)
        if r["ok"]:
            assert np.abs(res["pose"][p, :3] - r["rvec"]).max() < 1e-8 and np.abs(res["pose"][p, 3:] - r["tvec"]).max() < 1e-8
    fr.close(); to.close()


@pytest.mark.parametrize("rounds", [1, 5])
def test_refine_iterations_loop(ctx, rounds):
    """Vis/PnPRefineIterations > 0: RTAB-Map's refineModel loop after RANSAC (oracle/pnp.py:refine_model on top of
    cv2.solvePnPRansac)."""
    rng = np.random.default_rng(25)
    problems = [make_problem(rng, int(rng.integers(20, 300)), float(rng.uniform(0.1, 0.5)), noise=float(rng.uniform(0.2, 0.9)))
                for _ in range(20)]
    fr, to = upload_problems(ctx, problems, 5)
    K = pnp.K_matrix(FX, FY, CX, CY)
    prm = capi.pnp_params(FX, FY, CX, CY, refine_iterations=rounds)
    rv0, tv0 = np.array(prm.guess_rvec), np.array(prm.guess_tvec)
    res = capi.verify_pairs(ctx, fr, to, [[i, i] for i in range(20)], prm)
    base = capi.verify_pairs(ctx, fr, to, [[i, i] for i in range(20)], capi.pnp_params(FX, FY, CX, CY))
    n_changed = 0
    for p, (obj, img) in enumerate(problems):
        ok, rv, tv, inl = pnp.cv_solve_pnp_ransac(obj, img, K, rv0, tv0)
        if ok and len(inl) >= 5:
            rv, tv, inl = pnp.refine_model(obj, img, K, rv, tv, inl, 5, rounds)
        check_against(res, p, ok, rv, tv, inl, len(obj), 5)
        n_changed += int(not np.array_equal(res["inlier_mask"][p], base["inlier_mask"][p]))
    assert n_changed >= 3        # the loop does something on this data
    fr.close(); to.close()
