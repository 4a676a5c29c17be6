"""2-view bundle adjustment (SURVEY.md 8 row a8, R/src/myRegistrationVis.cpp:1192-1345) on the GPU, through the C ABI,
against oracle/ba.py (the restated RTAB-Map OptimizerG2O::optimizeBA / g2o algorithm; unpinned to the reference's
binaries, see the oracle's header) and the committed fixture tests/golden/ba.npz."""
import numpy as np
import pytest

from multi_robot_slam_separators_b200 import capi, synth
from oracle import ba
from tests import ba_problems as bp

pytestmark = pytest.mark.gpu

K = (bp.FX, bp.FY, bp.CX, bp.CY)
T_TOL = 1e-6        # transform entries (f32 output of an f64 optimisation; oracle and kernel sum in different orders)
CHI_TOL = 1e-6      # relative, final robust chi2


@pytest.fixture(scope="module")
def ctx():
    c = capi.Context(0)
    yield c
    c.close()


def upload(ctx, problems, F, rng, shuffle=True):
    """One (From, To) frame pair per problem; the problem's correspondences sit at random feature slots and in a random
    match-list order, padded with non-inlier matches (mask bit 0) that the kernel must ignore."""
    P = len(problems)
    kp2f, kp3f = np.zeros((P, F, 2), np.float32), np.ones((P, F, 3), np.float32)
    kp2t, kp3t = np.zeros((P, F, 2), np.float32), np.ones((P, F, 3), np.float32)
    desc = rng.integers(0, 256, (P, F, 32), dtype=np.uint8)
    cnt = np.full(P, F, np.int32)
    n_matches = np.zeros(P, np.int32)
    mfrom, mto = np.zeros((P, F), np.uint16), np.zeros((P, F), np.uint16)
    mask = np.zeros((P, (F + 31) // 32), np.uint32)
    order = []
    for p, pr in enumerate(problems):
        n = len(pr["kp3d_from"])
        M = min(F, n + int(rng.integers(0, 20)))
        sf = rng.permutation(F)[:M] if shuffle else np.arange(M)
        st = rng.permutation(F)[:M] if shuffle else np.arange(M)
        pos = np.sort(rng.permutation(M)[:n])            # match-list positions of the inliers, ascending = inlier order
        kp2f[p, sf[pos]], kp3f[p, sf[pos]] = pr["kp2d_from"], pr["kp3d_from"]
        kp2t[p, st[pos]], kp3t[p, st[pos]] = pr["kp2d_to"], pr["kp3d_to"]
        mfrom[p, :M], mto[p, :M], n_matches[p] = sf, st, M
        bits = np.zeros(mask.shape[1] * 32, np.uint8)
        bits[pos] = 1
        mask[p] = np.packbits(bits, bitorder="little").view(np.uint32)
        order.append(pos)
    fr, to = capi.Frames(ctx, P, F), capi.Frames(ctx, P, F)
    fr.upload(0, desc, kp2f, kp3f, cnt)
    to.upload(0, desc, kp2t, kp3t, cnt)
    res = dict(status=np.ones(P, np.int32), n_matches=n_matches, n_inliers=np.array([len(o) for o in order], np.int32),
               pose=np.zeros((P, 6)), transform=np.stack([pr["T0"].reshape(12) for pr in problems]).astype(np.float32),
               match_from=mfrom, match_to=mto, inlier_mask=mask,
               covariance=np.stack([pr["cov"].reshape(36) for pr in problems]))
    return fr, to, res, order


def pnp_prm(local_t=(0, 0, 0), min_inliers=5):
    prm = capi.pnp_params(*K, min_inliers=min_inliers)
    for i in range(3):
        prm.local_t[i] = float(local_t[i])
    return prm


def check_pair(out, p, pr, pos, M, min_inliers=5, **okw):
    ok, T, kept, res = ba.bundle_adjust_pair(pr["T0"], pr["kp2d_from"], pr["kp3d_from"], pr["kp2d_to"], pr["kp3d_to"], K,
                                             okw.pop("baseline", bp.BASELINE), bp.LOCAL_R, pr["local_t"],
                                             okw.pop("cov", pr["cov"]), min_inliers, **okw)
    got = capi.inlier_indices(out["inlier_mask"][p], M)
    assert res["ok"]
    assert np.array_equal(got, pos[kept]), (p, got, pos[kept])
    assert out["n_inliers"][p] == kept.sum()
    assert out["status"][p] == (1 if ok else 3), (p, out["status"][p])
    assert np.abs(out["transform"][p].reshape(3, 4) - T).max() <= T_TOL, (p, out["transform"][p].reshape(3, 4), T)
    assert abs(out["stats"][p, 1] - res["chi2"]) <= CHI_TOL * max(1.0, res["chi2"])
    # iterations: at convergence the stop (ten failed trials / zero gain) hangs on the last bits, so only the bound
    assert 1 <= out["stats"][p, 0] <= 5 + okw.get("iterations", 20) and out["stats"][p, 2] == (~kept).sum()
    if ok:
        # pose = (rvec, tvec) of From base -> To camera: consistent with the transform
        R = bp.rodrigues(out["pose"][p, :3])
        Tm = out["transform"][p].reshape(3, 4).astype(np.float64)
        Rc, tc = Tm[:, :3] @ bp.LOCAL_R, Tm[:, :3] @ pr["local_t"] + Tm[:, 3]
        assert np.abs(R - Rc.T).max() < 1e-6 and np.abs(out["pose"][p, 3:] + Rc.T @ tc).max() < 1e-5
    return ok, res


def test_fixture_and_seeded_problems_vs_oracle(ctx, golden):
    g = golden("ba")
    problems = []
    for i in range(int(g["n"])):
        problems.append({k: g[f"{k}_{i}"] for k in ("T0", "kp2d_from", "kp3d_from", "kp2d_to", "kp3d_to", "cov", "local_t")})
    rng = np.random.default_rng(31)
    for seed in range(100, 124):
        problems.append(bp.make(seed, int(rng.integers(6, 400)), n_out=int(rng.integers(0, 8)),
                                nan_frac=float(rng.uniform(0, 0.6))))
    # local_t is a parameter of the call: group the problems by it
    for lt in sorted({tuple(pr["local_t"]) for pr in problems}):
        group = [pr for pr in problems if tuple(pr["local_t"]) == lt]
        fr, to, res, order = upload(ctx, group, 512, np.random.default_rng(32))
        try:
            out = capi.bundle_adjust_pairs(ctx, fr, to, [[i, i] for i in range(len(group))], res, pnp_prm(lt),
                                           capi.ba_params(bp.BASELINE))
            n_ok = 0
            for p, pr in enumerate(group):
                ok, _ = check_pair(out, p, pr, order[p], res["n_matches"][p])
                n_ok += ok
            assert n_ok >= len(group) - 1
        finally:
            fr.close(); to.close()
    # the committed known answers themselves
    fr, to, res, order = upload(ctx, problems[:int(g["n"])][:5], 512, np.random.default_rng(33), shuffle=False)
    try:
        out = capi.bundle_adjust_pairs(ctx, fr, to, [[i, i] for i in range(5)], res, pnp_prm(), capi.ba_params(bp.BASELINE))
        for i in range(5):
            assert np.abs(out["transform"][i].reshape(3, 4) - g[f"T_{i}"]).max() <= T_TOL
            assert np.array_equal(capi.inlier_indices(out["inlier_mask"][i], res["n_matches"][i]), order[i][g[f"kept_{i}"]])
    finally:
        fr.close(); to.close()


@pytest.mark.parametrize("variant", ["no_kernel", "mono_only", "identity_information", "iterations_5", "variance_4"])
def test_parameters(ctx, variant):
    rng = np.random.default_rng(41)
    problems = [bp.make(200 + i, int(rng.integers(10, 250)), n_out=int(rng.integers(0, 6))) for i in range(10)]
    fr, to, res, order = upload(ctx, problems, 256, rng)
    b = capi.ba_params(bp.BASELINE)
    okw = {}
    if variant == "no_kernel":
        b.robust_kernel_delta, okw = 0.0, dict(robust_delta=0.0)
    elif variant == "mono_only":
        b.baseline, okw = 0.0, dict(baseline=0.0)
    elif variant == "identity_information":
        b.variance_ignored, okw = 1, dict(cov=None)
    elif variant == "iterations_5":
        b.iterations, okw = 5, dict(iterations=5)
    elif variant == "variance_4":
        b.pixel_variance, okw = 4.0, dict(pixel_variance=4.0)
    try:
        out = capi.bundle_adjust_pairs(ctx, fr, to, [[i, i] for i in range(10)], res, pnp_prm(), b,
                                       use_covariance=variant != "identity_information")
        for p, pr in enumerate(problems):
            check_pair(out, p, pr, order[p], res["n_matches"][p], **dict(okw))
    finally:
        fr.close(); to.close()


def test_few_inliers_left_and_untouched_pairs(ctx):
    rng = np.random.default_rng(42)
    problems = [bp.make(7, 8, n_out=2), bp.make(301, 50), bp.make(302, 50)]
    fr, to, res, order = upload(ctx, problems, 128, rng)
    res["status"][2] = 3                      # not a verified pair: the block must leave it alone
    before = {k: res[k][2].copy() for k in ("transform", "inlier_mask", "n_inliers", "pose")}
    try:
        out = capi.bundle_adjust_pairs(ctx, fr, to, [[0, 0], [1, 1], [2, 2]], res, pnp_prm(min_inliers=7), capi.ba_params(bp.BASELINE))
        ok, _ = check_pair(out, 0, problems[0], order[0], res["n_matches"][0], min_inliers=7)
        assert not ok and out["status"][0] == 3 and not out["transform"][0].any() and out["n_inliers"][0] == 6
        check_pair(out, 1, problems[1], order[1], res["n_matches"][1], min_inliers=7)
        assert out["status"][2] == 3
        for k, v in before.items():
            assert np.array_equal(out[k][2], v)
    finally:
        fr.close(); to.close()


def test_large_frames(ctx):
    """F = 2000 (BASELINE.json configs[3]): 150 KB of shared memory per pair."""
    rng = np.random.default_rng(43)
    problems = [bp.make(400 + i, n, n_out=10) for i, n in enumerate((1500, 900, 40))]
    fr, to, res, order = upload(ctx, problems, 2000, rng)
    try:
        out = capi.bundle_adjust_pairs(ctx, fr, to, [[i, i] for i in range(3)], res, pnp_prm(), capi.ba_params(bp.BASELINE))
        for p, pr in enumerate(problems):
            check_pair(out, p, pr, order[p], res["n_matches"][p])
    finally:
        fr.close(); to.close()


def test_inside_verify_pairs_and_the_two_call_driver(ctx):
    """sepb_ctx_set_bundle_adjustment: K2 -> K3 -> covariance -> bundle adjustment in one call equals the block applied
    to the same call's result without it; the two-call driver feeds the adjusted transform to the second call."""
    A, B = synth.make_robots(2, 36, n_feat=1000, dim=8, seed=321, overlap=0.5)
    fr, to = capi.frames_from_robot(ctx, A), capi.frames_from_robot(ctx, B)
    pairs = [[o, l] for l in range(36) for o in range(36) if A.place[o] == B.place[l]][:24]
    pairs += [[1, 0], [2, 5], [3, 7]]
    prm = capi.pnp_params(synth.FX, synth.FY, synth.CX, synth.CY)
    b = capi.ba_params(synth.BASELINE * synth.FY / synth.FX)
    try:
        plain = capi.verify_pairs(ctx, fr, to, pairs, prm)
        want = capi.bundle_adjust_pairs(ctx, fr, to, pairs, plain, prm, b)
        ctx.set_bundle_adjustment(b)
        fused = capi.verify_pairs(ctx, fr, to, pairs, prm)
        n_ok = int((plain["status"] == 1).sum())
        assert n_ok >= 5
        for k in ("status", "n_inliers", "transform", "inlier_mask", "pose"):
            assert np.array_equal(fused[k], want[k]), k
        assert np.array_equal(fused["covariance"], plain["covariance"])     # computed before the adjustment (:1077-1090)
        changed = sum(not np.array_equal(fused["transform"][p], plain["transform"][p]) for p in range(len(pairs))
                      if plain["status"][p] == 1)
        assert changed >= n_ok - 1
        # oracle on a verified pair, from the call's own outputs
        p = int(np.nonzero(plain["status"] == 1)[0][0])
        M = plain["n_matches"][p]
        inl = capi.inlier_indices(plain["inlier_mask"][p], M)
        i, j = pairs[p]
        f_idx, t_idx = plain["match_from"][p][inl], plain["match_to"][p][inl]
        ok, T, kept, res = ba.bundle_adjust_pair(plain["transform"][p], A.kp2d[i][f_idx], A.kp3d[i][f_idx], B.kp2d[j][t_idx],
                                                 B.kp3d[j][t_idx], (synth.FX, synth.FY, synth.CX, synth.CY), b.baseline,
                                                 bp.LOCAL_R, np.zeros(3), plain["covariance"][p].reshape(6, 6), 5)
        assert fused["status"][p] == (1 if ok else 3)
        assert np.array_equal(capi.inlier_indices(fused["inlier_mask"][p], M), inl[kept])
        assert np.abs(fused["transform"][p].reshape(3, 4) - T).max() <= T_TOL
        # two-call driver with the adjustment after each call
        gp = capi.guess_params(synth.WIDTH, synth.HEIGHT)
        two = capi.estimate_transformations(ctx, fr, to, pairs, prm, gp)
        ctx.set_bundle_adjustment(None)
        two_plain = capi.estimate_transformations(ctx, fr, to, pairs, prm, gp)
        assert np.array_equal(two["status_pass1"] == 1, fused["status"] == 1)
        both = (two["status"] == 1) & (two_plain["status"] == 1)
        assert both.sum() >= 5
        d = np.abs(two["transform"][both] - two_plain["transform"][both]).max(1)
        assert (d > 0).any() and np.median(d) < 0.05      # refined, not a different answer
    finally:
        ctx.set_bundle_adjustment(None)
        fr.close(); to.close()


def test_host_bound_stores_and_other_solver_options(ctx):
    """The adjustment needs the From pixels and the To 3-D points too: with host-bound stores (sepb_frames_bind_host_async)
    they are fetched before it runs -- same results as with whole uploads; and it composes with the P3P flags and the
    refine loop (it starts from whatever PnP produced)."""
    import torch
    A, B = synth.make_robots(2, 36, n_feat=1000, dim=8, seed=321, overlap=0.5)
    pin = lambda a: torch.from_numpy(np.ascontiguousarray(a)).pin_memory().numpy()  # noqa: E731
    hA = {k: pin(getattr(A, k)) for k in ("desc", "kp2d", "kp3d", "count", "octave")}
    hB = {k: pin(getattr(B, k)) for k in ("desc", "kp2d", "kp3d", "count", "octave")}
    pairs = [[o, l] for l in range(36) for o in range(36) if A.place[o] == B.place[l]][:20] + [[1, 0], [2, 5]]
    fa, fb = capi.frames_from_robot(ctx, A), capi.frames_from_robot(ctx, B)
    la, lb = capi.Frames(ctx, 36, 1000), capi.Frames(ctx, 36, 1000)
    b = capi.ba_params(synth.BASELINE * synth.FY / synth.FX)
    gp = capi.guess_params(synth.WIDTH, synth.HEIGHT)
    try:
        ctx.set_bundle_adjustment(b)
        for kw in (dict(), dict(flags=capi.PNP_P3P), dict(refine_iterations=2)):
            prm = capi.pnp_params(synth.FX, synth.FY, synth.CX, synth.CY, **kw)
            want = capi.verify_pairs(ctx, fa, fb, pairs, prm)
            want2 = capi.estimate_transformations(ctx, fa, fb, pairs, prm, gp)
            assert (want["status"] == 1).sum() >= 10
            la.upload(0, hA["desc"], hA["kp2d"], hA["kp3d"], hA["count"], hA["octave"], lazy=True)
            lb.upload(0, hB["desc"], hB["kp2d"], hB["kp3d"], hB["count"], hB["octave"], lazy=True)
            got = capi.verify_pairs(ctx, la, lb, pairs, prm)
            got2 = capi.estimate_transformations(ctx, la, lb, pairs, prm, gp)
            for g, w in ((got, want), (got2, want2)):
                for k in ("status", "n_matches", "n_inliers", "transform", "pose", "inlier_mask", "covariance"):
                    assert np.array_equal(g[k], w[k], equal_nan=True), (kw, k)
        # the adjustment really ran on top of each option: different from the same call without it
        ctx.set_bundle_adjustment(None)
        plain = capi.verify_pairs(ctx, fa, fb, pairs, prm)
        ok = (plain["status"] == 1) & (want["status"] == 1)
        assert ok.sum() >= 10 and (np.abs(plain["transform"][ok] - want["transform"][ok]).max(1) > 0).all()
    finally:
        ctx.set_bundle_adjustment(None)
        ctx.sync()
        for f in (fa, fb, la, lb):
            f.close()
