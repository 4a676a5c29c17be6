"""CPU checks of oracle/ba.py, the restatement of the reference's 2-view bundle adjustment block
(R/src/myRegistrationVis.cpp:1192-1345 -> RTAB-Map OptimizerG2O::optimizeBA -> g2o).  g2o is not installed here, so
the oracle is held to what can be checked without it: its optimum against scipy's generic least-squares solver on the
same residuals, stationarity of the robust cost, the outlier rule, and the committed fixture tests/golden/ba.npz."""
import numpy as np
import pytest
from scipy.optimize import least_squares

from tests import ba_problems as bp
from oracle import ba

K = (bp.FX, bp.FY, bp.CX, bp.CY)


def problem(seed, n, **kw):
    pr = bp.make(seed, n, **kw)
    pb = ba.build_problem(pr["T0"], pr["kp2d_from"], pr["kp3d_from"], pr["kp2d_to"], pr["kp3d_to"], K, bp.BASELINE,
                          bp.LOCAL_R, pr["local_t"], pr["cov"])
    return pr, pb


def residuals(pb, x, n):
    cam = ba.cam_update(pb["cam2"], x[:6])
    pts = x[6:].reshape(n, 3)
    out = []
    for c, cm in enumerate((pb["cam1"], cam)):
        uvr, _, _ = ba._project(cm, pb["K"], pb["baseline"], pts)
        e = uvr - pb["meas"][c]
        e[:, 2] = np.where(pb["stereo"][c], e[:, 2], 0.0)
        e[pb["level"][c] != 0] = 0.0
        out.append(e)
    return out, ba._link_error(pb, cam) * np.sqrt(np.diag(pb["info"]))


@pytest.mark.parametrize("seed,n", [(11, 30), (12, 60)])
def test_plain_least_squares_optimum_matches_scipy(seed, n):
    """No kernel: Levenberg-Marquardt + Schur must land on the minimiser scipy finds for the same residual vector."""
    pr, pb = problem(seed, n, n_out=0)
    res = ba.optimize_ba(pb, iterations=60, robust_delta=0.0)
    assert res["ok"] and res["outliers"] == []

    def f(x):
        e, el = residuals(pb, x, n)
        return np.concatenate([e[0].ravel(), e[1].ravel(), el])

    ls = least_squares(f, np.concatenate([np.zeros(6), pb["pts"].ravel()]), method="lm", xtol=1e-15, ftol=1e-15, gtol=1e-15)
    assert abs(2 * ls.cost - res["chi2"]) <= 1e-7 * res["chi2"]
    cam = ba.cam_update(pb["cam2"], ls.x[:6])
    assert np.abs(cam[0] - res["cam2"][0]).max() < 1e-8 and np.abs(cam[1] - res["cam2"][1]).max() < 1e-6
    assert np.abs(ls.x[6:].reshape(n, 3) - res["pts"]).max() < 1e-4


def test_robust_cost_is_stationary_and_outliers_leave():
    pr, pb = problem(13, 80, n_out=6)
    res = ba.optimize_ba(pb)            # 5 + 20 iterations, Huber(8)
    assert res["ok"]
    assert set(pr["outliers"].tolist()) <= set(res["outliers"])        # the gross ones are all flagged
    assert len(res["outliers"]) <= len(pr["outliers"]) + 8              # chi2 > 8 also takes the tail of the noise
    n = 80
    pb2 = dict(pb)
    pb2["cam2"] = res["cam2"]          # increments are taken around the solution

    def cost2(x):
        e, el = residuals(pb2, x, n)
        total = float(el @ el)
        for c in range(2):
            rho, _ = ba._huber((e[c] * e[c]).sum(1), 8.0)
            total += float(rho.sum())
        return total

    # numerical gradient of the robust cost over camera 2 (6) and a few point coordinates, at the solution
    x0 = np.concatenate([np.zeros(6), res["pts"].ravel()])
    c0 = cost2(x0)
    assert abs(c0 - res["chi2"]) <= 1e-9 * max(1.0, res["chi2"]) or c0 <= res["chi2"]   # chi2 was taken before the last marking
    for d in list(range(6)) + [6, 7, 8, 6 + 3 * 40, 6 + 3 * 79 + 2]:
        h = 1e-6 if d < 6 else 1e-5
        xp, xm = x0.copy(), x0.copy()
        xp[d] += h
        xm[d] -= h
        g = (cost2(xp) - cost2(xm)) / (2 * h)
        curv = (cost2(xp) - 2 * c0 + cost2(xm)) / (h * h)
        assert abs(g) <= 2e-3 * max(1.0, abs(curv)) * (1e-2 if d < 6 else 1.0) + 1e-4, (d, g, curv)


def test_improves_the_pose_and_is_deterministic():
    better = 0
    for seed, n, kw in bp.default_set():
        pr = bp.make(seed, n, **kw)
        args = (pr["T0"], pr["kp2d_from"], pr["kp3d_from"], pr["kp2d_to"], pr["kp3d_to"], K, bp.BASELINE, bp.LOCAL_R,
                pr["local_t"], pr["cov"], 5)
        ok, T, kept, res = ba.bundle_adjust_pair(*args)
        ok2, T2, kept2, _ = ba.bundle_adjust_pair(*args)
        assert ok and ok2 and np.array_equal(T, T2) and np.array_equal(kept, kept2)
        better += np.abs(T - pr["T_true"]).max() < np.abs(pr["T0"] - pr["T_true"]).max()
    assert better >= len(bp.default_set()) - 1


def test_too_few_inliers_after_the_outliers_left():
    pr = bp.make(7, 8, n_out=2)
    ok, T, kept, res = ba.bundle_adjust_pair(pr["T0"], pr["kp2d_from"], pr["kp3d_from"], pr["kp2d_to"], pr["kp3d_to"], K,
                                             bp.BASELINE, bp.LOCAL_R, pr["local_t"], pr["cov"], 7)
    assert res["ok"] and not ok and kept.sum() == 6 and not T.any()      # :1340-1345: transform nulled


def test_fixture(golden):
    g = golden("ba")
    for i in range(int(g["n"])):
        ok, T, kept, res = ba.bundle_adjust_pair(g[f"T0_{i}"], g[f"kp2d_from_{i}"], g[f"kp3d_from_{i}"], g[f"kp2d_to_{i}"],
                                                 g[f"kp3d_to_{i}"], K, bp.BASELINE, bp.LOCAL_R, g[f"local_t_{i}"],
                                                 g[f"cov_{i}"], 5)
        assert ok == bool(g[f"ok_{i}"]) and np.array_equal(kept, g[f"kept_{i}"])
        assert np.abs(T - g[f"T_{i}"]).max() <= 1e-6
        assert abs(res["chi2"] - float(g[f"chi2_{i}"])) <= 1e-6 * float(g[f"chi2_{i}"])


# ---- the product's arithmetic (csrc/ba_math.cuh, host build) against this oracle ---------------------------------------

def _harness():
    import ctypes as C
    from tests import host_harness
    H = host_harness.load()
    dp = lambda a: a.ctypes.data_as(C.c_void_p)  # noqa: E731
    H.hh_ba_point_system.restype = C.c_double
    H.hh_ba_point_system.argtypes = [C.c_void_p, C.c_double, C.c_double, C.c_double] + [C.c_void_p] * 6 + [C.c_uint] + [C.c_void_p] * 5
    H.hh_ba_cholesky6.argtypes = [C.c_void_p, C.c_double, C.c_void_p, C.c_void_p]
    return H, dp


def test_header_point_block_matches_the_oracles_linearisation():
    """ba_math.cuh:point_system (what the kernel accumulates per inlier) == oracle _linearize restricted to that point,
    stereo and mono edges, edges that left (level 1), Huber on and off."""
    H, dp = _harness()
    pr, pb = problem(21, 40, n_out=5, nan_frac=0.3)
    K4 = np.array(K, np.float64)
    cam2 = ba.cam_update(pb["cam2"], np.array([0.01, -0.02, 0.015, 1e-3, -2e-3, 1e-3]))
    pts = pb["pts"] + 0.01
    pb["level"][0, 3] = 1
    pb["level"][1, 7] = 1
    for delta in (8.0, 0.0):
        Vs, bls, Ws, A_or, bc_or, chi_or = ba._linearize(dict(pb, info=np.zeros((6, 6))), cam2, pts, 1.5, delta)
        A_sum, bc_sum, chi_sum = np.zeros(21), np.zeros(6), 0.0
        for i in range(40):
            meas = np.concatenate([pb["meas"][0, i], pb["meas"][1, i]]).astype(np.float32)
            flags = int(pb["level"][0, i]) | int(pb["level"][1, i]) << 1 | int(pb["stereo"][0, i]) << 2 | int(pb["stereo"][1, i]) << 3
            V, bl, W, A, bc = np.zeros(6), np.zeros(3), np.zeros(18), np.zeros(21), np.zeros(6)
            X = np.ascontiguousarray(pts[i])
            chi_sum += H.hh_ba_point_system(dp(K4), bp.BASELINE, 1.5, delta, dp(pb["cam1"][0]), dp(pb["cam1"][1]),
                                            dp(np.ascontiguousarray(cam2[0])), dp(np.ascontiguousarray(cam2[1])), dp(X),
                                            dp(meas), flags, dp(V), dp(bl), dp(W), dp(A), dp(bc))
            Vf = np.array([[V[0], V[1], V[2]], [V[1], V[3], V[4]], [V[2], V[4], V[5]]])
            scale = max(1.0, np.abs(Vs[i]).max())
            assert np.abs(Vf - Vs[i]).max() <= 1e-9 * scale and np.abs(bl - bls[i]).max() <= 1e-9 * scale
            assert np.abs(W.reshape(6, 3) - Ws[i]).max() <= 1e-9 * max(1.0, np.abs(Ws[i]).max())
            A_sum += A
            bc_sum += bc
        iu = np.triu_indices(6)
        # the oracle's camera block also holds the link edge: its information was zeroed above, so only its chi2 (0) remains
        assert np.abs(A_sum - A_or[iu]).max() <= 1e-9 * np.abs(A_or).max() and np.abs(bc_sum - bc_or).max() <= 1e-9 * np.abs(bc_or).max()
        assert abs(chi_sum - chi_or) <= 1e-9 * chi_or


def test_header_pose_algebra_and_small_solves():
    import ctypes as C
    H, dp = _harness()
    rng = np.random.default_rng(5)
    for _ in range(20):
        qa = ba.quat_normalize_rotation(rng.normal(size=4))
        qb = ba.quat_normalize_rotation(rng.normal(size=4))
        ta, tb = rng.normal(size=3), rng.normal(size=3)
        upd = rng.normal(0, 0.05, 6)
        mul7, inv7, upd7, link6 = np.zeros(7), np.zeros(7), np.zeros(7), np.zeros(6)
        H.hh_ba_se3(dp(qa), dp(ta), dp(qb), dp(tb), dp(upd), dp(mul7), dp(inv7), dp(upd7), dp(link6))
        m, iv, u = ba.se3_mul((qa, ta), (qb, tb)), ba.se3_inv((qa, ta)), ba.cam_update((qa, ta), upd)
        for got, want in ((mul7, m), (inv7, iv), (upd7, u)):
            assert np.abs(got[:4] - want[0]).max() < 1e-14 and np.abs(got[4:] - want[1]).max() < 1e-13
        d = ba.se3_mul(iv, ba.se3_mul(iv, (qb, tb)))
        assert np.abs(link6 - np.concatenate([d[1], d[0][:3]])).max() < 1e-13
        # rotation matrix <-> quaternion, both branches of Eigen's conversion (trace > 0 and each dominant diagonal)
        R = ba.quat_to_matrix(qa)
        q4, Rb = np.zeros(4), np.zeros(9)
        H.hh_ba_quat(dp(np.ascontiguousarray(R.ravel())), dp(q4), dp(Rb))
        assert np.abs(q4 - ba.quat_normalize_rotation(ba.quat_from_matrix(R))).max() < 1e-14 and np.abs(Rb.reshape(3, 3) - R).max() < 1e-14
        # 6 x 6 Cholesky (one reciprocal per column) against numpy
        M = rng.normal(size=(6, 6))
        S = M @ M.T + 0.1 * np.eye(6)
        rhs, x = rng.normal(size=6), np.zeros(6)
        assert H.hh_ba_cholesky6(dp(np.ascontiguousarray(S[np.triu_indices(6)])), 0.3, dp(rhs), dp(x)) == 1
        assert np.abs(x - np.linalg.solve(S + 0.3 * np.eye(6), rhs)).max() < 1e-10
    S = -np.eye(6)
    assert H.hh_ba_cholesky6(dp(np.ascontiguousarray(S[np.triu_indices(6)])), 0.0, dp(np.ones(6)), dp(np.zeros(6))) == 0
