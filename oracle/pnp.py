"""TEST INFRASTRUCTURE ONLY (oracle) -- never imported by the product path.

CPU restatement of the reference's PnP step (R/src/myRegistrationVis.cpp:1055-1112):
RTAB-Map `util3d::estimateMotion3DTo2D` -> `util3d::solvePnPRansac` -> `cv::solvePnPRansac`
(all unvendored).  Two forms are kept side by side:

* `cv_solve_pnp_ransac`  -- the library call itself (cv2 4.13.0), the pinned arithmetic;
* `restated_solve_pnp_ransac` -- the explicit RANSAC loop (RNG stream, 5-point EPnP
  hypotheses, f32 reprojection error, `>` acceptance, adaptive iteration count, final
  ITERATIVE refit seeded by the caller's guess) exposing / accepting the minimal-sample
  indices.  tests/test_oracle_pnp.py checks it reproduces the library call exactly.
"""
from __future__ import annotations

import math

import numpy as np

try:
    import cv2
except Exception:  # pragma: no cover
    cv2 = None

from .cvjacobi import CvRNG
from . import cv_epnp

DBL_MIN = 2.2250738585072014e-308


def K_matrix(fx, fy, cx, cy):
    return np.array([[fx, 0.0, cx], [0.0, fy, cy], [0.0, 0.0, 1.0]], dtype=np.float64)


def cv_round(x: float) -> int:
    """cvRound: round-half-to-even (SSE cvtsd2si)."""
    return int(np.rint(x))


def ransac_update_num_iters(p, ep, model_points, max_iters):
    """cv::RANSACUpdateNumIters (modules/calib3d/src/ptsetreg.cpp)."""
    p = max(p, 0.0)
    p = min(p, 1.0)
    ep = max(ep, 0.0)
    ep = min(ep, 1.0)
    num = max(1.0 - p, DBL_MIN)
    denom = 1.0 - math.pow(1.0 - ep, model_points)
    if denom < DBL_MIN:
        return 0
    num = math.log(num)
    denom = math.log(denom)
    return max_iters if (denom >= 0 or -num >= max_iters * (-denom)) else cv_round(num / denom)


def sample_index_stream(count, model_points, n_hyp):
    """RANSACPointSetRegistrator::getSubset index stream for `n_hyp` iterations
    (RNG((uint64)-1); per slot redraw until distinct).  Returns int32 [n_hyp, model_points]."""
    rng = CvRNG(0xFFFFFFFFFFFFFFFF)
    out = np.zeros((n_hyp, model_points), dtype=np.int32)
    for h in range(n_hyp):
        idx = []
        for i in range(model_points):
            v = rng.next() % count
            while v in idx:
                v = rng.next() % count
            idx.append(v)
        out[h] = idx
    return out


def reproj_err_f32(obj, img, rvec, tvec, K):
    """PnPRansacCallback::computeError: projectPoints -> f32, err = dx*dx + dy*dy in f32."""
    proj, _ = cv2.projectPoints(obj, rvec, tvec, K, None)
    proj = proj.reshape(-1, 2).astype(np.float32)
    d = img.astype(np.float32) - proj
    return d[:, 0] * d[:, 0] + d[:, 1] * d[:, 1]


def cv_solve_pnp_ransac(obj, img, K, rvec0, tvec0, iterations=300, reproj_error=2.0, confidence=0.99,
                        flags=0):
    """The library call exactly as RTAB-Map issues it (useExtrinsicGuess=true)."""
    obj = np.ascontiguousarray(obj, dtype=np.float32).reshape(-1, 3)
    img = np.ascontiguousarray(img, dtype=np.float32).reshape(-1, 2)
    rvec = np.array(rvec0, dtype=np.float64).reshape(3, 1).copy()
    tvec = np.array(tvec0, dtype=np.float64).reshape(3, 1).copy()
    ok, rvec, tvec, inl = cv2.solvePnPRansac(obj, img, K, None, rvec, tvec, True, iterations,
                                             reproj_error, confidence, None, flags)
    inl = np.zeros(0, np.int32) if inl is None else inl.ravel().astype(np.int32)
    return bool(ok), rvec.ravel(), tvec.ravel(), inl


def restated_solve_pnp_ransac(obj, img, K, rvec0, tvec0, iterations=300, reproj_error=2.0,
                              confidence=0.99, sample_indices=None, hyp_solver="cv2", refit="cv2", flags=0):
    """Explicit-loop restatement.  flags=ITERATIVE (0) => 5-point EPnP minimal sets + ITERATIVE refit from the guess;
    flags=SOLVEPNP_P3P (2) / AP3P (5) => 4-point P3P minimal sets (a set without a solution is a skipped iteration)
    + EPNP refit over the inliers.

    Returns dict(ok, rvec, tvec, inliers, best_h, n_iters_run, samples, model)."""
    obj = np.ascontiguousarray(obj, dtype=np.float32).reshape(-1, 3)
    img = np.ascontiguousarray(img, dtype=np.float32).reshape(-1, 2)
    n = obj.shape[0]
    p3p = flags in (2, 5)
    mp = 4 if p3p else 5
    fx, fy, cx, cy = K[0, 0], K[1, 1], K[0, 2], K[1, 2]
    assert n > mp, "n == model_points is the single-solve path; n < model_points never reaches here"
    niters = max(int(iterations), 1)
    if sample_indices is None:
        sample_indices = sample_index_stream(n, mp, niters)
    thr = np.float32(reproj_error * reproj_error)
    best_mask, best_good, best_h, best_model = None, 0, -1, None
    it = 0
    while it < niters:
        s = sample_indices[it]
        if p3p:
            ok, rv, tv = cv2.solvePnP(obj[s], img[s], K, None, flags=cv2.SOLVEPNP_P3P)
            if not ok:
                it += 1
                continue
            rv, tv = rv.ravel(), tv.ravel()
        elif hyp_solver == "cv2":
            ok, rv, tv = cv2.solvePnP(obj[s], img[s], K, None, flags=cv2.SOLVEPNP_EPNP)
            rv, tv = rv.ravel(), tv.ravel()
        else:
            rv, tv, _ = cv_epnp.solve_epnp(obj[s], img[s], fx, fy, cx, cy)
            rv, tv = np.array(rv), np.array(tv)
        if np.all(np.isfinite(rv)) and np.all(np.isfinite(tv)):
            err = reproj_err_f32(obj, img, rv, tv, K)
            mask = err <= thr
        else:
            mask = np.zeros(n, bool)          # cv2's P3P hands back a NaN model on a few degenerate sets: no inlier
        good = int(mask.sum())
        if good > max(best_good, mp - 1):
            best_mask, best_good, best_h, best_model = mask, good, it, (rv, tv)
            niters = ransac_update_num_iters(confidence, (n - good) / n, mp, niters)
        it += 1
    if best_mask is None:
        return dict(ok=False, rvec=np.array(rvec0, float), tvec=np.array(tvec0, float),
                    inliers=np.zeros(0, np.int32), best_h=-1, n_iters_run=it, samples=sample_indices)
    inl = np.nonzero(best_mask)[0].astype(np.int32)
    o64 = obj[inl].astype(np.float64)
    i64 = img[inl].astype(np.float64)
    if p3p:
        ok, rvec, tvec = cv2.solvePnP(o64, i64, K, None, flags=cv2.SOLVEPNP_EPNP)
        rvec, tvec = rvec.ravel(), tvec.ravel()
    elif refit == "cv2":
        rvec = np.array(rvec0, dtype=np.float64).reshape(3, 1).copy()
        tvec = np.array(tvec0, dtype=np.float64).reshape(3, 1).copy()
        ok, rvec, tvec = cv2.solvePnP(o64, i64, K, None, rvec, tvec, True, cv2.SOLVEPNP_ITERATIVE)
        rvec, tvec = rvec.ravel(), tvec.ravel()
    else:
        rvec, tvec, _ = lm_refine(o64, i64, K, rvec0, tvec0)
    return dict(ok=True, rvec=rvec, tvec=tvec, inliers=inl, best_h=best_h, n_iters_run=it,
                samples=sample_indices, model=best_model)


def refine_model(obj, img, K, rvec, tvec, inliers, min_inliers, refine_iterations, reproj_error=2.0, refine_sigma=3.0):
    """RTAB-Map util3d::solvePnPRansac with refineIterations > 0 -- its clone of PCL's SampleConsensus::refineModel
    (unvendored; restated from the published source; reached from
    R/src/myRegistrationVis.cpp:1085 with _PnPRefineIterations parsed at :104).  All error arithmetic in f32 like the
    C++ (std::vector<float> err, uNorm, uMean, uVariance).  Returns (rvec, tvec, inliers)."""
    f32 = np.float32
    obj = np.ascontiguousarray(obj, dtype=np.float32).reshape(-1, 3)
    img = np.ascontiguousarray(img, dtype=np.float32).reshape(-1, 2)
    n = obj.shape[0]
    error_threshold = f32(reproj_error)
    it = 0
    inlier_changed = False
    prev_inliers = [int(i) for i in inliers]
    new_inliers = []
    inliers_sizes = []
    rv = np.array(rvec, np.float64).reshape(3, 1).copy()
    tv = np.array(tvec, np.float64).reshape(3, 1).copy()
    while True:
        cont = False
        _, rv, tv = cv2.solvePnP(obj[prev_inliers], img[prev_inliers], K, None, rv, tv, True, cv2.SOLVEPNP_ITERATIVE)
        inliers_sizes.append(len(prev_inliers))
        proj, _ = cv2.projectPoints(obj, rv, tv, K, None)
        proj = proj.reshape(-1, 2).astype(np.float32)
        dx, dy = img[:, 0] - proj[:, 0], img[:, 1] - proj[:, 1]
        err = np.sqrt(dx * dx + dy * dy, dtype=np.float32)                 # uNorm, f32
        new_inliers = [i for i in range(n) if err[i] < error_threshold]
        if len(new_inliers) < min_inliers:
            it += 1
            if it >= refine_iterations:
                break
            cont = True
        if not cont:
            s = f32(0)
            for e in err:
                s = f32(s + e)
            mean = f32(s / f32(n))
            s2 = f32(0)
            for e in err:
                d = f32(e - mean)
                s2 = f32(s2 + f32(d * d))
            variance = f32(s2 / f32(n - 1)) if n > 1 else f32(0)
            error_threshold = min(f32(reproj_error), f32(f32(refine_sigma) * np.sqrt(variance, dtype=np.float32)))
            inlier_changed = False
            prev_inliers, new_inliers = new_inliers, prev_inliers
            if len(new_inliers) != len(prev_inliers):
                if len(inliers_sizes) >= 4 and inliers_sizes[-1] == inliers_sizes[-3] and inliers_sizes[-2] == inliers_sizes[-4]:
                    break
                inlier_changed = True
            else:
                inlier_changed = any(a != b for a, b in zip(prev_inliers, new_inliers))
        if inlier_changed:
            it += 1
            if it < refine_iterations:
                continue
        break
    return rv.ravel(), tv.ravel(), np.array(new_inliers, np.int32)


# ---------------------------------------------------------------------------------------
# SOLVEPNP_ITERATIVE with an extrinsic guess: CvLevMarq on the reprojection residuals.
def _rodrigues_with_jac(r):
    rx, ry, rz = (float(x) for x in r)
    theta = math.sqrt(rx * rx + ry * ry + rz * rz)
    Jm = np.zeros((3, 9))
    if theta < 2.220446049250313e-16:
        R = np.eye(3)
        Jm[0, 5] = Jm[1, 6] = Jm[2, 1] = -1.0
        Jm[0, 7] = Jm[1, 2] = Jm[2, 3] = 1.0
        return R, Jm
    c, s = math.cos(theta), math.sin(theta)
    c1 = 1.0 - c
    it = 1.0 / theta
    r = np.array([rx, ry, rz]) * it
    rrt = np.outer(r, r)
    r_x = np.array([[0, -r[2], r[1]], [r[2], 0, -r[0]], [-r[1], r[0], 0]])
    R = c * np.eye(3) + c1 * rrt + s * r_x
    I = np.eye(3).ravel()
    drrt = np.array([[r[0] + r[0], r[1], r[2], r[1], 0, 0, r[2], 0, 0],
                     [0, r[0], 0, r[0], r[1] + r[1], r[2], 0, r[2], 0],
                     [0, 0, r[0], 0, 0, r[1], r[0], r[1], r[2] + r[2]]])
    d_r_x = np.array([[0, 0, 0, 0, 0, -1, 0, 1, 0],
                      [0, 0, 1, 0, 0, 0, -1, 0, 0],
                      [0, -1, 0, 1, 0, 0, 0, 0, 0]], dtype=np.float64)
    for i in range(3):
        ri = r[i]
        a0 = -s * ri
        a1 = (s - 2 * c1 * it) * ri
        a2 = c1 * it
        a3 = (c - s * it) * ri
        a4 = s * it
        Jm[i] = a0 * I + a1 * rrt.ravel() + a2 * drrt[i] + a3 * r_x.ravel() + a4 * d_r_x[i]
    return R, Jm


def _project_with_jac(obj, rvec, tvec, fx, fy, cx, cy, want_jac):
    R, dRdr = _rodrigues_with_jac(rvec)
    X = obj @ R.T + np.asarray(tvec).reshape(1, 3)
    z = np.where(X[:, 2] != 0, 1.0 / X[:, 2], 1.0)
    x = X[:, 0] * z
    y = X[:, 1] * z
    proj = np.stack([x * fx + cx, y * fy + cy], axis=1)
    if not want_jac:
        return proj, None
    n = obj.shape[0]
    Jac = np.zeros((2 * n, 6))
    # d/dt
    Jac[0::2, 3] = fx * z
    Jac[0::2, 5] = fx * (-x * z)
    Jac[1::2, 4] = fy * z
    Jac[1::2, 5] = fy * (-y * z)
    for j in range(3):
        dR = dRdr[j].reshape(3, 3)
        dX = obj @ dR.T
        Jac[0::2, j] = fx * z * (dX[:, 0] - x * dX[:, 2])
        Jac[1::2, j] = fy * z * (dX[:, 1] - y * dX[:, 2])
    return proj, Jac


def lm_refine(obj64, img64, K, rvec0, tvec0, max_iter=20, eps=1.1920928955078125e-07):
    """cvFindExtrinsicCameraParams2 with useExtrinsicGuess: CvLevMarq(6, 2n, 20 iters, FLT_EPSILON,
    completeSymm) over param=[rvec,tvec]; residual = proj - measured; DECOMP_SVD solves."""
    fx, fy, cx, cy = K[0, 0], K[1, 1], K[0, 2], K[1, 2]
    obj64 = np.asarray(obj64, dtype=np.float64).reshape(-1, 3)
    img64 = np.asarray(img64, dtype=np.float64).reshape(-1, 2)
    param = np.concatenate([np.asarray(rvec0, float).ravel(), np.asarray(tvec0, float).ravel()])
    prev_param = param.copy()
    lambda_lg10 = -3
    iters = 0
    state = "STARTED"
    prev_err_norm = 0.0
    JtJ = JtErr = None

    def residual(p, want_jac):
        proj, Jac = _project_with_jac(obj64, p[:3], p[3:], fx, fy, cx, cy, want_jac)
        return (proj - img64).ravel(), Jac

    def step():
        lam = math.exp(lambda_lg10 * math.log(10.0))
        A = JtJ.copy()
        A[np.diag_indices(6)] *= 1.0 + lam
        ok, x = cv2.solve(A, JtErr.reshape(6, 1), flags=cv2.DECOMP_SVD)
        return prev_param - x.ravel()

    err, Jac = residual(param, True)
    while True:
        if state == "STARTED":
            state = "CALC_J"
            err, Jac = residual(param, True)
            continue
        if state == "CALC_J":
            JtJ = Jac.T @ Jac
            JtErr = Jac.T @ err
            prev_param = param.copy()
            param = step()
            if iters == 0:
                prev_err_norm = float(np.linalg.norm(err))
            err, _ = residual(param, False)
            state = "CHECK_ERR"
            continue
        # CHECK_ERR
        err_norm = float(np.linalg.norm(err))
        if err_norm > prev_err_norm:
            lambda_lg10 += 1
            if lambda_lg10 <= 16:
                param = step()
                err, _ = residual(param, False)
                continue
        lambda_lg10 = max(lambda_lg10 - 1, -16)
        iters += 1
        rel = np.linalg.norm(param - prev_param) / (np.linalg.norm(prev_param) + 2.220446049250313e-16)
        if iters >= max_iter or rel < eps:
            break
        prev_err_norm = err_norm
        err, Jac = residual(param, True)
        state = "CALC_J"
    R, _ = _rodrigues_with_jac(param[:3])
    return param[:3].copy(), param[3:].copy(), R


def pose_covariance(T12, from3d, to3d, inliers):
    """Covariance as RTAB-Map's util3d::estimateMotion3DTo2D fills it (restated from the
    published source; reached from R/src/myRegistrationVis.cpp:1077-1090) and Registration clamps it
    (R/src/myRegistration.cpp:279-295).  T12: f32 3x4 transform (From-base <- To-base); from3d / to3d: [M,3] f32 points
    of the correspondences; inliers: indices into them.  All geometry in f32, like cv::Point3f / Eigen::Vector4f."""
    T = np.asarray(T12, np.float32).reshape(3, 4)
    cov = np.eye(6)
    d2s, angs = [], []
    f32 = np.float32
    for i in inliers:
        b = np.asarray(to3d[i], f32)
        if not np.all(np.isfinite(b)):
            continue
        a = np.asarray(from3d[i], f32)
        n = [f32(f32(f32(f32(T[r, 0] * b[0]) + f32(T[r, 1] * b[1])) + f32(T[r, 2] * b[2])) + T[r, 3]) for r in range(3)]
        d = [f32(a[k] - n[k]) for k in range(3)]
        d2s.append(f32(f32(f32(d[0] * d[0]) + f32(d[1] * d[1])) + f32(d[2] * d[2])))
        v1 = [f32(a[k] - T[k, 3]) for k in range(3)]
        v2 = [f32(n[k] - T[k, 3]) for k in range(3)]
        n1 = np.sqrt(f32(f32(f32(v1[0] * v1[0]) + f32(v1[1] * v1[1])) + f32(v1[2] * v1[2])), dtype=f32)
        n2 = np.sqrt(f32(f32(f32(v2[0] * v2[0]) + f32(v2[1] * v2[1])) + f32(v2[2] * v2[2])), dtype=f32)
        rad = f32(f32(f32(f32(v1[0] / n1) * f32(v2[0] / n2)) + f32(f32(v1[1] / n1) * f32(v2[1] / n2))) + f32(f32(v1[2] / n1) * f32(v2[2] / n2)))
        rad = min(max(rad, f32(-1)), f32(1))
        angs.append(np.arccos(rad, dtype=f32))
    if d2s:
        vl = 2.1981 * float(sorted(d2s)[len(d2s) >> 2])
        va = 2.1981 * float(sorted(angs)[len(angs) >> 2])
        for i in range(3):
            cov[i, i] = max(vl, 1e-9)
            cov[3 + i, 3 + i] = max(va, 1e-9)
    return cov
