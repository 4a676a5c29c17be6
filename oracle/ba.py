"""TEST INFRASTRUCTURE ONLY (oracle) -- never imported by the product path.

CPU restatement of the 2-view bundle adjustment block of the reference, R/src/myRegistrationVis.cpp:1192-1345:
`Optimizer::create(kTypeG2O)->optimizeBA(1, poses, links, models, points3DMap, wordReferences, &sbaOutliers)`.

PARITY UNPINNED: the optimiser lives in RTAB-Map (`OptimizerG2O::optimizeBA`) on top of g2o, neither of which is under
/root/reference nor installed here, so nothing in this file was checked against the reference's binaries.  It restates
the published algorithm [from memory of RTAB-Map 0.17-0.20 and g2o]:

* vertices: `VertexCam` 1 = From camera (pose `localTransform`, FIXED: rootId = 1), `VertexCam` 2 = To camera
  (pose `transforms[0] * localTransform`), both with `setKcam(fx, fy, cx, cy, baseline)`; one marginalised
  `VertexSBAPointXYZ` per inlier, initialised with the From frame's 3-D point (From base frame = world);
* edges: per inlier and camera an `EdgeProjectP2SC` (measurement `(u, v, u - disparity)`, disparity =
  baseline * fx / depth, when the key point's depth in that camera is finite and > 0) or an `EdgeProjectP2MC`
  (measurement `(u, v)`), information I / g2o/PixelVariance, `RobustKernelHuber(g2o/RobustKernelDelta)`;
  one `EdgeSBACam` 1 -> 2 with measurement `local^-1 * transforms[0] * local` and information covariance^-1;
* `OptimizationAlgorithmLevenberg(BlockSolver_6_3)`: lambda_0 = 1e-5 * max |diag H|, gain ratio
  rho = (chi2 - chi2_new) / (dx . (lambda dx + b) + 1e-3), good step: lambda *= max(1/3, min(1 - (2 rho - 1)^3, 2/3)),
  bad step: lambda *= ni, ni *= 2, at most 10 trials; Schur complement over the points;
* schedule (robust kernel on): 5 iterations, every edge whose chi2 exceeds the kernel's delta (NOT delta^2: the
  reference compares chi2 with `robustKernel()->delta()`) goes to level 1 and its word id to `sbaOutliers`,
  `initializeOptimization(0)`, Optimizer/Iterations (20) more iterations, same marking again;
* failure (null transform): NaN / > 1e12 robust chi2.

Afterwards the caller drops the outlier ids from the inliers, fails the pair when fewer than Vis/MinInliers remain
(:1340-1345) and otherwise takes camera 2's pose * localTransform^-1 as the transform (:1348).
"""
from __future__ import annotations

import numpy as np

HUGE = 1.7976931348623157e308


# ---- Eigen / g2o::SE3Quat arithmetic (quaternions as (x, y, z, w)) ------------------------------------------------

def quat_from_matrix(m):
    """Eigen::Quaterniond(Matrix3d)."""
    m = np.asarray(m, np.float64)
    q = np.zeros(4)
    t = m[0, 0] + m[1, 1] + m[2, 2]
    if t > 0:
        t = np.sqrt(t + 1.0)
        q[3] = 0.5 * t
        t = 0.5 / t
        q[0] = (m[2, 1] - m[1, 2]) * t
        q[1] = (m[0, 2] - m[2, 0]) * t
        q[2] = (m[1, 0] - m[0, 1]) * t
    else:
        i = 0
        if m[1, 1] > m[0, 0]:
            i = 1
        if m[2, 2] > m[i, i]:
            i = 2
        j = (i + 1) % 3
        k = (j + 1) % 3
        t = np.sqrt(m[i, i] - m[j, j] - m[k, k] + 1.0)
        q[i] = 0.5 * t
        t = 0.5 / t
        q[3] = (m[k, j] - m[j, k]) * t
        q[j] = (m[j, i] + m[i, j]) * t
        q[k] = (m[k, i] + m[i, k]) * t
    return q


def quat_normalize_rotation(q):
    """SE3Quat::normalizeRotation: w >= 0, unit norm."""
    q = np.array(q, np.float64)
    if q[3] < 0:
        q = -q
    return q / np.sqrt(q @ q)


def quat_to_matrix(q):
    """Eigen::Quaterniond::toRotationMatrix."""
    x, y, z, w = q
    tx, ty, tz = 2 * x, 2 * y, 2 * z
    twx, twy, twz = tx * w, ty * w, tz * w
    txx, txy, txz = tx * x, ty * x, tz * x
    tyy, tyz, tzz = ty * y, tz * y, tz * z
    return np.array([[1 - (tyy + tzz), txy - twz, txz + twy],
                     [txy + twz, 1 - (txx + tzz), tyz - twx],
                     [txz - twy, tyz + twx, 1 - (txx + tyy)]])


def quat_mul(a, b):
    ax, ay, az, aw = a
    bx, by, bz, bw = b
    return np.array([aw * bx + ax * bw + ay * bz - az * by,
                     aw * by + ay * bw + az * bx - ax * bz,
                     aw * bz + az * bw + ax * by - ay * bx,
                     aw * bw - ax * bx - ay * by - az * bz])


def quat_rotate(q, v):
    return quat_to_matrix(q) @ v


def se3_mul(a, b):
    """SE3Quat::operator* -- (q, t) pairs."""
    return quat_normalize_rotation(quat_mul(a[0], b[0])), a[1] + quat_rotate(a[0], b[1])


def se3_inv(a):
    qi = np.array([-a[0][0], -a[0][1], -a[0][2], a[0][3]])
    return qi, quat_rotate(qi, -a[1])


def cam_update(cam, upd):
    """SBACam::update: t += upd[:3]; r *= (upd[3:], sqrt(1 - |upd[3:]|^2)) normalised."""
    q, t = cam
    v = np.asarray(upd[3:6], np.float64)
    w2 = 1.0 - v @ v
    qr = np.array([v[0], v[1], v[2], np.sqrt(w2) if w2 > 0 else np.nan])
    qr = qr / np.sqrt(qr @ qr)
    qn = quat_mul(q, qr)
    return qn / np.sqrt(qn @ qn), t + np.asarray(upd[:3], np.float64)


# ---- the problem as RTAB-Map builds it (myRegistrationVis.cpp:1203-1318) ------------------------------------------

def depth_in_camera(p3, local_R, local_t):
    """util3d::transformPoint(pt, localTransform^-1).z, f32 like rtabmap::Transform."""
    Ri = np.asarray(local_R, np.float32).T
    ti = -(Ri @ np.asarray(local_t, np.float32))
    p = np.asarray(p3, np.float32)
    return (p @ Ri.T + ti)[..., 2].astype(np.float32)


def build_problem(T12, from_kp2d, from_kp3d, to_kp2d, to_kp3d, K, baseline, local_R, local_t, covariance=None):
    """Inputs are per inlier, in inlier order: From / To pixels [N, 2] f32, From / To 3-D points [N, 3] f32 (base
    frames; To points may be NaN), T12 the 3x4 f32 transform found by PnP, `baseline` the one VertexCam gets
    (stereo baseline * fy / fx, :1247-1252).  Returns a dict the optimiser works on."""
    fx, fy, cx, cy = K
    local_R = np.asarray(local_R, np.float64).reshape(3, 3)
    local_t = np.asarray(local_t, np.float64)
    T = np.asarray(T12, np.float32).reshape(3, 4).astype(np.float64)
    n = len(from_kp3d)
    pts = np.asarray(from_kp3d, np.float32).astype(np.float64)
    meas = np.zeros((2, n, 3))
    stereo = np.zeros((2, n), bool)
    for c, (k2, k3) in enumerate(((from_kp2d, from_kp3d), (to_kp2d, to_kp3d))):
        k2 = np.asarray(k2, np.float32)
        depth = depth_in_camera(k3, local_R, local_t).astype(np.float64)
        ok = np.isfinite(depth) & (depth > 0) & (baseline > 0)
        with np.errstate(all="ignore"):
            disparity = (baseline * fx / depth).astype(np.float32)
        meas[c, :, 0] = k2[:, 0]
        meas[c, :, 1] = k2[:, 1]
        meas[c, :, 2] = np.where(ok, (k2[:, 0] - disparity).astype(np.float64), 0.0)
        stereo[c] = ok
    cam1 = (quat_normalize_rotation(quat_from_matrix(local_R)), local_t.copy())
    R2 = T[:, :3] @ local_R
    t2 = T[:, :3] @ local_t + T[:, 3]
    cam2 = (quat_normalize_rotation(quat_from_matrix(R2)), t2)
    # EdgeSBACam measurement: local^-1 * T * local (camera 1 -> camera 2), between cameras, not base_link
    Rm = local_R.T @ R2
    tm = local_R.T @ (t2 - local_t)
    link = (quat_normalize_rotation(quat_from_matrix(Rm)), tm)
    if covariance is None:
        info = np.eye(6)                       # Optimizer/VarianceIgnored
    else:
        info = np.diag(1.0 / np.diag(np.asarray(covariance, np.float64).reshape(6, 6)))   # the path's covariance is diagonal
    return dict(K=(float(fx), float(fy), float(cx), float(cy)), baseline=float(baseline), cam1=cam1, cam2=cam2,
                pts=pts, meas=meas, stereo=stereo, level=np.zeros((2, n), np.int32), link_inv=se3_inv(link), info=info,
                local_R=local_R, local_t=local_t)


def _project(cam, K, baseline, X):
    """EdgeProjectP2SC / P2MC::computeError up to the measurement: (u, v, u_right), camera-frame points."""
    fx, fy, cx, cy = K
    R = quat_to_matrix(cam[0])
    pc = X @ R - (R.T @ cam[1])          # w2n * (X, 1): rows R^T X - R^T t
    with np.errstate(all="ignore"):
        iz = 1.0 / pc[:, 2]
        uvr = np.stack([fx * pc[:, 0] * iz + cx, fy * pc[:, 1] * iz + cy, fx * (pc[:, 0] - baseline) * iz + cx], 1)
    return uvr, pc, R


def _huber(chi2, delta):
    """RobustKernelHuber::robustify: (rho, rho')."""
    if delta <= 0:
        return chi2, np.ones_like(chi2)
    with np.errstate(all="ignore"):
        s = np.sqrt(chi2)
        big = chi2 > delta * delta
        return np.where(big, 2 * s * delta - delta * delta, chi2), np.where(big, delta / s, 1.0)


def _link_error(pb, cam2):
    """EdgeSBACam::computeError: delta = measurement^-1 * (cam1^-1 * cam2); (translation, quaternion vector)."""
    d = se3_mul(pb["link_inv"], se3_mul(se3_inv(pb["cam1"]), cam2))
    return np.concatenate([d[1], d[0][:3]])


def edge_chi2(pb, cam2, pts, var):
    """chi2 of every projection edge [2, N] (inactive ones included) and of the camera link."""
    out = np.zeros((2, len(pts)))
    for c, cam in enumerate((pb["cam1"], cam2)):
        uvr, _, _ = _project(cam, pb["K"], pb["baseline"], pts)
        e = uvr - pb["meas"][c]
        e[:, 2] = np.where(pb["stereo"][c], e[:, 2], 0.0)
        out[c] = (e * e).sum(1) / var
    el = _link_error(pb, cam2)
    return out, float(el @ pb["info"] @ el)


def robust_chi2(pb, cam2, pts, var, delta):
    """SparseOptimizer::activeRobustChi2 over the level-0 edges."""
    chi, link = edge_chi2(pb, cam2, pts, var)
    rho, _ = _huber(chi, delta)
    return float(np.where(pb["level"] == 0, rho, 0.0).sum() + link)


# ---- linear system and Levenberg-Marquardt (g2o::OptimizationAlgorithmLevenberg + BlockSolver_6_3, Schur) ---------

def _linearize(pb, cam2, pts, var, delta):
    """buildSystem over the level-0 edges: per point V [N,3,3], bl [N,3], W [N,6,3]; camera block A [6,6], bc [6]
    (b = -J^T rho' Omega e); robust chi2."""
    n = len(pts)
    V = np.zeros((n, 3, 3))
    bl = np.zeros((n, 3))
    W = np.zeros((n, 6, 3))
    A = np.zeros((6, 6))
    bc = np.zeros(6)
    chi_total = 0.0
    fx, fy, _, _ = pb["K"]
    b = pb["baseline"]
    for c, cam in enumerate((pb["cam1"], cam2)):
        uvr, pc, R = _project(cam, pb["K"], b, pts)
        e = uvr - pb["meas"][c]
        st = pb["stereo"][c]
        e[:, 2] = np.where(st, e[:, 2], 0.0)
        chi = (e * e).sum(1) / var
        rho, w = _huber(chi, delta)
        act = pb["level"][c] == 0
        chi_total += float(np.where(act, rho, 0.0).sum())
        w = np.where(act, w, 0.0) / var
        with np.errstate(all="ignore"):
            iz = 1.0 / pc[:, 2]
            iz2 = iz * iz
        Jpc = np.zeros((n, 3, 3))                       # d(u, v, u_r) / d(camera-frame point)
        Jpc[:, 0, 0] = fx * iz
        Jpc[:, 0, 2] = -fx * pc[:, 0] * iz2
        Jpc[:, 1, 1] = fy * iz
        Jpc[:, 1, 2] = -fy * pc[:, 1] * iz2
        Jpc[:, 2, 0] = np.where(st, fx * iz, 0.0)
        Jpc[:, 2, 2] = np.where(st, -fx * (pc[:, 0] - b) * iz2, 0.0)
        Jp = Jpc @ R.T                                  # d / d(world point)
        V += w[:, None, None] * (Jp.transpose(0, 2, 1) @ Jp)
        bl -= w[:, None] * np.einsum("nki,nk->ni", Jp, e)
        if c == 1:
            D = np.zeros((n, 3, 3))                     # d(camera-frame point) / d(quaternion vector): 2 pc x e_k
            D[:, 1, 0] = 2 * pc[:, 2]
            D[:, 2, 0] = -2 * pc[:, 1]
            D[:, 0, 1] = -2 * pc[:, 2]
            D[:, 2, 1] = 2 * pc[:, 0]
            D[:, 0, 2] = 2 * pc[:, 1]
            D[:, 1, 2] = -2 * pc[:, 0]
            Jc = np.concatenate([-Jp, Jpc @ D], axis=2)  # [N, 3, 6]
            W += w[:, None, None] * (Jc.transpose(0, 2, 1) @ Jp)
            A += np.einsum("n,nki,nkj->ij", w, Jc, Jc)
            bc -= np.einsum("n,nki,nk->i", w, Jc, e)
    # EdgeSBACam 1 -> 2, no kernel, numeric Jacobian (BaseBinaryEdge::linearizeOplus: central differences, 1e-9)
    el = _link_error(pb, cam2)
    Jl = np.zeros((6, 6))
    for d in range(6):
        step = np.zeros(6)
        step[d] = 1e-9
        Jl[:, d] = (_link_error(pb, cam_update(cam2, step)) - _link_error(pb, cam_update(cam2, -step))) * (1.0 / 2e-9)
    A += Jl.T @ pb["info"] @ Jl
    bc -= Jl.T @ pb["info"] @ el
    chi_total += float(el @ pb["info"] @ el)
    return V, bl, W, A, bc, chi_total


def _cholesky_solve(S, rhs):
    n = len(rhs)
    L = np.zeros((n, n))
    for j in range(n):
        d = S[j, j] - L[j, :j] @ L[j, :j]
        if not (d > 0):
            return None
        L[j, j] = np.sqrt(d)
        for i in range(j + 1, n):
            L[i, j] = (S[i, j] - L[i, :j] @ L[j, :j]) / L[j, j]
    y = np.zeros(n)
    for i in range(n):
        y[i] = (rhs[i] - L[i, :i] @ y[:i]) / L[i, i]
    x = np.zeros(n)
    for i in reversed(range(n)):
        x[i] = (y[i] - L[i + 1:, i] @ x[i + 1:]) / L[i, i]
    return x


def _inv3(M):
    """Eigen 3x3 inverse (cofactors), batched."""
    a, b, c = M[:, 0, 0], M[:, 0, 1], M[:, 0, 2]
    d, e, f = M[:, 1, 0], M[:, 1, 1], M[:, 1, 2]
    g, h, i = M[:, 2, 0], M[:, 2, 1], M[:, 2, 2]
    C = np.stack([np.stack([e * i - f * h, c * h - b * i, b * f - c * e], 1),
                  np.stack([f * g - d * i, a * i - c * g, c * d - a * f], 1),
                  np.stack([d * h - e * g, b * g - a * h, a * e - b * d], 1)], 1)
    det = a * C[:, 0, 0] + b * C[:, 1, 0] + c * C[:, 2, 0]
    with np.errstate(all="ignore"):
        return C / det[:, None, None]


def levenberg(pb, cam2, pts, n_iter, var, delta, stats=None):
    """SparseOptimizer::optimize(n_iter) with OptimizationAlgorithmLevenberg; returns (cam2, pts, iterations)."""
    lam, ni = 0.0, 2.0
    done = 0
    active_pt = (pb["level"] == 0).any(0)
    for it in range(n_iter):
        V, bl, W, A, bc, cur_chi = _linearize(pb, cam2, pts, var, delta)
        if it == 0:
            diag = np.abs(np.diag(A)).max()
            if active_pt.any():
                diag = max(diag, np.abs(np.einsum("nii->ni", V))[active_pt].max())
            lam, ni = 1e-5 * diag, 2.0
        rho, qmax = 0.0, 0
        while True:
            Vi = _inv3(V + lam * np.eye(3))
            Vi[~active_pt] = 0.0
            WVi = W @ Vi
            S = A + lam * np.eye(6) - np.einsum("nij,nkj->ik", WVi, W)
            rhs = bc - np.einsum("nij,nj->i", WVi, bl)
            xp = _cholesky_solve(S, rhs)
            if xp is not None:
                xl = np.einsum("nij,nj->ni", Vi, bl - np.einsum("nij,i->nj", W, xp))
                cam_new, pts_new = cam_update(cam2, xp), pts + xl
                tmp_chi = robust_chi2(pb, cam_new, pts_new, var, delta)
                scale = float(xp @ (lam * xp + bc) + (xl * (lam * xl + bl))[active_pt].sum())
            else:
                tmp_chi, scale = HUGE, 0.0
            rho = (cur_chi - tmp_chi) / (scale + 1e-3)
            if rho > 0 and np.isfinite(tmp_chi):
                alpha = min(1.0 - (2 * rho - 1) ** 3, 2.0 / 3.0)
                lam *= max(1.0 / 3.0, alpha)
                ni = 2.0
                cur_chi = tmp_chi
                cam2, pts = cam_new, pts_new
            else:
                lam *= ni
                ni *= 2
                if not np.isfinite(lam):
                    break
            qmax += 1
            if not (rho < 0 and qmax < 10):
                break
        done += 1
        if stats is not None:
            stats.append((it, cur_chi, lam, qmax))
        if qmax == 10 or rho == 0 or not np.isfinite(lam):
            break
    return cam2, pts, done


def optimize_ba(pb, iterations=20, pixel_variance=1.0, robust_delta=8.0):
    """OptimizerG2O::optimizeBA for this 2-camera graph.  Returns dict(ok, cam2=(q, t), pts, outliers (sorted point
    indices), chi2, iterations)."""
    cam2, pts = pb["cam2"], pb["pts"].copy()
    pb["level"][:] = 0
    outliers = set()
    total = 0
    rounds = 2 if robust_delta > 0 else 1
    for r in range(rounds):
        cam2, pts, done = levenberg(pb, cam2, pts, 5 if (r == 0 and robust_delta > 0) else iterations, pixel_variance,
                                    robust_delta)
        total += done
        chi2 = robust_chi2(pb, cam2, pts, pixel_variance, robust_delta)
        if np.isnan(chi2) or (r == 0 and (chi2 > 1e12 or not np.isfinite(chi2))):
            return dict(ok=False, outliers=[], chi2=chi2, iterations=total)
        if robust_delta > 0:
            chi, _ = edge_chi2(pb, cam2, pts, pixel_variance)
            mark = (pb["level"] == 0) & (chi > robust_delta)
            pb["level"][mark] = 1
            outliers |= set(np.nonzero(mark.any(0))[0].tolist())
    if chi2 > 1e12:      # activeRobustChi2() of the last stage's active set, as computed before the marking
        return dict(ok=False, outliers=[], chi2=chi2, iterations=total)
    return dict(ok=True, cam2=cam2, pts=pts, outliers=sorted(outliers), chi2=chi2, iterations=total)


def transform_of(pb, cam2):
    """camera 2's pose * localTransform^-1 as a 3x4 f32 transform (From base <- To base)."""
    R = quat_to_matrix(cam2[0]) @ pb["local_R"].T
    t = cam2[1] - R @ pb["local_t"]
    return np.concatenate([R, t[:, None]], 1).astype(np.float32)


def bundle_adjust_pair(T12, from_kp2d, from_kp3d, to_kp2d, to_kp3d, K, baseline, local_R, local_t, covariance,
                       min_inliers, iterations=20, pixel_variance=1.0, robust_delta=8.0):
    """The block :1192-1370 for one pair: (ok, T 3x4 f32 or zeros, kept [N] bool, result dict)."""
    pb = build_problem(T12, from_kp2d, from_kp3d, to_kp2d, to_kp3d, K, baseline, local_R, local_t, covariance)
    res = optimize_ba(pb, iterations, pixel_variance, robust_delta)
    n = len(from_kp3d)
    kept = np.ones(n, bool)
    if not res["ok"]:
        return False, np.zeros((3, 4), np.float32), kept, res
    kept[res["outliers"]] = False
    if kept.sum() < min_inliers:
        return False, np.zeros((3, 4), np.float32), kept, res
    return True, transform_of(pb, res["cam2"]), kept, res
