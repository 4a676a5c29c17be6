"""TEST INFRASTRUCTURE ONLY (oracle) -- never imported by the product path.

Pure-Python IEEE-double restatement of the minimal solver that the reference's PnP step
reaches:  R/src/myRegistrationVis.cpp:1077-1090 -> (RTAB-Map, unvendored)
util3d::estimateMotion3DTo2D -> cv::solvePnPRansac -> PnPRansacCallback::runKernel ->
cv::solvePnP(SOLVEPNP_EPNP) -> epnp::compute_pose (OpenCV calib3d, unvendored; algorithm
of Lepetit/Moreno-Noguer/Fua "EPnP", IJCV 2009) plus cv::Rodrigues and cv::projectPoints.

Pinned against cv2 4.13.0 `cv2.solvePnP(flags=SOLVEPNP_EPNP)` / `cv2.Rodrigues` /
`cv2.projectPoints` in tests/test_oracle_epnp.py (bit-for-bit where only
+,-,*,/,sqrt are involved).  The operation ORDER below is part of the specification: for a
5-point minimal set M^T M (12x12) has a 2-dimensional exact null space whose basis is decided
by rounding noise, so any re-association changes the basis EPnP linearises in.
"""
from __future__ import annotations

import math

import numpy as np

from . import cvjacobi as J


def f32(x):
    return float(np.float32(x))


def undistort_normalise_f32(img_pts, fx, fy, cx, cy):
    """cv::undistortPoints with zero distortion on CV_32FC2 input: double math, f32 store."""
    ifx, ify = 1.0 / fx, 1.0 / fy
    out = []
    for (u, v) in img_pts:
        x = (float(u) - cx) * ifx
        y = (float(v) - cy) * ify
        out.append((f32(x), f32(y)))
    return out


def _dot3(a, b):
    return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]


def _dist2(a, b):
    return (a[0] - b[0]) * (a[0] - b[0]) + (a[1] - b[1]) * (a[1] - b[1]) + (a[2] - b[2]) * (a[2] - b[2])


def mul_transposed_ata(M):
    """cv::mulTransposed(M, dst, aTa=true) (MulTransposedR<double,double>) + completeSymm."""
    rows = len(M)
    cols = len(M[0])
    dst = [[0.0] * cols for _ in range(cols)]
    for i in range(cols):
        for j in range(i, cols):
            s = 0.0
            for k in range(rows):
                s += M[k][i] * M[k][j]
            dst[i][j] = s
    for i in range(cols):
        for j in range(i):
            dst[i][j] = dst[j][i]
    return dst


def _qr_solve(A, b, nr, nc):
    """epnp::qr_solve (Householder; including the row-scan quirk for eta)."""
    A = list(A)
    b = list(b)
    A1 = [0.0] * nr
    A2 = [0.0] * nr
    for k in range(nc):
        kk = k * nc + k
        p = kk
        eta = abs(A[p])
        for _i in range(k + 1, nr):
            elt = abs(A[p])
            if eta < elt:
                eta = elt
            p += nc
        if eta == 0:
            return None
        inv_eta = 1.0 / eta
        sum2 = 0.0
        p = kk
        for _i in range(k, nr):
            A[p] *= inv_eta
            sum2 += A[p] * A[p]
            p += nc
        sigma = math.sqrt(sum2)
        if A[kk] < 0:
            sigma = -sigma
        A[kk] += sigma
        A1[k] = sigma * A[kk]
        A2[k] = -eta * sigma
        for j in range(k + 1, nc):
            s = 0.0
            p = kk
            for _i in range(k, nr):
                s += A[p] * A[p + j - k]
                p += nc
            tau = s / A1[k]
            p = kk
            for _i in range(k, nr):
                A[p + j - k] -= tau * A[p]
                p += nc
    for j in range(nc):
        jj = j * nc + j
        tau = 0.0
        p = jj
        for i in range(j, nr):
            tau += A[p] * b[i]
            p += nc
        tau /= A1[j]
        p = jj
        for i in range(j, nr):
            b[i] -= tau * A[p]
            p += nc
    X = [0.0] * nc
    X[nc - 1] = b[nc - 1] / A2[nc - 1]
    for i in range(nc - 2, -1, -1):
        s = 0.0
        p = i * nc + (i + 1)
        for j in range(i + 1, nc):
            s += A[p] * X[j]
            p += 1
        X[i] = (b[i] - s) / A2[i]
    return X


class EPnP:
    def __init__(self, obj_pts, norm_pts, fx, fy, cx, cy):
        self.fu, self.fv, self.uc, self.vc = float(fx), float(fy), float(cx), float(cy)
        self.n = len(obj_pts)
        self.pws = [[float(c) for c in p] for p in obj_pts]
        self.us = [(float(p[0]) * self.fu + self.uc, float(p[1]) * self.fv + self.vc) for p in norm_pts]

    # -- control points -------------------------------------------------------------------
    def choose_control_points(self):
        n = self.n
        cws = [[0.0] * 3 for _ in range(4)]
        for i in range(n):
            for j in range(3):
                cws[0][j] += self.pws[i][j]
        for j in range(3):
            cws[0][j] /= n
        PW0 = [[self.pws[i][j] - cws[0][j] for j in range(3)] for i in range(n)]
        ptp = mul_transposed_ata(PW0)
        # cvSVD(PW0tPW0, DC, UCt, 0, MODIFY_A | U_T): rows of UCt are the left vectors.
        W, Ut, _ = J.jacobi_svd_rows([[ptp[r][c] for r in range(3)] for c in range(3)])
        for i in range(1, 4):
            k = math.sqrt(W[i - 1] / n)
            for j in range(3):
                cws[i][j] = cws[0][j] + k * Ut[i - 1][j]
        self.cws = cws

    def compute_barycentric_coordinates(self):
        cws = self.cws
        cc = [[cws[j][i] - cws[0][i] for j in range(1, 4)] for i in range(3)]
        w, u, vt = J.svd(np.array(cc))
        ci = np.stack([J.svd_backsubst(w, u, vt, np.eye(3)[:, c]) for c in range(3)], axis=1)
        ci = [[float(ci[r, c]) for c in range(3)] for r in range(3)]
        self.alphas = []
        for i in range(self.n):
            pi = self.pws[i]
            a = [0.0] * 4
            for j in range(3):
                a[1 + j] = (ci[j][0] * (pi[0] - cws[0][0]) + ci[j][1] * (pi[1] - cws[0][1])
                            + ci[j][2] * (pi[2] - cws[0][2]))
            a[0] = 1.0 - a[1] - a[2] - a[3]
            self.alphas.append(a)

    def build_M(self):
        M = []
        for i in range(self.n):
            a = self.alphas[i]
            u, v = self.us[i]
            r1 = [0.0] * 12
            r2 = [0.0] * 12
            for k in range(4):
                r1[3 * k] = a[k] * self.fu
                r1[3 * k + 1] = 0.0
                r1[3 * k + 2] = a[k] * (self.uc - u)
                r2[3 * k] = 0.0
                r2[3 * k + 1] = a[k] * self.fv
                r2[3 * k + 2] = a[k] * (self.vc - v)
            M.append(r1)
            M.append(r2)
        return M

    # -- betas ----------------------------------------------------------------------------
    @staticmethod
    def compute_L_6x10(ut):
        v = [ut[11], ut[10], ut[9], ut[8]]
        dv = [[[0.0] * 3 for _ in range(6)] for _ in range(4)]
        for i in range(4):
            a, b = 0, 1
            for j in range(6):
                dv[i][j][0] = v[i][3 * a] - v[i][3 * b]
                dv[i][j][1] = v[i][3 * a + 1] - v[i][3 * b + 1]
                dv[i][j][2] = v[i][3 * a + 2] - v[i][3 * b + 2]
                b += 1
                if b > 3:
                    a += 1
                    b = a + 1
        L = []
        for i in range(6):
            L.append([
                _dot3(dv[0][i], dv[0][i]),
                2.0 * _dot3(dv[0][i], dv[1][i]),
                _dot3(dv[1][i], dv[1][i]),
                2.0 * _dot3(dv[0][i], dv[2][i]),
                2.0 * _dot3(dv[1][i], dv[2][i]),
                _dot3(dv[2][i], dv[2][i]),
                2.0 * _dot3(dv[0][i], dv[3][i]),
                2.0 * _dot3(dv[1][i], dv[3][i]),
                2.0 * _dot3(dv[2][i], dv[3][i]),
                _dot3(dv[3][i], dv[3][i]),
            ])
        return L

    def compute_rho(self):
        c = self.cws
        return [_dist2(c[0], c[1]), _dist2(c[0], c[2]), _dist2(c[0], c[3]),
                _dist2(c[1], c[2]), _dist2(c[1], c[3]), _dist2(c[2], c[3])]

    @staticmethod
    def find_betas_approx_1(L, rho):
        A = [[L[i][0], L[i][1], L[i][3], L[i][6]] for i in range(6)]
        b4 = J.solve_svd(np.array(A), np.array(rho))
        b4 = [float(x) for x in b4]
        betas = [0.0] * 4
        if b4[0] < 0:
            betas[0] = math.sqrt(-b4[0])
            betas[1] = -b4[1] / betas[0]
            betas[2] = -b4[2] / betas[0]
            betas[3] = -b4[3] / betas[0]
        else:
            betas[0] = math.sqrt(b4[0])
            betas[1] = b4[1] / betas[0]
            betas[2] = b4[2] / betas[0]
            betas[3] = b4[3] / betas[0]
        return betas

    @staticmethod
    def find_betas_approx_2(L, rho):
        A = [[L[i][0], L[i][1], L[i][2]] for i in range(6)]
        b3 = [float(x) for x in J.solve_svd(np.array(A), np.array(rho))]
        betas = [0.0] * 4
        if b3[0] < 0:
            betas[0] = math.sqrt(-b3[0])
            betas[1] = math.sqrt(-b3[2]) if b3[2] < 0 else 0.0
        else:
            betas[0] = math.sqrt(b3[0])
            betas[1] = math.sqrt(b3[2]) if b3[2] > 0 else 0.0
        if b3[1] < 0:
            betas[0] = -betas[0]
        return betas

    @staticmethod
    def find_betas_approx_3(L, rho):
        A = [[L[i][0], L[i][1], L[i][2], L[i][3], L[i][4]] for i in range(6)]
        b5 = [float(x) for x in J.solve_svd(np.array(A), np.array(rho))]
        betas = [0.0] * 4
        if b5[0] < 0:
            betas[0] = math.sqrt(-b5[0])
            betas[1] = math.sqrt(-b5[2]) if b5[2] < 0 else 0.0
        else:
            betas[0] = math.sqrt(b5[0])
            betas[1] = math.sqrt(b5[2]) if b5[2] > 0 else 0.0
        if b5[1] < 0:
            betas[0] = -betas[0]
        betas[2] = b5[3] / betas[0] if betas[0] != 0 else math.copysign(math.inf, b5[3]) if b5[3] != 0 else math.nan
        betas[3] = 0.0
        return betas

    @staticmethod
    def gauss_newton(L, rho, betas):
        betas = list(betas)
        stale = [0.0] * 4
        for _ in range(5):
            A = []
            b = []
            for i in range(6):
                r = L[i]
                A += [
                    2 * r[0] * betas[0] + r[1] * betas[1] + r[3] * betas[2] + r[6] * betas[3],
                    r[1] * betas[0] + 2 * r[2] * betas[1] + r[4] * betas[2] + r[7] * betas[3],
                    r[3] * betas[0] + r[4] * betas[1] + 2 * r[5] * betas[2] + r[8] * betas[3],
                    r[6] * betas[0] + r[7] * betas[1] + r[8] * betas[2] + 2 * r[9] * betas[3],
                ]
                b.append(rho[i] - (
                    r[0] * betas[0] * betas[0] +
                    r[1] * betas[0] * betas[1] +
                    r[2] * betas[1] * betas[1] +
                    r[3] * betas[0] * betas[2] +
                    r[4] * betas[1] * betas[2] +
                    r[5] * betas[2] * betas[2] +
                    r[6] * betas[0] * betas[3] +
                    r[7] * betas[1] * betas[3] +
                    r[8] * betas[2] * betas[3] +
                    r[9] * betas[3] * betas[3]))
            x = _qr_solve(A, b, 6, 4)
            if x is None:
                # qr_solve returns early leaving the caller's x[4] (zero-initialised per call) untouched
                x = stale
            stale = x
            for i in range(4):
                betas[i] += x[i]
        return betas

    # -- pose from betas ------------------------------------------------------------------
    def compute_R_and_t(self, ut, betas):
        n = self.n
        ccs = [[0.0] * 3 for _ in range(4)]
        for i in range(4):
            v = ut[11 - i]
            for j in range(4):
                for k in range(3):
                    ccs[j][k] += betas[i] * v[3 * j + k]
        pcs = []
        for i in range(n):
            a = self.alphas[i]
            pcs.append([a[0] * ccs[0][j] + a[1] * ccs[1][j] + a[2] * ccs[2][j] + a[3] * ccs[3][j]
                        for j in range(3)])
        if pcs[0][2] < 0.0:
            ccs = [[-x for x in c] for c in ccs]
            pcs = [[-x for x in p] for p in pcs]
        pc0 = [0.0] * 3
        pw0 = [0.0] * 3
        for i in range(n):
            for j in range(3):
                pc0[j] += pcs[i][j]
                pw0[j] += self.pws[i][j]
        for j in range(3):
            pc0[j] /= n
            pw0[j] /= n
        abt = [[0.0] * 3 for _ in range(3)]
        for i in range(n):
            pc, pw = pcs[i], self.pws[i]
            for j in range(3):
                abt[j][0] += (pc[j] - pc0[j]) * (pw[0] - pw0[0])
                abt[j][1] += (pc[j] - pc0[j]) * (pw[1] - pw0[1])
                abt[j][2] += (pc[j] - pc0[j]) * (pw[2] - pw0[2])
        # cvSVD(ABt, D, U, V, MODIFY_A): U = left vectors as columns, V = right vectors as columns
        W, Ut_, Vt_ = J.jacobi_svd_rows([[abt[r][c] for r in range(3)] for c in range(3)])
        U = [[Ut_[c][r] for c in range(3)] for r in range(3)]
        V = [[Vt_[c][r] for c in range(3)] for r in range(3)]
        R = [[_dot3(U[i], V[j]) for j in range(3)] for i in range(3)]
        det = (R[0][0] * R[1][1] * R[2][2] + R[0][1] * R[1][2] * R[2][0] + R[0][2] * R[1][0] * R[2][1]
               - R[0][2] * R[1][1] * R[2][0] - R[0][1] * R[1][0] * R[2][2] - R[0][0] * R[1][2] * R[2][1])
        if det < 0:
            R[2] = [-R[2][0], -R[2][1], -R[2][2]]
        t = [pc0[0] - _dot3(R[0], pw0), pc0[1] - _dot3(R[1], pw0), pc0[2] - _dot3(R[2], pw0)]
        # reprojection error
        sum2 = 0.0
        for i in range(n):
            pw = self.pws[i]
            Xc = _dot3(R[0], pw) + t[0]
            Yc = _dot3(R[1], pw) + t[1]
            inv_Zc = 1.0 / (_dot3(R[2], pw) + t[2])
            ue = self.uc + self.fu * Xc * inv_Zc
            ve = self.vc + self.fv * Yc * inv_Zc
            u, v = self.us[i]
            sum2 += math.sqrt((u - ue) * (u - ue) + (v - ve) * (v - ve))
        return sum2 / n, R, t

    def compute_pose(self, return_debug=False):
        self.choose_control_points()
        self.compute_barycentric_coordinates()
        M = self.build_M()
        mtm = mul_transposed_ata(M)
        # cvSVD(MtM, D, Ut, 0, MODIFY_A | U_T): Jacobi runs on rows of MtM^T (== columns of MtM)
        W, ut, _ = J.jacobi_svd_rows([[mtm[r][c] for r in range(12)] for c in range(12)])
        L = self.compute_L_6x10(ut)
        rho = self.compute_rho()
        res = []
        for finder in (self.find_betas_approx_1, self.find_betas_approx_2, self.find_betas_approx_3):
            b = finder(L, rho)
            b = self.gauss_newton(L, rho, b)
            res.append(self.compute_R_and_t(ut, b))
        N = 0
        if res[1][0] < res[0][0]:
            N = 1
        if res[2][0] < res[N][0]:
            N = 2
        if return_debug:
            return res[N][1], res[N][2], dict(ut=ut, W=W, mtm=mtm, errs=[r[0] for r in res], N=N + 1)
        return res[N][1], res[N][2]


# ---------------------------------------------------------------------------------------
def rodrigues_mat_to_vec(R):
    """cv::Rodrigues 3x3 -> rvec (orthonormalises through SVD first)."""
    Rm = np.asarray(R, dtype=np.float64)
    W, Ut_, Vt_ = J.jacobi_svd_rows(Rm.T.tolist())
    U = [[Ut_[c][r] for c in range(3)] for r in range(3)]
    # R = U * Vt  (cv::Matx multiply: row-by-column, sequential sum)
    Rn = [[U[i][0] * Vt_[0][j] + U[i][1] * Vt_[1][j] + U[i][2] * Vt_[2][j] for j in range(3)] for i in range(3)]
    rx = Rn[2][1] - Rn[1][2]
    ry = Rn[0][2] - Rn[2][0]
    rz = Rn[1][0] - Rn[0][1]
    s = math.sqrt((rx * rx + ry * ry + rz * rz) * 0.25)
    c = (Rn[0][0] + Rn[1][1] + Rn[2][2] - 1) * 0.5
    c = 1.0 if c > 1.0 else (-1.0 if c < -1.0 else c)
    theta = math.acos(c)
    if s < 1e-5:
        if c > 0:
            return [0.0, 0.0, 0.0]
        t = (Rn[0][0] + 1) * 0.5
        rx = math.sqrt(max(t, 0.0))
        t = (Rn[1][1] + 1) * 0.5
        ry = math.sqrt(max(t, 0.0)) * (-1.0 if Rn[0][1] < 0 else 1.0)
        t = (Rn[2][2] + 1) * 0.5
        rz = math.sqrt(max(t, 0.0)) * (-1.0 if Rn[0][2] < 0 else 1.0)
        if abs(rx) < abs(ry) and abs(rx) < abs(rz) and ((Rn[1][2] > 0) != (ry * rz > 0)):
            rz = -rz
        nrm = math.sqrt(rx * rx + ry * ry + rz * rz)
        theta /= nrm
        return [rx * theta, ry * theta, rz * theta]
    vth = 1.0 / (2 * s)
    vth *= theta
    return [rx * vth, ry * vth, rz * vth]


def rodrigues_vec_to_mat(r):
    """cv::Rodrigues rvec -> 3x3."""
    rx, ry, rz = (float(x) for x in r)
    theta = math.sqrt(rx * rx + ry * ry + rz * rz)
    if theta < 2.220446049250313e-16:
        return [[1.0, 0.0, 0.0], [0.0, 1.0, 0.0], [0.0, 0.0, 1.0]]
    c = math.cos(theta)
    s = math.sin(theta)
    c1 = 1.0 - c
    itheta = 1.0 / theta if theta else 0.0
    rx *= itheta
    ry *= itheta
    rz *= itheta
    rrt = [[rx * rx, rx * ry, rx * rz], [rx * ry, ry * ry, ry * rz], [rx * rz, ry * rz, rz * rz]]
    rxm = [[0.0, -rz, ry], [rz, 0.0, -rx], [-ry, rx, 0.0]]
    eye = [[1.0, 0.0, 0.0], [0.0, 1.0, 0.0], [0.0, 0.0, 1.0]]
    # R = c*I + c1*rrt + s*r_x   (Matx expression evaluated left to right)
    return [[(c * eye[i][j] + c1 * rrt[i][j]) + s * rxm[i][j] for j in range(3)] for i in range(3)]


def project_points_f32(obj_pts, R, t, fx, fy, cx, cy):
    """cv::projectPoints without distortion, f32 output (double internals)."""
    out = []
    for p in obj_pts:
        X, Y, Z = float(p[0]), float(p[1]), float(p[2])
        x = R[0][0] * X + R[0][1] * Y + R[0][2] * Z + t[0]
        y = R[1][0] * X + R[1][1] * Y + R[1][2] * Z + t[1]
        z = R[2][0] * X + R[2][1] * Y + R[2][2] * Z + t[2]
        z = 1.0 / z if z != 0 else 1.0
        x *= z
        y *= z
        out.append((np.float32(x * fx + cx), np.float32(y * fy + cy)))
    return out


def solve_epnp(obj_pts_f32, img_pts_f32, fx, fy, cx, cy):
    """cv2.solvePnP(obj, img, K, None, flags=SOLVEPNP_EPNP) -> (rvec, tvec, R)."""
    norm = undistort_normalise_f32(img_pts_f32, fx, fy, cx, cy)
    e = EPnP(obj_pts_f32, norm, fx, fy, cx, cy)
    R, t = e.compute_pose()
    return rodrigues_mat_to_vec(R), t, R
