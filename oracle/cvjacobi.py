"""TEST INFRASTRUCTURE ONLY (oracle) -- never imported by the product path.

Pure-Python (IEEE double, no FMA) restatement of the small-matrix SVD that OpenCV's
calib3d code reaches for matrices with fewer than 25 rows: the LAPACK HAL declines those
and `cv::SVD` falls back to OpenCV's own one-sided Jacobi (Hestenes) sweep.  This is
third-party arithmetic that the reference reaches through

    R/src/myRegistrationVis.cpp:1077-1090  util3d::estimateMotion3DTo2D
        -> cv::solvePnPRansac -> epnp::compute_pose -> cvSVD / cvSolve(CV_SVD) / cvInvert(CV_SVD)

OpenCV is NOT vendored under /root/reference (reference image pins OpenCV 3.2; this
container has cv2 4.13.0).  The restatement is pinned bit-for-bit against
`cv2.SVDecomp` in tests/test_oracle_cvjacobi.py.

Functions operate on lists-of-lists / numpy float64 arrays and go through Python floats
so that every operation is a single correctly rounded IEEE-754 double operation.
"""
from __future__ import annotations

import math

import numpy as np

DBL_EPSILON = 2.220446049250313e-16
DBL_MIN = 2.2250738585072014e-308

def cv_hypot(a: float, b: float) -> float:
    """OpenCV's file-local `hypot` template in core/src/lapack.cpp (NOT libm's hypot:
    unqualified lookup inside namespace cv finds the local template first).  Pinned
    empirically: only this form reproduces cv2.SVDecomp bit-for-bit."""
    a = abs(a)
    b = abs(b)
    if a > b:
        b /= a
        return a * math.sqrt(1 + b * b)
    if b > 0:
        a /= b
        return b * math.sqrt(1 + a * a)
    return 0.0


class CvRNG:
    """cv::RNG multiply-with-carry generator (used for zero singular values and RANSAC)."""

    COEFF = 4164903690

    def __init__(self, state: int = 0xFFFFFFFFFFFFFFFF):
        self.state = state & 0xFFFFFFFFFFFFFFFF

    def next(self) -> int:
        self.state = ((self.state & 0xFFFFFFFF) * self.COEFF + (self.state >> 32)) & 0xFFFFFFFFFFFFFFFF
        return self.state & 0xFFFFFFFF

    def uniform_int(self, a: int, b: int) -> int:
        return a if a == b else self.next() % (b - a) + a


def jacobi_svd_rows(At, n1=None):
    """One-sided Jacobi on the ROWS of At (n rows of length m), in place.

    Returns (W, At_normalised, Vt): W descending; rows of At become the unit left
    vectors (of A = At^T), rows of Vt the matching right vectors.
    """
    At = [[float(x) for x in row] for row in At]
    n = len(At)
    m = len(At[0])
    if n1 is None:
        n1 = n
    eps = DBL_EPSILON * 10
    minval = DBL_MIN
    W = [0.0] * n
    Vt = [[0.0] * n for _ in range(n)]
    for i in range(n):
        sd = 0.0
        for k in range(m):
            t = At[i][k]
            sd += t * t
        W[i] = sd
        Vt[i][i] = 1.0
    max_iter = max(m, 30)
    for _ in range(max_iter):
        changed = False
        for i in range(n - 1):
            for j in range(i + 1, n):
                Ai, Aj = At[i], At[j]
                a, p, b = W[i], 0.0, W[j]
                for k in range(m):
                    p += Ai[k] * Aj[k]
                if abs(p) <= eps * math.sqrt(a * b):
                    continue
                p *= 2.0
                beta = a - b
                gamma = cv_hypot(p, beta)
                if beta < 0:
                    delta = (gamma - beta) * 0.5
                    s = math.sqrt(delta / gamma)
                    c = p / (gamma * s * 2)
                else:
                    c = math.sqrt((gamma + beta) / (gamma * 2))
                    s = p / (gamma * c * 2)
                a = b = 0.0
                for k in range(m):
                    t0 = c * Ai[k] + s * Aj[k]
                    t1 = -s * Ai[k] + c * Aj[k]
                    Ai[k] = t0
                    Aj[k] = t1
                    a += t0 * t0
                    b += t1 * t1
                W[i] = a
                W[j] = b
                changed = True
                Vi, Vj = Vt[i], Vt[j]
                for k in range(n):
                    t0 = c * Vi[k] + s * Vj[k]
                    t1 = -s * Vi[k] + c * Vj[k]
                    Vi[k] = t0
                    Vj[k] = t1
        if not changed:
            break
    for i in range(n):
        sd = 0.0
        for k in range(m):
            t = At[i][k]
            sd += t * t
        W[i] = math.sqrt(sd)
    for i in range(n - 1):
        j = i
        for k in range(i + 1, n):
            if W[j] < W[k]:
                j = k
        if i != j:
            W[i], W[j] = W[j], W[i]
            At[i], At[j] = At[j], At[i]
            Vt[i], Vt[j] = Vt[j], Vt[i]
    rng = CvRNG(0x12345678)
    for i in range(n1):
        sd = W[i] if i < n else 0.0
        ii = 0
        while ii < 100 and sd <= minval:
            val0 = 1.0 / m
            for k in range(m):
                At[i][k] = val0 if (rng.next() & 256) != 0 else -val0
            for _it in range(2):
                for j in range(i):
                    sd = 0.0
                    for k in range(m):
                        sd += At[i][k] * At[j][k]
                    asum = 0.0
                    for k in range(m):
                        t = At[i][k] - sd * At[j][k]
                        At[i][k] = t
                        asum += abs(t)
                    asum = 1.0 / asum if asum > eps * 100 else 0.0
                    for k in range(m):
                        At[i][k] *= asum
            sd = 0.0
            for k in range(m):
                t = At[i][k]
                sd += t * t
            sd = math.sqrt(sd)
            ii += 1
        s = 1.0 / sd if sd > minval else 0.0
        for k in range(m):
            At[i][k] *= s
    return W, At, Vt


def svd(A):
    """cv::SVD::compute(A) for m >= n: returns (w[n], U[m][n], Vt[n][n]) like cv2.SVDecomp."""
    A = np.asarray(A, dtype=np.float64)
    m, n = A.shape
    assert m >= n, "only the m >= n branch is restated (all call sites on this path)"
    W, Ut, Vt = jacobi_svd_rows(A.T.tolist())
    return np.array(W), np.array(Ut).T.copy(), np.array(Vt)


def svd_backsubst(w, u, vt, b):
    """cv::SVD::backSubst for a single right-hand side (SVBkSb): x = V diag(1/w) U^T b.

    threshold = eps * 2 * sum(w); singular values <= threshold are skipped.
    """
    w = [float(x) for x in np.asarray(w).ravel()]
    u = np.asarray(u, dtype=np.float64)
    vt = np.asarray(vt, dtype=np.float64)
    b = [float(x) for x in np.asarray(b).ravel()]
    m, nm = u.shape
    n = vt.shape[1]
    threshold = 0.0
    for i in range(nm):
        threshold += w[i]
    threshold *= DBL_EPSILON * 2
    x = [0.0] * n
    for i in range(nm):
        wi = w[i]
        if abs(wi) <= threshold:
            continue
        wi = 1.0 / wi
        s = 0.0
        for k in range(m):
            s += float(u[k, i]) * b[k]
        s *= wi
        for k in range(n):
            x[k] = x[k] + s * float(vt[i, k])
    return np.array(x)


def solve_svd(A, b):
    """cvSolve(A, b, x, CV_SVD) for m >= n, single rhs."""
    w, u, vt = svd(A)
    return svd_backsubst(w, u, vt, b)
