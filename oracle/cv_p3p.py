"""TEST INFRASTRUCTURE ONLY (oracle) -- never imported by the product path.

Restatement of OpenCV's 4-point P3P minimal solver (modules/calib3d/src/p3p.cpp + polynom_solver.cpp of the
OpenCV 3.2 the reference image pins; unvendored): Gao/Hou/Tang/Chang's complete solution
classification -- quartic in x = |PA|/|PC| solved in closed form (Ferrari via the resolvent cubic), the three
lengths turned into camera-frame points, absolute orientation by Horn's quaternion (4x4 symmetric Jacobi), the
fourth point picks the root with the smallest reprojection error.  It is reached from
R/src/myRegistrationVis.cpp:1084 through `_PnPFlags` (parsed :103) = SOLVEPNP_P3P, or whenever exactly four
correspondences reach cv::solvePnPRansac (the `npoints == 4` branch).

Plain Python floats (IEEE double, one rounding per operation like the C++), same statement order.  `pow`, `acos`
and `cos` are libm calls: the last bit may differ between libm builds, so parity of this path is numerical
(<= 1e-9 on well-conditioned sets), not bit-for-bit.  Pinned against cv2.solveP3P / cv2.solvePnP(SOLVEPNP_P3P)
in tests/test_oracle_cv.py.
"""
from __future__ import annotations

import math

import numpy as np

CV_PI = 3.1415926535897932384626433832795


def solve_deg2(a, b, c):
    delta = b * b - 4 * a * c
    if delta < 0:
        return []
    inv_2a = 0.5 / a
    if delta == 0:
        x = -b * inv_2a
        return [x]
    sq = math.sqrt(delta)
    return [(-b + sq) * inv_2a, (-b - sq) * inv_2a]


def solve_deg3(a, b, c, d):
    """Real roots of a x^3 + b x^2 + c x + d (MathWorld's cubic formula, as polynom_solver.cpp)."""
    if a == 0:
        if b == 0:
            if c == 0:
                return []
            return [-d / c]
        return solve_deg2(b, c, d)
    inv_a = 1.0 / a
    b_a = inv_a * b
    b_a2 = b_a * b_a
    c_a = inv_a * c
    d_a = inv_a * d
    Q = (3 * c_a - b_a2) / 9
    R = (9 * b_a * c_a - 27 * d_a - 2 * b_a * b_a2) / 54
    Q3 = Q * Q * Q
    D = Q3 + R * R
    b_a_3 = (1.0 / 3.0) * b_a
    if Q == 0:
        if R == 0:
            return [-b_a_3, -b_a_3, -b_a_3]
        return [math.pow(2 * R, 1 / 3.0) - b_a_3] if R > 0 else [float("nan")]
    if D <= 0:
        theta = math.acos(R / math.sqrt(-Q3))
        sqrt_Q = math.sqrt(-Q)
        return [2 * sqrt_Q * math.cos(theta / 3.0) - b_a_3,
                2 * sqrt_Q * math.cos((theta + 2 * CV_PI) / 3.0) - b_a_3,
                2 * sqrt_Q * math.cos((theta + 4 * CV_PI) / 3.0) - b_a_3]
    AD = math.pow(abs(R) + math.sqrt(D), 1.0 / 3.0) * (1 if R > 0 else (-1 if R < 0 else 0))
    BD = 0.0 if AD == 0 else -Q / AD
    return [AD + BD - b_a_3]


def solve_deg4(a, b, c, d, e):
    """Real roots of the quartic (MathWorld's quartic formula, as polynom_solver.cpp); 0, 2 or 4 values."""
    if a == 0:
        return solve_deg3(b, c, d, e)
    inv_a = 1.0 / a
    b *= inv_a
    c *= inv_a
    d *= inv_a
    e *= inv_a
    b2 = b * b
    bc = b * c
    b3 = b2 * b
    r = solve_deg3(1.0, -c, d * b - 4 * e, 4 * c * e - d * d - b2 * e)
    if not r:
        return []
    r0 = r[0]
    R2 = 0.25 * b2 - c + r0
    if not (R2 >= 0):
        return []
    R = math.sqrt(R2)
    roots = []
    if R < 10e-12:
        temp = r0 * r0 - 4 * e
        if temp < 0:
            D2 = E2 = -1.0
        else:
            st = math.sqrt(temp)
            D2 = 0.75 * b2 - 2 * c + 2 * st
            E2 = D2 - 4 * st
    else:
        inv_R = 1.0 / R
        u = 0.75 * b2 - 2 * c - R2
        v = 0.25 * inv_R * (4 * bc - 8 * d - b3)
        D2 = u + v
        E2 = u - v
    b_4 = 0.25 * b
    R_2 = 0.5 * R
    if D2 >= 0:
        Dv = math.sqrt(D2)
        x0 = R_2 + 0.5 * Dv - b_4
        roots += [x0, x0 - Dv]
    if E2 >= 0:
        Ev = math.sqrt(E2)
        x2 = -R_2 + 0.5 * Ev - b_4
        roots += [x2, x2 - Ev]
    return roots


def jacobi_4x4(A):
    """p3p::jacobi_4x4: cyclic Jacobi eigen-decomposition of a symmetric 4x4 (upper triangle used, A destroyed).
    Returns (D eigenvalues, U row-major 16 with eigenvectors in columns)."""
    A = list(A)
    U = [1.0 if (i // 4) == (i % 4) else 0.0 for i in range(16)]
    B = [A[0], A[5], A[10], A[15]]
    D = list(B)
    Z = [0.0] * 4
    for it in range(50):
        s_ = abs(A[1]) + abs(A[2]) + abs(A[3]) + abs(A[6]) + abs(A[7]) + abs(A[11])
        if s_ == 0.0:
            return D, U
        tresh = 0.2 * s_ / 16.0 if it < 3 else 0.0
        for i in range(3):
            p = 5 * i + 1
            for j in range(i + 1, 4):
                Aij = A[p]
                eps_machine = 100.0 * abs(Aij)
                if it > 3 and abs(D[i]) + eps_machine == abs(D[i]) and abs(D[j]) + eps_machine == abs(D[j]):
                    A[p] = 0.0
                elif abs(Aij) > tresh:
                    hh = D[j] - D[i]
                    if abs(hh) + eps_machine == abs(hh):
                        t = Aij / hh
                    else:
                        theta = 0.5 * hh / Aij
                        t = 1.0 / (abs(theta) + math.sqrt(1.0 + theta * theta))
                        if theta < 0.0:
                            t = -t
                    hh = t * Aij
                    Z[i] -= hh
                    Z[j] += hh
                    D[i] -= hh
                    D[j] += hh
                    A[p] = 0.0
                    c = 1.0 / math.sqrt(1 + t * t)
                    s = t * c
                    tau = s / (1.0 + c)
                    for k in range(0, i):
                        g, h = A[k * 4 + i], A[k * 4 + j]
                        A[k * 4 + i] = g - s * (h + g * tau)
                        A[k * 4 + j] = h + s * (g - h * tau)
                    for k in range(i + 1, j):
                        g, h = A[i * 4 + k], A[k * 4 + j]
                        A[i * 4 + k] = g - s * (h + g * tau)
                        A[k * 4 + j] = h + s * (g - h * tau)
                    for k in range(j + 1, 4):
                        g, h = A[i * 4 + k], A[j * 4 + k]
                        A[i * 4 + k] = g - s * (h + g * tau)
                        A[j * 4 + k] = h + s * (g - h * tau)
                    for k in range(4):
                        g, h = U[k * 4 + i], U[k * 4 + j]
                        U[k * 4 + i] = g - s * (h + g * tau)
                        U[k * 4 + j] = h + s * (g - h * tau)
                p += 1
        for i in range(4):
            B[i] += Z[i]
        D = list(B)
        Z = [0.0] * 4
    return D, U


def align(M_end, P0, P1, P2):
    """p3p::align: rigid transform (R, T) taking the world points P0..P2 onto the camera-frame points M_end[0..2]
    (Horn's closed form: quaternion = dominant eigenvector of the 4x4 matrix built from the cross-covariance)."""
    X0, Y0, Z0 = P0
    X1, Y1, Z1 = P1
    X2, Y2, Z2 = P2
    C_end = [(M_end[0][i] + M_end[1][i] + M_end[2][i]) / 3 for i in range(3)]
    C_start = [(X0 + X1 + X2) / 3, (Y0 + Y1 + Y2) / 3, (Z0 + Z1 + Z2) / 3]
    s = [0.0] * 9
    for j in range(3):
        s[0 * 3 + j] = (X0 * M_end[0][j] + X1 * M_end[1][j] + X2 * M_end[2][j]) / 3 - C_end[j] * C_start[0]
        s[1 * 3 + j] = (Y0 * M_end[0][j] + Y1 * M_end[1][j] + Y2 * M_end[2][j]) / 3 - C_end[j] * C_start[1]
        s[2 * 3 + j] = (Z0 * M_end[0][j] + Z1 * M_end[1][j] + Z2 * M_end[2][j]) / 3 - C_end[j] * C_start[2]
    Qs = [0.0] * 16
    Qs[0 * 4 + 0] = s[0 * 3 + 0] + s[1 * 3 + 1] + s[2 * 3 + 2]
    Qs[1 * 4 + 1] = s[0 * 3 + 0] - s[1 * 3 + 1] - s[2 * 3 + 2]
    Qs[2 * 4 + 2] = s[1 * 3 + 1] - s[2 * 3 + 2] - s[0 * 3 + 0]
    Qs[3 * 4 + 3] = s[2 * 3 + 2] - s[0 * 3 + 0] - s[1 * 3 + 1]
    Qs[1 * 4 + 0] = Qs[0 * 4 + 1] = s[1 * 3 + 2] - s[2 * 3 + 1]
    Qs[2 * 4 + 0] = Qs[0 * 4 + 2] = s[2 * 3 + 0] - s[0 * 3 + 2]
    Qs[3 * 4 + 0] = Qs[0 * 4 + 3] = s[0 * 3 + 1] - s[1 * 3 + 0]
    Qs[2 * 4 + 1] = Qs[1 * 4 + 2] = s[1 * 3 + 0] + s[0 * 3 + 1]
    Qs[3 * 4 + 1] = Qs[1 * 4 + 3] = s[2 * 3 + 0] + s[0 * 3 + 2]
    Qs[3 * 4 + 2] = Qs[2 * 4 + 3] = s[2 * 3 + 1] + s[1 * 3 + 2]
    evs, U = jacobi_4x4(Qs)
    i_ev = 0
    ev_max = evs[0]
    for i in range(1, 4):
        if evs[i] > ev_max:
            i_ev = i
            ev_max = evs[i]
    q = [U[i * 4 + i_ev] for i in range(4)]
    q02, q12, q22, q32 = q[0] * q[0], q[1] * q[1], q[2] * q[2], q[3] * q[3]
    q0_1, q0_2, q0_3 = q[0] * q[1], q[0] * q[2], q[0] * q[3]
    q1_2, q1_3, q2_3 = q[1] * q[2], q[1] * q[3], q[2] * q[3]
    R = [[q02 + q12 - q22 - q32, 2.0 * (q1_2 - q0_3), 2.0 * (q1_3 + q0_2)],
         [2.0 * (q1_2 + q0_3), q02 + q22 - q12 - q32, 2.0 * (q2_3 - q0_1)],
         [2.0 * (q1_3 - q0_2), 2.0 * (q2_3 + q0_1), q02 + q32 - q12 - q22]]
    T = [C_end[i] - (R[i][0] * C_start[0] + R[i][1] * C_start[1] + R[i][2] * C_start[2]) for i in range(3)]
    return R, T


def solve_for_lengths(distances, cosines):
    """p3p::solve_for_lengths: |PA|, |PB|, |PC| for up to four real configurations (Gao et al., main branch)."""
    p = cosines[0] * 2
    q = cosines[1] * 2
    r = cosines[2] * 2
    inv_d22 = 1.0 / (distances[2] * distances[2])
    a = inv_d22 * (distances[0] * distances[0])
    b = inv_d22 * (distances[1] * distances[1])
    a2, b2, p2, q2, r2 = a * a, b * b, p * p, q * q, r * r
    pr = p * r
    pqr = q * pr
    if p2 + q2 + r2 - pqr - 1 == 0:
        return []
    ab = a * b
    a_2 = 2 * a
    A = -2 * b + b2 + a2 + 1 + ab * (2 - r2) - a_2
    if A == 0:
        return []
    a_4 = 4 * a
    B = q * (-2 * (ab + a2 + 1 - b) + r2 * ab + a_4) + pr * (b - b2 + ab)
    C = q2 + b2 * (r2 + p2 - 2) - b * (p2 + pqr) - ab * (r2 + pqr) + (a2 - a_2) * (2 + q2) + 2
    D = pr * (ab - b2 + b) + q * ((p2 - 2) * b + 2 * (ab - a2) + a_4 - 2)
    E = 1 + 2 * (b - a - ab) + b2 - b * p2 + a2
    temp = (p2 * (a - 1 + b) + r2 * (a - 1 - b) + pqr - a * pqr)
    b0 = b * temp * temp
    if b0 == 0:
        return []
    roots = solve_deg4(A, B, C, D, E)
    if not roots:
        return []
    r3 = r2 * r
    pr2 = p * r2
    r3q = r3 * q
    inv_b0 = 1.0 / b0
    out = []
    for x in roots:
        if not (x > 0):
            continue
        x2 = x * x
        b1 = (((1 - a - b) * x2 + (q * a - q) * x + 1 - a + b) *
              (((r3 * (a2 + ab * (2 - r2) - a_2 + b2 - 2 * b + 1)) * x +
                (r3q * (2 * (b - a2) + a_4 + ab * (r2 - 2) - 2) + pr2 * (1 + a2 + 2 * (ab - a - b) + r2 * (b - b2) + b2))) * x2 +
               (r3 * (q2 * (1 - 2 * a + a2) + r2 * (b2 - ab) - a_4 + 2 * (a2 - b2) + 2) + r * p2 * (b2 + 2 * (ab - b - a) + 1 + a2) +
                pr2 * q * (a_4 + 2 * (b - ab - a2) - 2 - r2 * b)) * x +
               2 * r3q * (a_2 - b - a2 + ab - 1) + pr2 * (q2 - a_4 + 2 * (a2 - b2) + r2 * b + q2 * (a2 - a_2) + 2) +
               p2 * (p * (2 * (ab - a - b) + a2 + b2 + 1) + 2 * q * r * (b + a_2 - a2 - ab - 1))))
        if not (b1 > 0):
            continue
        y = inv_b0 * b1
        v = x2 + y * y - x * y * r
        if not (v > 0):
            continue
        Z = distances[2] / math.sqrt(v)
        out.append((x * Z, y * Z, Z))
    return out


def normalized_f32(img, fx, fy, cx, cy):
    """cv::undistortPoints with zero distortion on CV_32FC2 input: (u - cx) * (1/fx) in double, stored back as f32."""
    img = np.asarray(img, np.float32).reshape(-1, 2)
    ifx, ify = 1.0 / fx, 1.0 / fy
    return [(float(np.float32((float(u) - cx) * ifx)), float(np.float32((float(v) - cy) * ify))) for u, v in img]


def solve_p3p_3(obj3, mu, fx, fy, cx, cy):
    """p3p::solve (three points): list of (R, t) candidates.  `mu`: three pixel positions (double)."""
    inv_fx, inv_fy, cx_fx, cy_fy = 1.0 / fx, 1.0 / fy, cx / fx, cy / fy
    rays = []
    for (u, v) in mu:
        u = inv_fx * u - cx_fx
        v = inv_fy * v - cy_fy
        norm = math.sqrt(u * u + v * v + 1)
        k = 1.0 / norm
        rays.append((u * k, v * k, k))
    P = [tuple(float(x) for x in o) for o in obj3]

    def dist(i, j):
        return math.sqrt((P[i][0] - P[j][0]) * (P[i][0] - P[j][0]) + (P[i][1] - P[j][1]) * (P[i][1] - P[j][1]) +
                         (P[i][2] - P[j][2]) * (P[i][2] - P[j][2]))

    def cosang(i, j):
        return rays[i][0] * rays[j][0] + rays[i][1] * rays[j][1] + rays[i][2] * rays[j][2]

    distances = [dist(1, 2), dist(0, 2), dist(0, 1)]
    cosines = [cosang(1, 2), cosang(0, 2), cosang(0, 1)]
    sols = []
    for L in solve_for_lengths(distances, cosines):
        M = [[L[i] * rays[i][0], L[i] * rays[i][1], L[i] * rays[i][2]] for i in range(3)]
        sols.append(align(M, P[0], P[1], P[2]))
    return sols


def solve_p3p(obj4, img4, fx, fy, cx, cy):
    """cv::solvePnP(flags=SOLVEPNP_P3P) on exactly four correspondences (OpenCV 3.2 form): the first three give
    the candidates, the fourth picks the one with the smallest squared pixel error.  obj4 [4,3], img4 [4,2] f32.
    Returns (ok, R 3x3 list, t list)."""
    nrm = normalized_f32(img4, fx, fy, cx, cy)
    px = [(x * fx + cx, y * fy + cy) for x, y in nrm]          # p3p::extract_points
    obj = np.asarray(obj4, np.float32).reshape(4, 3)
    sols = solve_p3p_3(obj[:3], px[:3], fx, fy, cx, cy)
    if not sols:
        return False, None, None
    X3, Y3, Z3 = (float(v) for v in obj[3])
    mu3, mv3 = px[3]
    ns, min_reproj = 0, 0.0
    for i, (R, t) in enumerate(sols):
        X3p = R[0][0] * X3 + R[0][1] * Y3 + R[0][2] * Z3 + t[0]
        Y3p = R[1][0] * X3 + R[1][1] * Y3 + R[1][2] * Z3 + t[1]
        Z3p = R[2][0] * X3 + R[2][1] * Y3 + R[2][2] * Z3 + t[2]
        mu3p = cx + fx * X3p / Z3p
        mv3p = cy + fy * Y3p / Z3p
        reproj = (mu3p - mu3) * (mu3p - mu3) + (mv3p - mv3) * (mv3p - mv3)
        if i == 0 or min_reproj > reproj:
            ns, min_reproj = i, reproj
    return True, sols[ns][0], sols[ns][1]
