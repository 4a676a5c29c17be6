// 4-point minimal solver of K3: what cv::solvePnP(SOLVEPNP_P3P) returns for exactly four correspondences, the
// RANSAC kernel cv::solvePnPRansac switches to when `flags` is P3P / AP3P (R/src/myRegistrationVis.cpp:103 parses
// Vis/PnPFlags into _PnPFlags, :1084 hands it on) or when exactly four correspondences arrive (`npoints == 4`).
//
// OpenCV's P3P is NOT replayed operation for operation, unlike EPnP (epnp.cuh): the solver differs between the
// OpenCV 3.2 of the reference image (Gao's quartic; restated in oracle/cv_p3p.py, it loses up to several pixels on
// ~1 % of random sets) and the cv2 4.13 the parity tests are pinned to (accurate to the f32 rounding of the
// normalised image points).  P3P has no rank-deficient step -- each real configuration is an isolated root -- so any
// solver that is exact to rounding yields the same poses to ~1e-9 and the same inlier sets.  This one is Persson &
// Nordberg's "Lambda Twist" (ECCV 2018): the three distance constraints form two conics D1, D2 in the depths; the
// cubic det(D1 + g D2) = 0 gives a degenerate conic that factors into two lines, each line intersected with one
// constraint is a quadratic in a depth ratio -> up to four depth triples, polished by Gauss-Newton on the three
// constraints, then R, t in closed form from the two triangles.  The fourth correspondence picks the candidate
// (smallest squared reprojection error over the four points, as cv::solveP3P orders its results).
#pragma once
#include "cvmath.cuh"

namespace cvm {
namespace p3p_detail {

CVM_HD inline void cross3(const double* a, const double* b, double* c) {
  c[0] = a[1] * b[2] - a[2] * b[1];
  c[1] = a[2] * b[0] - a[0] * b[2];
  c[2] = a[0] * b[1] - a[1] * b[0];
}

// real roots of x^2 + b x + c (cancellation-free form); false if complex
CVM_HD inline bool root2real(double b, double c, double& r1, double& r2) {
  const double v = b * b - 4.0 * c;
  if (v < 0) {
    r1 = r2 = 0.5 * b;
    return false;
  }
  const double y = sqrt(v);
  if (b < 0) {
    r1 = 0.5 * (-b + y);
    r2 = 0.5 * (-b - y);
  } else {
    r1 = 2.0 * c / (-b + y);
    r2 = 2.0 * c / (-b - y);
  }
  return true;
}

// one real root of r^3 + b r^2 + c r + d, started where the cubic is steep and polished by Newton's method
CVM_HD inline double cubick(double b, double c, double d) {
  double r0;
  if (b * b >= 3.0 * c) {
    const double v = sqrt(b * b - 3.0 * c);
    const double t1 = (-b - v) / 3.0;
    double k = ((t1 + b) * t1 + c) * t1 + d;
    if (k > 0.0) {
      r0 = t1 - sqrt(-k / (3.0 * t1 + b));
    } else {
      const double t2 = (-b + v) / 3.0;
      k = ((t2 + b) * t2 + c) * t2 + d;
      r0 = t2 + sqrt(-k / (3.0 * t2 + b));
    }
  } else {
    r0 = -b / 3.0;
    if (fabs((3.0 * r0 + 2.0 * b) * r0 + c) < 1e-4) r0 += 1;
  }
  CVM_NOUNROLL
  for (int cnt = 0; cnt < 50; ++cnt) {
    const double fx = ((r0 + b) * r0 + c) * r0 + d;
    if (cnt < 7 || fabs(fx) > 1e-13) {
      const double fpx = (3.0 * r0 + 2.0 * b) * r0 + c;
      r0 -= fx / fpx;
    } else {
      break;
    }
  }
  return r0;
}

// eigen-decomposition of a symmetric 3x3 (row-major x) that has one zero eigenvalue: E columns = eigenvectors of
// L[0], L[1] (|L0| >= |L1|) and of 0
CVM_HD inline void eig_with_known0(const double* x, double* E, double* L) {
  L[2] = 0;
  double v3[3] = {x[3] * x[7] - x[6] * x[4], x[6] * x[1] - x[7] * x[0], x[4] * x[0] - x[3] * x[1]};
  const double n3 = 1.0 / sqrt(v3[0] * v3[0] + v3[1] * v3[1] + v3[2] * v3[2]);
  v3[0] *= n3; v3[1] *= n3; v3[2] *= n3;
  const double x01_sq = x[1] * x[1];
  const double b = -x[0] - x[4] - x[8];
  const double c = -x01_sq - x[2] * x[2] - x[5] * x[5] + x[0] * (x[4] + x[8]) + x[4] * x[8];
  double e1, e2;
  root2real(b, c, e1, e2);
  if (fabs(e1) < fabs(e2)) {
    const double tmp = e1;
    e1 = e2;
    e2 = tmp;
  }
  L[0] = e1;
  L[1] = e2;
  const double mx0011 = -x[0] * x[4];
  const double prec_0 = x[1] * x[5] - x[2] * x[4];
  const double prec_1 = x[1] * x[2] - x[0] * x[5];
  for (int k = 0; k < 2; ++k) {
    const double e = k ? e2 : e1;
    const double tmp = 1.0 / (e * (x[0] + x[4]) + mx0011 - e * e + x01_sq);
    double a1 = -(e * x[2] + prec_0) * tmp;
    double a2 = -(e * x[5] + prec_1) * tmp;
    const double rnorm = 1.0 / sqrt(a1 * a1 + a2 * a2 + 1.0);
    a1 *= rnorm;
    a2 *= rnorm;
    E[k] = a1;
    E[3 + k] = a2;
    E[6 + k] = rnorm;
  }
  E[2] = v3[0];
  E[5] = v3[1];
  E[8] = v3[2];
}

// Gauss-Newton on the three distance constraints r_ij(l) = l_i^2 + l_j^2 + b_ij l_i l_j - a_ij
CVM_HD inline void refine_lengths(double* Lv, double a12, double a13, double a23, double b12, double b13, double b23) {
  CVM_NOUNROLL
  for (int it = 0; it < 5; ++it) {
    const double l1 = Lv[0], l2 = Lv[1], l3 = Lv[2];
    const double r1 = l1 * l1 + l2 * l2 + b12 * l1 * l2 - a12;
    const double r2 = l1 * l1 + l3 * l3 + b13 * l1 * l3 - a13;
    const double r3 = l2 * l2 + l3 * l3 + b23 * l2 * l3 - a23;
    if (fabs(r1) + fabs(r2) + fabs(r3) < 1e-10) break;
    const double v0 = 2.0 * l1 + b12 * l2, v1 = 2.0 * l2 + b12 * l1;
    const double v3 = 2.0 * l1 + b13 * l3, v5 = 2.0 * l3 + b13 * l1;
    const double v7 = 2.0 * l2 + b23 * l3, v8 = 2.0 * l3 + b23 * l2;
    const double det = 1.0 / (-v0 * v5 * v7 - v1 * v3 * v8);
    const double n1 = l1 - det * (-v5 * v7 * r1 - v1 * v8 * r2 + v1 * v5 * r3);
    const double n2 = l2 - det * (-v3 * v8 * r1 + v0 * v8 * r2 - v0 * v5 * r3);
    const double n3 = l3 - det * (v3 * v7 * r1 - v0 * v7 * r2 - v1 * v3 * r3);
    const double q1 = n1 * n1 + n2 * n2 + b12 * n1 * n2 - a12;
    const double q2 = n1 * n1 + n3 * n3 + b13 * n1 * n3 - a13;
    const double q3 = n2 * n2 + n3 * n3 + b23 * n2 * n3 - a23;
    if (fabs(q1) + fabs(q2) + fabs(q3) > fabs(r1) + fabs(r2) + fabs(r3)) break;
    Lv[0] = n1;
    Lv[1] = n2;
    Lv[2] = n3;
  }
}

}  // namespace p3p_detail

// y: three UNIT bearing vectors (camera frame), x: the three world points.  Up to four (R row-major, t) with
// l_i y_i = R x_i + t.  Returns the number of solutions.
CVM_HD CVM_NOINLINE inline int p3p_lambdatwist(const double (*y)[3], const double (*x)[3], double (*Rs)[9], double (*ts)[3]) {
  using namespace p3p_detail;
  const double b12 = -2.0 * (y[0][0] * y[1][0] + y[0][1] * y[1][1] + y[0][2] * y[1][2]);
  const double b13 = -2.0 * (y[0][0] * y[2][0] + y[0][1] * y[2][1] + y[0][2] * y[2][2]);
  const double b23 = -2.0 * (y[1][0] * y[2][0] + y[1][1] * y[2][1] + y[1][2] * y[2][2]);
  double d12[3], d13[3], d23[3], d12xd13[3];
  for (int i = 0; i < 3; ++i) {
    d12[i] = x[0][i] - x[1][i];
    d13[i] = x[0][i] - x[2][i];
    d23[i] = x[1][i] - x[2][i];
  }
  cross3(d12, d13, d12xd13);
  const double a12 = d12[0] * d12[0] + d12[1] * d12[1] + d12[2] * d12[2];
  const double a13 = d13[0] * d13[0] + d13[1] * d13[1] + d13[2] * d13[2];
  const double a23 = d23[0] * d23[0] + d23[1] * d23[1] + d23[2] * d23[2];
  const double c31 = -0.5 * b13, c23 = -0.5 * b23, c12 = -0.5 * b12;
  const double blob = c12 * c23 * c31 - 1.0;
  const double s31_sq = 1.0 - c31 * c31, s23_sq = 1.0 - c23 * c23, s12_sq = 1.0 - c12 * c12;
  double p3 = a13 * (a23 * s31_sq - a13 * s23_sq);
  double p2 = 2.0 * blob * a23 * a13 + a13 * (2.0 * a12 + a13) * s23_sq + a23 * (a23 - a12) * s31_sq;
  double p1 = a23 * (a13 - a23) * s12_sq - a12 * a12 * s23_sq - 2.0 * a12 * (blob * a23 + a13 * s23_sq);
  double p0 = a12 * (a12 * s23_sq - a23 * s12_sq);
  p3 = 1.0 / p3;
  p2 *= p3;
  p1 *= p3;
  p0 *= p3;
  const double g = cubick(p2, p1, p0);
  const double A[9] = {a23 * (1.0 - g), (a23 * b12) * 0.5, (a23 * b13 * g) * (-0.5),
                       (a23 * b12) * 0.5, a23 - a12 + a13 * g, b23 * (a13 * g - a12) * 0.5,
                       (a23 * b13 * g) * (-0.5), b23 * (a13 * g - a12) * 0.5, g * (a13 - a23) - a12};
  double V[9], L[3];
  eig_with_known0(A, V, L);
  const double vq = -L[1] / L[0];
  const double v = sqrt(vq > 0 ? vq : 0.0);
  int valid = 0;
  double Ls[4][3];
  for (int sgn = 0; sgn < 2; ++sgn) {
    const double s = sgn ? -v : v;
    const double w2 = 1.0 / (s * V[1] - V[0]);
    const double w0 = (V[3] - s * V[4]) * w2;
    const double w1 = (V[6] - s * V[7]) * w2;
    const double a = 1.0 / ((a13 - a12) * w1 * w1 - a12 * b13 * w1 - a12);
    const double b = (a13 * b12 * w1 - a12 * b13 * w0 - 2.0 * w0 * w1 * (a12 - a13)) * a;
    const double c = ((a13 - a12) * w0 * w0 + a13 * b12 * w0 + a13) * a;
    if (b * b - 4.0 * c >= 0) {
      double tau[2];
      root2real(b, c, tau[0], tau[1]);
      for (int k = 0; k < 2; ++k) {
        if (tau[k] > 0 && valid < 4) {
          const double d = a23 / (tau[k] * (b23 + tau[k]) + 1.0);
          const double l2 = sqrt(d);
          const double l3 = tau[k] * l2;
          const double l1 = w0 * l2 + w1 * l3;
          if (l1 >= 0) {
            Ls[valid][0] = l1;
            Ls[valid][1] = l2;
            Ls[valid][2] = l3;
            ++valid;
          }
        }
      }
    }
  }
  // polish; a triple that does not satisfy the three constraints afterwards is no configuration at all (the cubic's
  // root or the line split degenerated: cv2 hands back a NaN model there, which scores no inlier either)
  {
    int kept = 0;
    const double tol = 1e-9 * (a12 + a13 + a23);
    for (int i = 0; i < valid; ++i) {
      refine_lengths(Ls[i], a12, a13, a23, b12, b13, b23);
      const double l1 = Ls[i][0], l2 = Ls[i][1], l3 = Ls[i][2];
      const double res = fabs(l1 * l1 + l2 * l2 + b12 * l1 * l2 - a12) + fabs(l1 * l1 + l3 * l3 + b13 * l1 * l3 - a13) +
                         fabs(l2 * l2 + l3 * l3 + b23 * l2 * l3 - a23);
      if (!(res <= tol)) continue;
      Ls[kept][0] = l1;
      Ls[kept][1] = l2;
      Ls[kept][2] = l3;
      ++kept;
    }
    valid = kept;
  }
  // X = [d12 d13 d12xd13]^-1 (columns), by cofactors
  double Xi[9];
  {
    const double m[9] = {d12[0], d13[0], d12xd13[0], d12[1], d13[1], d12xd13[1], d12[2], d13[2], d12xd13[2]};
    const double det = m[0] * (m[4] * m[8] - m[5] * m[7]) - m[1] * (m[3] * m[8] - m[5] * m[6]) + m[2] * (m[3] * m[7] - m[4] * m[6]);
    const double id = 1.0 / det;
    Xi[0] = (m[4] * m[8] - m[5] * m[7]) * id;
    Xi[1] = (m[2] * m[7] - m[1] * m[8]) * id;
    Xi[2] = (m[1] * m[5] - m[2] * m[4]) * id;
    Xi[3] = (m[5] * m[6] - m[3] * m[8]) * id;
    Xi[4] = (m[0] * m[8] - m[2] * m[6]) * id;
    Xi[5] = (m[2] * m[3] - m[0] * m[5]) * id;
    Xi[6] = (m[3] * m[7] - m[4] * m[6]) * id;
    Xi[7] = (m[1] * m[6] - m[0] * m[7]) * id;
    Xi[8] = (m[0] * m[4] - m[1] * m[3]) * id;
  }
  int n_out = 0;
  for (int i = 0; i < valid; ++i) {
    double ry1[3], yd1[3], yd2[3], yx[3];
    for (int k = 0; k < 3; ++k) {
      ry1[k] = y[0][k] * Ls[i][0];
      yd1[k] = ry1[k] - y[1][k] * Ls[i][1];
      yd2[k] = ry1[k] - y[2][k] * Ls[i][2];
    }
    cross3(yd1, yd2, yx);
    double* R = Rs[n_out];
    bool finite = true;
    for (int r = 0; r < 3; ++r)
      for (int c = 0; c < 3; ++c) {
        R[3 * r + c] = yd1[r] * Xi[c] + yd2[r] * Xi[3 + c] + yx[r] * Xi[6 + c];
        finite = finite && (R[3 * r + c] == R[3 * r + c]) && fabs(R[3 * r + c]) < 1e300;
      }
    if (!finite) continue;   // degenerate (collinear world points, repeated rays)
    for (int r = 0; r < 3; ++r)
      ts[n_out][r] = ry1[r] - (R[3 * r] * x[0][0] + R[3 * r + 1] * x[0][1] + R[3 * r + 2] * x[0][2]);
    ++n_out;
  }
  return n_out;
}

// cv::solvePnP(SOLVEPNP_P3P) on four correspondences.  o: 4 object points (f32 xyz), im: 4 image points (f32 pixels).
// undistortPoints (no distortion) stores the normalised points as f32, like the EPnP path.  false: no solution.
CVM_HD CVM_NOINLINE inline bool p3p4(const float* o, const float* im, const Intrinsics& K, double* R, double* t) {
  const double ifx = 1. / K.fx, ify = 1. / K.fy;
  double y[3][3], x[3][3], nrm[4][2];
  for (int i = 0; i < 4; ++i) {
    nrm[i][0] = (double)(float)(((double)im[2 * i] - K.cx) * ifx);
    nrm[i][1] = (double)(float)(((double)im[2 * i + 1] - K.cy) * ify);
  }
  for (int i = 0; i < 3; ++i) {
    const double n = 1.0 / sqrt(nrm[i][0] * nrm[i][0] + nrm[i][1] * nrm[i][1] + 1.0);
    y[i][0] = nrm[i][0] * n;
    y[i][1] = nrm[i][1] * n;
    y[i][2] = n;
    for (int k = 0; k < 3; ++k) x[i][k] = (double)o[3 * i + k];
  }
  double Rs[4][9], ts[4][3];
  const int n = p3p_lambdatwist(y, x, Rs, ts);
  int best = -1;
  double best_err = 0;
  for (int s = 0; s < n; ++s) {
    double err = 0;
    for (int i = 0; i < 4; ++i) {
      const double X = (double)o[3 * i], Y = (double)o[3 * i + 1], Z = (double)o[3 * i + 2];
      const double xc = Rs[s][0] * X + Rs[s][1] * Y + Rs[s][2] * Z + ts[s][0];
      const double yc = Rs[s][3] * X + Rs[s][4] * Y + Rs[s][5] * Z + ts[s][1];
      const double zc = Rs[s][6] * X + Rs[s][7] * Y + Rs[s][8] * Z + ts[s][2];
      const double iz = zc ? 1. / zc : 1;
      const double du = (double)im[2 * i] - (xc * iz * K.fx + K.cx), dv = (double)im[2 * i + 1] - (yc * iz * K.fy + K.cy);
      err += du * du + dv * dv;
    }
    if (!(err == err)) continue;
    if (best < 0 || err < best_err) {
      best = s;
      best_err = err;
    }
  }
  if (best < 0) return false;
  for (int i = 0; i < 9; ++i) R[i] = Rs[best][i];
  for (int i = 0; i < 3; ++i) t[i] = ts[best][i];
  return true;
}

}  // namespace cvm
