// Arithmetic of the 2-view bundle adjustment (bundle_adjust.cu): g2o's SE3Quat / SBACam operations, the stereo / mono
// projection edge with its Jacobians, the per-point block of the normal equations, the small solves.  __host__ __device__
// so that tests/host_harness can check it on the CPU against oracle/ba.py; the kernel is the only product user.
#pragma once
#include <math.h>
#include <stdint.h>

#if defined(__CUDACC__)
#define BA_HD __host__ __device__
#define BA_HDI __host__ __device__ __forceinline__
#else
#define BA_HD
#define BA_HDI inline
#endif

namespace ba {

struct BaKernelParams {
  double fx, fy, cx, cy, baseline, var, delta;
  double LR[9], Lt[3];   // camera localTransform (optical -> base)
  int iterations, min_inliers, use_cov;
};

struct Se3 {
  double q[4];   // x, y, z, w
  double t[3];
};

struct Cam {
  Se3 p;          // camera -> world
  double R[9];    // rotation of p, row-major
  double rt[3];   // R^T t
};

// ---- Eigen / g2o::SE3Quat arithmetic ------------------------------------------------------------------------------
BA_HD inline void quat_from_matrix(const double* m, double* q) {   // Eigen::Quaterniond(Matrix3d)
  double t = m[0] + m[4] + m[8];
  if (t > 0) {
    t = sqrt(t + 1.0);
    q[3] = 0.5 * t;
    t = 0.5 / t;
    q[0] = (m[7] - m[5]) * t;
    q[1] = (m[2] - m[6]) * t;
    q[2] = (m[3] - m[1]) * t;
  } else {
    int i = 0;
    if (m[4] > m[0]) i = 1;
    if (m[8] > m[4 * i]) i = 2;
    const int j = (i + 1) % 3, k = (j + 1) % 3;
    t = sqrt(m[4 * i] - m[4 * j] - m[4 * k] + 1.0);
    q[i] = 0.5 * t;
    t = 0.5 / t;
    q[3] = (m[3 * k + j] - m[3 * j + k]) * t;
    q[j] = (m[3 * j + i] + m[3 * i + j]) * t;
    q[k] = (m[3 * k + i] + m[3 * i + k]) * t;
  }
}

BA_HD inline void quat_normalize_rotation(double* q) {   // SE3Quat::normalizeRotation
  const double s = (q[3] < 0 ? -1.0 : 1.0) / sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
  for (int i = 0; i < 4; ++i) q[i] *= s;
}

BA_HD inline void quat_to_matrix(const double* q, double* R) {   // Eigen::Quaterniond::toRotationMatrix
  const double x = q[0], y = q[1], z = q[2], w = q[3];
  const double tx = 2 * x, ty = 2 * y, tz = 2 * z;
  const double twx = tx * w, twy = ty * w, twz = tz * w, txx = tx * x, txy = ty * x, txz = tz * x, tyy = ty * y,
               tyz = tz * y, tzz = tz * z;
  R[0] = 1 - (tyy + tzz); R[1] = txy - twz; R[2] = txz + twy;
  R[3] = txy + twz; R[4] = 1 - (txx + tzz); R[5] = tyz - twx;
  R[6] = txz - twy; R[7] = tyz + twx; R[8] = 1 - (txx + tyy);
}

BA_HD inline void quat_mul(const double* a, const double* b, double* o) {
  const double x = a[3] * b[0] + a[0] * b[3] + a[1] * b[2] - a[2] * b[1];
  const double y = a[3] * b[1] + a[1] * b[3] + a[2] * b[0] - a[0] * b[2];
  const double z = a[3] * b[2] + a[2] * b[3] + a[0] * b[1] - a[1] * b[0];
  const double w = a[3] * b[3] - a[0] * b[0] - a[1] * b[1] - a[2] * b[2];
  o[0] = x; o[1] = y; o[2] = z; o[3] = w;
}

BA_HD inline void quat_rotate(const double* q, const double* v, double* o) {
  double R[9];
  quat_to_matrix(q, R);
  const double x = R[0] * v[0] + R[1] * v[1] + R[2] * v[2], y = R[3] * v[0] + R[4] * v[1] + R[5] * v[2],
               z = R[6] * v[0] + R[7] * v[1] + R[8] * v[2];
  o[0] = x; o[1] = y; o[2] = z;
}

BA_HD inline Se3 se3_mul(const Se3& a, const Se3& b) {   // SE3Quat::operator*
  Se3 r;
  double rt[3];
  quat_rotate(a.q, b.t, rt);
  for (int i = 0; i < 3; ++i) r.t[i] = a.t[i] + rt[i];
  quat_mul(a.q, b.q, r.q);
  quat_normalize_rotation(r.q);
  return r;
}

BA_HD inline Se3 se3_inv(const Se3& a) {
  Se3 r;
  r.q[0] = -a.q[0]; r.q[1] = -a.q[1]; r.q[2] = -a.q[2]; r.q[3] = a.q[3];
  const double nt[3] = {-a.t[0], -a.t[1], -a.t[2]};
  quat_rotate(r.q, nt, r.t);
  return r;
}

BA_HD inline void cam_finish(Cam& c) {
  quat_to_matrix(c.p.q, c.R);
  for (int k = 0; k < 3; ++k) c.rt[k] = c.R[k] * c.p.t[0] + c.R[3 + k] * c.p.t[1] + c.R[6 + k] * c.p.t[2];
}

// SBACam::update: t += u[0:3]; r *= (u[3:6], sqrt(1 - |u[3:6]|^2)) normalised; r normalised
BA_HD inline Se3 se3_update(const Se3& c, const double* u) {
  Se3 r;
  for (int i = 0; i < 3; ++i) r.t[i] = c.t[i] + u[i];
  double qr[4] = {u[3], u[4], u[5], sqrt(1.0 - (u[3] * u[3] + u[4] * u[4] + u[5] * u[5]))};   // NaN past the unit ball
  double s = 1.0 / sqrt(qr[0] * qr[0] + qr[1] * qr[1] + qr[2] * qr[2] + qr[3] * qr[3]);
  for (int i = 0; i < 4; ++i) qr[i] *= s;
  quat_mul(c.q, qr, r.q);
  s = 1.0 / sqrt(r.q[0] * r.q[0] + r.q[1] * r.q[1] + r.q[2] * r.q[2] + r.q[3] * r.q[3]);
  for (int i = 0; i < 4; ++i) r.q[i] *= s;
  return r;
}

// EdgeSBACam::computeError: delta = measurement^-1 * (cam1^-1 * cam2) -> (translation, quaternion vector)
BA_HD inline void link_error(const Se3& link_inv, const Se3& cam1_inv, const Se3& cam2, double* e) {
  const Se3 d = se3_mul(link_inv, se3_mul(cam1_inv, cam2));
  for (int i = 0; i < 3; ++i) { e[i] = d.t[i]; e[3 + i] = d.q[i]; }
}

// ---- one projection edge (EdgeProjectP2SC / P2MC) -----------------------------------------------------------------
struct Edge {
  double e[3];     // error (third component 0 for a mono edge)
  double Jpc[9];   // d(u, v, u_right) / d(camera-frame point)
  double pc[3];
  double chi;      // e^T Omega e
};

BA_HDI void edge_error(const BaKernelParams& P, const Cam& c, const double* X, const float* meas,
                                           bool stereo, Edge& g, bool jac) {
  // camera-frame point: w2n * (X, 1) = R^T X - R^T t
  const double px = c.R[0] * X[0] + c.R[3] * X[1] + c.R[6] * X[2] - c.rt[0];
  const double py = c.R[1] * X[0] + c.R[4] * X[1] + c.R[7] * X[2] - c.rt[1];
  const double pz = c.R[2] * X[0] + c.R[5] * X[1] + c.R[8] * X[2] - c.rt[2];
  const double iz = 1.0 / pz;
  g.pc[0] = px; g.pc[1] = py; g.pc[2] = pz;
  g.e[0] = P.fx * px * iz + P.cx - (double)meas[0];
  g.e[1] = P.fy * py * iz + P.cy - (double)meas[1];
  g.e[2] = stereo ? P.fx * (px - P.baseline) * iz + P.cx - (double)meas[2] : 0.0;
  g.chi = (g.e[0] * g.e[0] + g.e[1] * g.e[1] + g.e[2] * g.e[2]) / P.var;
  if (jac) {
    const double iz2 = iz * iz;
    g.Jpc[0] = P.fx * iz; g.Jpc[1] = 0; g.Jpc[2] = -P.fx * px * iz2;
    g.Jpc[3] = 0; g.Jpc[4] = P.fy * iz; g.Jpc[5] = -P.fy * py * iz2;
    g.Jpc[6] = stereo ? P.fx * iz : 0.0; g.Jpc[7] = 0; g.Jpc[8] = stereo ? -P.fx * (px - P.baseline) * iz2 : 0.0;
  }
}

// RobustKernelHuber::robustify -> rho (returned) and rho' (w)
BA_HDI double huber(double chi, double delta, double& w) {
  w = 1.0;
  if (!(delta > 0) || !(chi > delta * delta)) return chi;
  const double s = sqrt(chi);
  w = delta / s;
  return 2 * s * delta - delta * delta;
}

// the point's share of the linear system over its level-0 edges: V (3x3 symmetric: 00 01 02 11 12 22), bl, W (6x3);
// camera block A (21, upper triangle by rows) and bc are ADDED to; returns the robust chi2 of the edges
struct PtSys {
  double V[6], bl[3], W[18];
};

BA_HD inline double point_system(const BaKernelParams& P, const Cam& c1, const Cam& c2, const double* X, const float* meas,
                               unsigned flags /*bit0,1: level of edge 1,2; bit2,3: stereo*/, PtSys& s, double* A,
                               double* bc) {
  for (int i = 0; i < 6; ++i) s.V[i] = 0;
  for (int i = 0; i < 3; ++i) s.bl[i] = 0;
  for (int i = 0; i < 18; ++i) s.W[i] = 0;
  double chi_sum = 0;
#pragma unroll 1
  for (int c = 0; c < 2; ++c) {
    if ((flags >> c) & 1u) continue;
    const Cam& cam = c ? c2 : c1;
    Edge g;
    edge_error(P, cam, X, meas + 3 * c, (flags >> (2 + c)) & 1u, g, true);
    double w;
    chi_sum += huber(g.chi, P.delta, w);
    w /= P.var;
    double Jp[9];   // d / d(world point) = Jpc R^T
    for (int r = 0; r < 3; ++r)
      for (int j = 0; j < 3; ++j)
        Jp[3 * r + j] = g.Jpc[3 * r] * cam.R[3 * j] + g.Jpc[3 * r + 1] * cam.R[3 * j + 1] + g.Jpc[3 * r + 2] * cam.R[3 * j + 2];
    int k = 0;
    for (int i = 0; i < 3; ++i)
      for (int j = i; j < 3; ++j, ++k) s.V[k] += w * (Jp[i] * Jp[j] + Jp[3 + i] * Jp[3 + j] + Jp[6 + i] * Jp[6 + j]);
    for (int i = 0; i < 3; ++i) s.bl[i] -= w * (Jp[i] * g.e[0] + Jp[3 + i] * g.e[1] + Jp[6 + i] * g.e[2]);
    if (c == 1) {
      // d(camera-frame point) / d(quaternion vector) = 2 pc x e_k
      const double px = g.pc[0], py = g.pc[1], pz = g.pc[2];
      const double D[9] = {0, -2 * pz, 2 * py, 2 * pz, 0, -2 * px, -2 * py, 2 * px, 0};
      double Jc[18];   // [3][6] = [-Jp | Jpc D]
      for (int r = 0; r < 3; ++r)
        for (int j = 0; j < 3; ++j) {
          Jc[6 * r + j] = -Jp[3 * r + j];
          Jc[6 * r + 3 + j] = g.Jpc[3 * r] * D[j] + g.Jpc[3 * r + 1] * D[3 + j] + g.Jpc[3 * r + 2] * D[6 + j];
        }
      for (int i = 0; i < 6; ++i)
        for (int j = 0; j < 3; ++j)
          s.W[3 * i + j] += w * (Jc[i] * Jp[j] + Jc[6 + i] * Jp[3 + j] + Jc[12 + i] * Jp[6 + j]);
      if (A) {
        k = 0;
        for (int i = 0; i < 6; ++i)
          for (int j = i; j < 6; ++j, ++k) A[k] += w * (Jc[i] * Jc[j] + Jc[6 + i] * Jc[6 + j] + Jc[12 + i] * Jc[12 + j]);
        for (int i = 0; i < 6; ++i) bc[i] -= w * (Jc[i] * g.e[0] + Jc[6 + i] * g.e[1] + Jc[12 + i] * g.e[2]);
      }
    }
  }
  return chi_sum;
}

// (V + lambda I)^-1 by cofactors (Eigen's 3x3 inverse), symmetric storage
BA_HDI void inv3_sym(const double* V, double lam, double* I) {
  const double a = V[0] + lam, b = V[1], c = V[2], e = V[3] + lam, f = V[4], i = V[5] + lam;
  const double c00 = e * i - f * f, c01 = c * f - b * i, c02 = b * f - c * e;
  const double c11 = a * i - c * c, c12 = c * b - a * f, c22 = a * e - b * b;
  const double inv = 1.0 / (a * c00 + b * c01 + c * c02);
  I[0] = c00 * inv; I[1] = c01 * inv; I[2] = c02 * inv; I[3] = c11 * inv; I[4] = c12 * inv; I[5] = c22 * inv;
}

BA_HDI void sym3_mul(const double* S, const double* v, double* o) {
  o[0] = S[0] * v[0] + S[1] * v[1] + S[2] * v[2];
  o[1] = S[1] * v[0] + S[3] * v[1] + S[4] * v[2];
  o[2] = S[2] * v[0] + S[4] * v[1] + S[5] * v[2];
}

// dense 6x6 Cholesky solve (the reduced camera system); false when a pivot is not positive
BA_HD inline bool cholesky6(const double* S /*21, upper by rows*/, double lam, const double* rhs, double* x) {
  double L[36];
  int k = 0;
  for (int i = 0; i < 6; ++i)
    for (int j = i; j < 6; ++j, ++k) L[6 * j + i] = S[k] + (i == j ? lam : 0.0);   // lower triangle holds S
  for (int j = 0; j < 6; ++j) {
    double d = L[7 * j];
    for (int p = 0; p < j; ++p) d -= L[6 * j + p] * L[6 * j + p];
    if (!(d > 0)) return false;
    // one reciprocal per column instead of a division per entry: this solve is serial work of one thread while the CTA
    // waits (ncu: `barrier` is the kernel's top stall), and an FP64 division is a ~40-instruction dependent chain
    const double inv = 1.0 / sqrt(d);
    L[7 * j] = inv;   // the diagonal holds 1 / L_jj
    for (int i = j + 1; i < 6; ++i) {
      double v = L[6 * i + j];
      for (int p = 0; p < j; ++p) v -= L[6 * i + p] * L[6 * j + p];
      L[6 * i + j] = v * inv;
    }
  }
  double y[6];
  for (int i = 0; i < 6; ++i) {
    double v = rhs[i];
    for (int p = 0; p < i; ++p) v -= L[6 * i + p] * y[p];
    y[i] = v * L[7 * i];
  }
  for (int i = 5; i >= 0; --i) {
    double v = y[i];
    for (int p = i + 1; p < 6; ++p) v -= L[6 * p + i] * x[p];
    x[i] = v * L[7 * i];
  }
  return true;
}

}  // namespace ba
