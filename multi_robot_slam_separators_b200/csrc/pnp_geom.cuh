// cv::Rodrigues (both directions, with the Jacobian the ITERATIVE refit needs) and the
// reprojection-error arithmetic of PnPRansacCallback::computeError (cv::projectPoints -> f32,
// err = dx*dx + dy*dy in f32).  Same replay rules as cvmath.cuh.
#pragma once
#include "cvmath.cuh"

namespace cvm {

// 3x3 -> rvec: orthonormalise through SVD (R = U Vt), then axis-angle.  scratch: >= 21 doubles.
CVM_HD CVM_NOINLINE inline void rodrigues_mat_to_vec(Heap scratch, const double* Rin, double* r) {
  Heap At = scratch.at(0), W = scratch.at(9), Vt = scratch.at(12);
  for (int i = 0; i < 3; ++i)
    for (int c = 0; c < 3; ++c) At[c * 3 + i] = Rin[3 * i + c];
  jacobi_svd<3, 3, true>(At, W, Vt);
  double R[9];
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j)
      R[3 * i + j] = At[0 * 3 + i] * Vt[0 * 3 + j] + At[1 * 3 + i] * Vt[1 * 3 + j] + At[2 * 3 + i] * Vt[2 * 3 + j];
  double rx = R[7] - R[5], ry = R[2] - R[6], rz = R[3] - R[1];
  const double s = sqrt((rx * rx + ry * ry + rz * rz) * 0.25);
  double c = (R[0] + R[4] + R[8] - 1) * 0.5;
  c = c > 1. ? 1. : c < -1. ? -1. : c;
  double theta = acos(c);
  if (s < 1e-5) {
    if (c > 0) {
      rx = ry = rz = 0;
    } else {
      double t = (R[0] + 1) * 0.5;
      rx = sqrt(t > 0. ? t : 0.);
      t = (R[4] + 1) * 0.5;
      ry = sqrt(t > 0. ? t : 0.) * (R[1] < 0 ? -1. : 1.);
      t = (R[8] + 1) * 0.5;
      rz = sqrt(t > 0. ? t : 0.) * (R[2] < 0 ? -1. : 1.);
      if (fabs(rx) < fabs(ry) && fabs(rx) < fabs(rz) && (R[5] > 0) != (ry * rz > 0)) rz = -rz;
      theta /= sqrt(rx * rx + ry * ry + rz * rz);
      rx *= theta; ry *= theta; rz *= theta;
    }
  } else {
    double vth = 1 / (2 * s);
    vth *= theta;
    rx *= vth; ry *= vth; rz *= vth;
  }
  r[0] = rx; r[1] = ry; r[2] = rz;
}

// rvec -> 3x3 (row-major); if J != nullptr also dR/dr as 3 rows of 9 (cv layout)
CVM_HD inline void rodrigues_vec_to_mat(const double* rv, double* R, double* J) {
  double rx = rv[0], ry = rv[1], rz = rv[2];
  const double theta = sqrt(rx * rx + ry * ry + rz * rz);
  if (theta < kDblEps) {
    for (int i = 0; i < 9; ++i) R[i] = (i % 4 == 0) ? 1. : 0.;
    if (J) {
      for (int i = 0; i < 27; ++i) J[i] = 0;
      J[5] = J[15] = J[19] = -1;
      J[7] = J[11] = J[21] = 1;
    }
    return;
  }
  const double c = cos(theta), s = sin(theta), c1 = 1. - c, itheta = theta ? 1. / theta : 0.;
  rx *= itheta; ry *= itheta; rz *= itheta;
  const double rrt[9] = {rx * rx, rx * ry, rx * rz, rx * ry, ry * ry, ry * rz, rx * rz, ry * rz, rz * rz};
  const double r_x[9] = {0, -rz, ry, rz, 0, -rx, -ry, rx, 0};
  for (int k = 0; k < 9; ++k) R[k] = (c * ((k % 4 == 0) ? 1. : 0.) + c1 * rrt[k]) + s * r_x[k];
  if (J) {
    const double drrt[27] = {rx + rx, ry, rz, ry, 0, 0, rz, 0, 0, 0, rx, 0, rx, ry + ry, rz, 0, rz, 0,
                             0, 0, rx, 0, 0, ry, rx, ry, rz + rz};
    const double d_r_x[27] = {0, 0, 0, 0, 0, -1, 0, 1, 0, 0, 0, 1, 0, 0, 0, -1, 0, 0, 0, -1, 0, 1, 0, 0, 0, 0, 0};
    for (int i = 0; i < 3; ++i) {
      const double ri = i == 0 ? rx : i == 1 ? ry : rz;
      const double a0 = -s * ri, a1 = (s - 2 * c1 * itheta) * ri, a2 = c1 * itheta;
      const double a3 = (c - s * itheta) * ri, a4 = s * itheta;
      for (int k = 0; k < 9; ++k)
        J[i * 9 + k] = a0 * ((k % 4 == 0) ? 1. : 0.) + a1 * rrt[k] + a2 * drrt[i * 9 + k] + a3 * r_x[k] + a4 * d_r_x[i * 9 + k];
    }
  }
}

// cv::projectPoints for one point, zero distortion, double internals.
CVM_HD inline void project_point(const double* R, const double* t, const Intrinsics& K, double X, double Y, double Z,
                                 double& u, double& v) {
  double x = R[0] * X + R[1] * Y + R[2] * Z + t[0];
  double y = R[3] * X + R[4] * Y + R[5] * Z + t[1];
  double z = R[6] * X + R[7] * Y + R[8] * Z + t[2];
  z = z ? 1. / z : 1;
  x *= z;
  y *= z;
  u = x * K.fx + K.cx;
  v = y * K.fy + K.cy;
}

// f32 squared reprojection error exactly as computeError forms it
CVM_HD inline float reproj_err_f32(const double* R, const double* t, const Intrinsics& K, const float* o3,
                                   const float* i2) {
  double u, v;
  project_point(R, t, K, (double)o3[0], (double)o3[1], (double)o3[2], u, v);
  const float dx = i2[0] - (float)u, dy = i2[1] - (float)v;
#if defined(__CUDA_ARCH__)
  return __fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy));
#else
  return dx * dx + dy * dy;
#endif
}

// cv::RANSACUpdateNumIters
CVM_HD inline int ransac_update_num_iters(double p, double ep, int model_points, int max_iters) {
  p = p > 0. ? p : 0.;
  p = p < 1. ? p : 1.;
  ep = ep > 0. ? ep : 0.;
  ep = ep < 1. ? ep : 1.;
  double num = 1. - p > kDblMin ? 1. - p : kDblMin;
  double denom = 1. - pow(1. - ep, (double)model_points);
  if (denom < kDblMin) return 0;
  num = log(num);
  denom = log(denom);
  return denom >= 0 || -num >= max_iters * (-denom) ? max_iters : (int)rint(num / denom);
}

}  // namespace cvm
