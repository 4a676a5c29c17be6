// 5-point EPnP minimal solver: operation-for-operation replay of cv::solvePnP(SOLVEPNP_EPNP) as
// PnPRansacCallback::runKernel reaches it (undistortPoints -> epnp::compute_pose), see cvmath.cuh.
// Per-thread scratch lives in a `cvm::Heap` of kEpnpHeap doubles (174: five one-warp CTAs of K3 fit one SM):
//   [0,144)   12x12 M^T M / its Jacobi sweep; afterwards rows 8..11 ([96,144)) are the null-space
//             vectors, [0,60) holds L_6x10 and [60,96) is small-SVD scratch
//   [144,174) control points cws (4x3) until rho is formed, then W of the big SVD, then W / Vt of the small SVDs
// The barycentric alphas (5x4) live in the caller's local memory: they are read once per point while M^T M
// is accumulated and once per candidate pose.
#pragma once
#include "cvmath.cuh"

namespace cvm {

constexpr int kEpnpHeap = 174;
constexpr int kN = 5;  // model points

namespace detail {
constexpr int H_AT = 0, H_L = 0, H_SAT = 60, H_UT = 96, H_CWS = 144, H_W = 144, H_SW = 144, H_SVT = 149;

CVM_HD inline double dot3(const double* a, const double* b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }

// cvSolve(A[6 x NC], rho, x, CV_SVD) with A's columns picked from L_6x10
template <int NC>
CVM_HD CVM_NOINLINE inline void solve_cols(Heap h, const int (&cols)[NC], const double* rho, double* x) {
  Heap At = h.at(H_SAT), W = h.at(H_SW), Vt = h.at(H_SVT);
  for (int c = 0; c < NC; ++c)
    for (int r = 0; r < 6; ++r) At[c * 6 + r] = h[H_L + r * 10 + cols[c]];
  jacobi_svd<6, NC, true>(At, W, Vt);
  svd_backsubst<6, NC>(At, W, Vt, rho, x);
}

// epnp::qr_solve for the 6x4 Gauss-Newton system (A row-major 6x4, b[6]) -> X[4]; false if singular
CVM_HD inline bool qr_solve_6x4(double* A, double* b, double* X) {
  const int nr = 6, nc = 4;
  double A1[4], A2[4];
  CVM_UNROLL
  for (int k = 0; k < nc; ++k) {
    const int kk = k * nc + k;
    int p = kk;
    double eta = fabs(A[p]);
    CVM_UNROLL
    for (int i = k + 1; i < nr; ++i) {  // (sic) scans rows k..nr-2
      const double elt = fabs(A[p]);
      if (eta < elt) eta = elt;
      p += nc;
    }
    if (eta == 0) return false;
    const double inv_eta = 1. / eta;
    double sum2 = 0;
    p = kk;
    CVM_UNROLL
    for (int i = k; i < nr; ++i) {
      A[p] *= inv_eta;
      sum2 += A[p] * A[p];
      p += nc;
    }
    double sigma = sqrt(sum2);
    if (A[kk] < 0) sigma = -sigma;
    A[kk] += sigma;
    A1[k] = sigma * A[kk];
    A2[k] = -eta * sigma;
    CVM_UNROLL
    for (int j = k + 1; j < nc; ++j) {
      double sum = 0;
      p = kk;
      CVM_UNROLL
      for (int i = k; i < nr; ++i) {
        sum += A[p] * A[p + j - k];
        p += nc;
      }
      const double tau = sum / A1[k];
      p = kk;
      CVM_UNROLL
      for (int i = k; i < nr; ++i) {
        A[p + j - k] -= tau * A[p];
        p += nc;
      }
    }
  }
  CVM_UNROLL
  for (int j = 0; j < nc; ++j) {
    const int jj = j * nc + j;
    double tau = 0;
    int p = jj;
    CVM_UNROLL
    for (int i = j; i < nr; ++i) {
      tau += A[p] * b[i];
      p += nc;
    }
    tau /= A1[j];
    p = jj;
    CVM_UNROLL
    for (int i = j; i < nr; ++i) {
      b[i] -= tau * A[p];
      p += nc;
    }
  }
  X[nc - 1] = b[nc - 1] / A2[nc - 1];
  CVM_UNROLL
  for (int i = nc - 2; i >= 0; --i) {
    double sum = 0;
    int p = i * nc + (i + 1);
    CVM_UNROLL
    for (int j = i + 1; j < nc; ++j) {
      sum += A[p] * X[j];
      ++p;
    }
    X[i] = (b[i] - sum) / A2[i];
  }
  return true;
}

CVM_HD CVM_NOINLINE inline void gauss_newton(Heap h, const double* rho, double* betas) {
  double x_stale[4] = {0, 0, 0, 0};  // epnp::gauss_newton's local x[4] = {}
  CVM_NOUNROLL
  for (int it = 0; it < 5; ++it) {
    double A[24], b[6];
    CVM_UNROLL
    for (int i = 0; i < 6; ++i) {
      double r[10];
      CVM_UNROLL
      for (int c = 0; c < 10; ++c) r[c] = h[H_L + i * 10 + c];
      A[i * 4 + 0] = 2 * r[0] * betas[0] + r[1] * betas[1] + r[3] * betas[2] + r[6] * betas[3];
      A[i * 4 + 1] = r[1] * betas[0] + 2 * r[2] * betas[1] + r[4] * betas[2] + r[7] * betas[3];
      A[i * 4 + 2] = r[3] * betas[0] + r[4] * betas[1] + 2 * r[5] * betas[2] + r[8] * betas[3];
      A[i * 4 + 3] = r[6] * betas[0] + r[7] * betas[1] + r[8] * betas[2] + 2 * r[9] * betas[3];
      b[i] = rho[i] - (r[0] * betas[0] * betas[0] + r[1] * betas[0] * betas[1] + r[2] * betas[1] * betas[1] +
                       r[3] * betas[0] * betas[2] + r[4] * betas[1] * betas[2] + r[5] * betas[2] * betas[2] +
                       r[6] * betas[0] * betas[3] + r[7] * betas[1] * betas[3] + r[8] * betas[2] * betas[3] +
                       r[9] * betas[3] * betas[3]);
    }
    double x[4];
    if (qr_solve_6x4(A, b, x)) {
      CVM_UNROLL
      for (int i = 0; i < 4; ++i) x_stale[i] = x[i];
    }  // singular: qr_solve returns early and X keeps its previous content
    CVM_UNROLL
    for (int i = 0; i < 4; ++i) betas[i] += x_stale[i];
  }
}

// compute_ccs / compute_pcs / solve_for_sign / estimate_R_and_t / reprojection_error
CVM_HD CVM_NOINLINE inline double compute_R_and_t(Heap h, const double* alpha, const float* o, const double* us, const Intrinsics& K, const double* betas,
                                     double* R, double* t) {
  double ccs[4][3] = {{0, 0, 0}, {0, 0, 0}, {0, 0, 0}, {0, 0, 0}};
  for (int i = 0; i < 4; ++i) {
    const int row = H_UT + (3 - i) * 12;  // ut row 11 - i
    for (int j = 0; j < 4; ++j)
      for (int k = 0; k < 3; ++k) ccs[j][k] += betas[i] * h[row + 3 * j + k];
  }
  double pcs[kN][3];
  for (int i = 0; i < kN; ++i) {
    const double a0 = alpha[4 * i], a1 = alpha[4 * i + 1], a2 = alpha[4 * i + 2], a3 = alpha[4 * i + 3];
    for (int j = 0; j < 3; ++j) pcs[i][j] = a0 * ccs[0][j] + a1 * ccs[1][j] + a2 * ccs[2][j] + a3 * ccs[3][j];
  }
  if (pcs[0][2] < 0.0) {
    for (int i = 0; i < kN; ++i)
      for (int j = 0; j < 3; ++j) pcs[i][j] = -pcs[i][j];
  }
  double pc0[3] = {0, 0, 0}, pw0[3] = {0, 0, 0};
  for (int i = 0; i < kN; ++i)
    for (int j = 0; j < 3; ++j) {
      pc0[j] += pcs[i][j];
      pw0[j] += (double)o[3 * i + j];
    }
  for (int j = 0; j < 3; ++j) {
    pc0[j] /= kN;
    pw0[j] /= kN;
  }
  double abt[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
  for (int i = 0; i < kN; ++i)
    for (int j = 0; j < 3; ++j) {
      abt[3 * j] += (pcs[i][j] - pc0[j]) * ((double)o[3 * i] - pw0[0]);
      abt[3 * j + 1] += (pcs[i][j] - pc0[j]) * ((double)o[3 * i + 1] - pw0[1]);
      abt[3 * j + 2] += (pcs[i][j] - pc0[j]) * ((double)o[3 * i + 2] - pw0[2]);
    }
  // cvSVD(ABt, D, U, V, MODIFY_A): Jacobi on rows of ABt^T; U(i,c) = At[c][i], V(j,c) = Vt[c][j]
  Heap At = h.at(H_SAT), W = h.at(H_SW), Vt = h.at(H_SVT);
  for (int r = 0; r < 3; ++r)
    for (int c = 0; c < 3; ++c) At[c * 3 + r] = abt[3 * r + c];
  jacobi_svd<3, 3, true>(At, W, Vt);
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j)
      R[3 * i + j] = At[0 * 3 + i] * Vt[0 * 3 + j] + At[1 * 3 + i] * Vt[1 * 3 + j] + At[2 * 3 + i] * Vt[2 * 3 + j];
  const double det = R[0] * R[4] * R[8] + R[1] * R[5] * R[6] + R[2] * R[3] * R[7] - R[2] * R[4] * R[6] -
                     R[1] * R[3] * R[8] - R[0] * R[5] * R[7];
  if (det < 0) {
    R[6] = -R[6];
    R[7] = -R[7];
    R[8] = -R[8];
  }
  t[0] = pc0[0] - dot3(R, pw0);
  t[1] = pc0[1] - dot3(R + 3, pw0);
  t[2] = pc0[2] - dot3(R + 6, pw0);
  double sum2 = 0;
  for (int i = 0; i < kN; ++i) {
    const double pw[3] = {(double)o[3 * i], (double)o[3 * i + 1], (double)o[3 * i + 2]};
    const double Xc = dot3(R, pw) + t[0];
    const double Yc = dot3(R + 3, pw) + t[1];
    const double inv_Zc = 1.0 / (dot3(R + 6, pw) + t[2]);
    const double ue = K.cx + K.fx * Xc * inv_Zc;
    const double ve = K.cy + K.fy * Yc * inv_Zc;
    const double u = us[2 * i], v = us[2 * i + 1];
    sum2 += sqrt((u - ue) * (u - ue) + (v - ve) * (v - ve));
  }
  return sum2 / kN;
}
// L_6x10 from the four null-space vectors (rows 8..11 of the swept matrix), the three beta initialisations
// (find_betas_approx_1/2/3) each followed by Gauss-Newton, and the choice of the candidate pose with the smallest
// mean reprojection error.  pose(betas, R, t) -> error is the caller's compute_ccs .. reprojection_error chain.
template <class PoseFn>
CVM_HD inline void solve_betas_and_pose(Heap h, const double* rho, PoseFn&& pose, double* R, double* t) {
  // ---- L_6x10 and rho
  {
    double L[60];
    for (int i = 0; i < 6; ++i) {
      // pairs (a,b): (0,1),(0,2),(0,3),(1,2),(1,3),(2,3)
      const int a = i < 3 ? 0 : (i < 5 ? 1 : 2);
      const int b = i < 3 ? i + 1 : (i < 5 ? i - 1 : 3);
      double dv[4][3];
      for (int v = 0; v < 4; ++v) {
        const int row = H_UT + (3 - v) * 12;
        for (int c = 0; c < 3; ++c) dv[v][c] = h[row + 3 * a + c] - h[row + 3 * b + c];
      }
      L[i * 10 + 0] = dot3(dv[0], dv[0]);
      L[i * 10 + 1] = 2.0 * dot3(dv[0], dv[1]);
      L[i * 10 + 2] = dot3(dv[1], dv[1]);
      L[i * 10 + 3] = 2.0 * dot3(dv[0], dv[2]);
      L[i * 10 + 4] = 2.0 * dot3(dv[1], dv[2]);
      L[i * 10 + 5] = dot3(dv[2], dv[2]);
      L[i * 10 + 6] = 2.0 * dot3(dv[0], dv[3]);
      L[i * 10 + 7] = 2.0 * dot3(dv[1], dv[3]);
      L[i * 10 + 8] = 2.0 * dot3(dv[2], dv[3]);
      L[i * 10 + 9] = dot3(dv[3], dv[3]);
    }
    for (int i = 0; i < 60; ++i) h[H_L + i] = L[i];  // rows 0..4 of the swept matrix are dead by now
  }
  double Rs[3][9], ts[3][3], errs[3];
  // ---- approx 1: [B11 B12 B13 B14]
  {
    const int cols[4] = {0, 1, 3, 6};
    double b4[4], betas[4];
    solve_cols<4>(h, cols, rho, b4);
    if (b4[0] < 0) {
      betas[0] = sqrt(-b4[0]);
      betas[1] = -b4[1] / betas[0];
      betas[2] = -b4[2] / betas[0];
      betas[3] = -b4[3] / betas[0];
    } else {
      betas[0] = sqrt(b4[0]);
      betas[1] = b4[1] / betas[0];
      betas[2] = b4[2] / betas[0];
      betas[3] = b4[3] / betas[0];
    }
    gauss_newton(h, rho, betas);
    errs[0] = pose(betas, Rs[0], ts[0]);
  }
  // ---- approx 2: [B11 B12 B22]
  {
    const int cols[3] = {0, 1, 2};
    double b3[3], betas[4];
    solve_cols<3>(h, cols, rho, b3);
    if (b3[0] < 0) {
      betas[0] = sqrt(-b3[0]);
      betas[1] = (b3[2] < 0) ? sqrt(-b3[2]) : 0.0;
    } else {
      betas[0] = sqrt(b3[0]);
      betas[1] = (b3[2] > 0) ? sqrt(b3[2]) : 0.0;
    }
    if (b3[1] < 0) betas[0] = -betas[0];
    betas[2] = 0.0;
    betas[3] = 0.0;
    gauss_newton(h, rho, betas);
    errs[1] = pose(betas, Rs[1], ts[1]);
  }
  // ---- approx 3: [B11 B12 B22 B13 B23]
  {
    const int cols[5] = {0, 1, 2, 3, 4};
    double b5[5], betas[4];
    solve_cols<5>(h, cols, rho, b5);
    if (b5[0] < 0) {
      betas[0] = sqrt(-b5[0]);
      betas[1] = (b5[2] < 0) ? sqrt(-b5[2]) : 0.0;
    } else {
      betas[0] = sqrt(b5[0]);
      betas[1] = (b5[2] > 0) ? sqrt(b5[2]) : 0.0;
    }
    if (b5[1] < 0) betas[0] = -betas[0];
    betas[2] = b5[3] / betas[0];
    betas[3] = 0.0;
    gauss_newton(h, rho, betas);
    errs[2] = pose(betas, Rs[2], ts[2]);
  }
  int N = 0;
  if (errs[1] < errs[0]) N = 1;
  if (errs[2] < errs[N]) N = 2;
  for (int i = 0; i < 9; ++i) R[i] = Rs[N][i];
  for (int i = 0; i < 3; ++i) t[i] = ts[N][i];
}
}  // namespace detail

// o: 5 object points (f32 xyz), im: 5 image points (f32 pixel uv).  Outputs R (row-major), t.
CVM_HD CVM_NOINLINE inline void epnp5(Heap h, const float* o, const float* im, const Intrinsics& K, double* R, double* t) {
  using namespace detail;
  // undistortPoints (no distortion): double math, stored as f32; epnp::init_points rescales in double
  const double ifx = 1. / K.fx, ify = 1. / K.fy;
  double us[2 * kN];
  for (int i = 0; i < kN; ++i) {
    const float xn = (float)(((double)im[2 * i] - K.cx) * ifx);
    const float yn = (float)(((double)im[2 * i + 1] - K.cy) * ify);
    us[2 * i] = (double)xn * K.fx + K.cx;
    us[2 * i + 1] = (double)yn * K.fy + K.cy;
  }
  // ---- choose_control_points
  double c0[3] = {0, 0, 0};
  for (int i = 0; i < kN; ++i)
    for (int j = 0; j < 3; ++j) c0[j] += (double)o[3 * i + j];
  for (int j = 0; j < 3; ++j) c0[j] /= kN;
  {
    // PW0^T PW0 (mulTransposed: upper triangle, sequential over points) then symmetric completion
    double ptp[9];
    for (int a = 0; a < 3; ++a)
      for (int b = a; b < 3; ++b) {
        double s = 0;
        for (int k = 0; k < kN; ++k) s += ((double)o[3 * k + a] - c0[a]) * ((double)o[3 * k + b] - c0[b]);
        ptp[3 * a + b] = s;
      }
    ptp[3] = ptp[1]; ptp[6] = ptp[2]; ptp[7] = ptp[5];
    Heap At = h.at(0), W = h.at(9), Vt = h.at(12);
    for (int r = 0; r < 3; ++r)
      for (int c = 0; c < 3; ++c) At[c * 3 + r] = ptp[3 * r + c];
    jacobi_svd<3, 3, true>(At, W, Vt);
    for (int j = 0; j < 3; ++j) h[H_CWS + j] = c0[j];
    for (int i = 1; i < 4; ++i) {
      const double k = sqrt(W[i - 1] / kN);
      for (int j = 0; j < 3; ++j) h[H_CWS + 3 * i + j] = c0[j] + k * At[(i - 1) * 3 + j];
    }
  }
  // ---- compute_barycentric_coordinates: cvInvert(CC, CV_SVD)
  double alpha[4 * kN];
  {
    Heap At = h.at(0), W = h.at(9), Vt = h.at(12);
    // CC[i][j-1] = cws[j][i] - cws[0][i]; Jacobi runs on rows of CC^T: At[c][r] = CC[r][c]
    for (int i = 0; i < 3; ++i)
      for (int j = 1; j < 4; ++j) At[(j - 1) * 3 + i] = h[H_CWS + 3 * j + i] - h[H_CWS + i];
    jacobi_svd<3, 3, true>(At, W, Vt);
    double ci[9];
    for (int c = 0; c < 3; ++c) {
      const double e[3] = {c == 0 ? 1. : 0., c == 1 ? 1. : 0., c == 2 ? 1. : 0.};
      double x[3];
      svd_backsubst<3, 3>(At, W, Vt, e, x);
      ci[c] = x[0]; ci[3 + c] = x[1]; ci[6 + c] = x[2];
    }
    for (int i = 0; i < kN; ++i) {
      const double px = (double)o[3 * i], py = (double)o[3 * i + 1], pz = (double)o[3 * i + 2];
      double a[4];
      for (int j = 0; j < 3; ++j)
        a[1 + j] = ci[3 * j] * (px - h[H_CWS]) + ci[3 * j + 1] * (py - h[H_CWS + 1]) + ci[3 * j + 2] * (pz - h[H_CWS + 2]);
      a[0] = 1.0 - a[1] - a[2] - a[3];
      for (int j = 0; j < 4; ++j) alpha[4 * i + j] = a[j];
    }
  }
  // rho (squared control-point distances) first: the control points share their slots with W of the big SVD
  double rho[6];
  {
    double c[4][3];
    for (int i = 0; i < 4; ++i)
      for (int j = 0; j < 3; ++j) c[i][j] = h[H_CWS + 3 * i + j];
    const int pa[6] = {0, 0, 0, 1, 1, 2}, pb[6] = {1, 2, 3, 2, 3, 3};
    for (int i = 0; i < 6; ++i) {
      const double d0 = c[pa[i]][0] - c[pb[i]][0], d1 = c[pa[i]][1] - c[pb[i]][1], d2 = c[pa[i]][2] - c[pb[i]][2];
      rho[i] = d0 * d0 + d1 * d1 + d2 * d2;
    }
  }
  // ---- M^T M (mulTransposed over the 10 x 12 M of fill_M, rows in order u0,v0,u1,v1,...)
  {
    // Every entry is s = 0; for each point: s += m1[a] * m1[b]; s += m2[a] * m2[b] -- accumulated here point by point
    // (same additions in the same order per entry), so that the alphas are read once per point and stay in registers.
    Heap At = h.at(H_AT);
    for (int a = 0; a < 12; ++a)
      for (int b = a; b < 12; ++b) At[a * 12 + b] = 0;
    CVM_NOUNROLL
    for (int i = 0; i < kN; ++i) {
      const double u = us[2 * i], v = us[2 * i + 1];
      double m1[12], m2[12];   // row 2i: [a*fu, 0, a*(uc-u)] x 4, row 2i+1: [0, a*fv, a*(vc-v)] x 4
      CVM_UNROLL
      for (int k4 = 0; k4 < 4; ++k4) {
        const double aa = alpha[4 * i + k4];
        m1[3 * k4] = aa * K.fx; m1[3 * k4 + 1] = 0.0; m1[3 * k4 + 2] = aa * (K.cx - u);
        m2[3 * k4] = 0.0; m2[3 * k4 + 1] = aa * K.fy; m2[3 * k4 + 2] = aa * (K.cy - v);
      }
      CVM_UNROLL
      for (int a = 0; a < 12; ++a) {
        CVM_UNROLL
        for (int b = a; b < 12; ++b) {
          double s = At[a * 12 + b];
          s += m1[a] * m1[b];
          s += m2[a] * m2[b];
          At[a * 12 + b] = s;
        }
      }
    }
    for (int a = 0; a < 12; ++a)
      for (int b = a + 1; b < 12; ++b) At[b * 12 + a] = At[a * 12 + b];  // symmetric: row a of MtM^T == row a of MtM
    K3_T0(t_j);
#if defined(__CUDA_ARCH__) && SEPB_JACOBI_ILP
    jacobi12_sweeps_ilp(At, h.at(H_W));   // same arithmetic, three row-disjoint rotations in flight (cvmath.cuh)
#else
    jacobi_svd<12, 12, false>(At, h.at(H_W), h.at(H_W));
#endif
    K3_ACC(5, t_j);
  }
  solve_betas_and_pose(h, rho, [&](const double* betas, double* Rc, double* tc) {
    return compute_R_and_t(h, alpha, o, us, K, betas, Rc, tc);
  }, R, t);
}

// ---- n-point EPnP: cv::solvePnP(SOLVEPNP_EPNP) over a whole consensus set -- the refit cv::solvePnPRansac runs
// when `flags` is P3P / AP3P (the minimal solver cannot take more than four points, so OpenCV swaps in EPnP there).
// Same statements as epnp5 with the point loops run to n; the barycentric alphas of a point are recomputed (same
// operations, same values) wherever epnp keeps them in its n x 4 array.  pts(i, o3, uv): object point and PIXEL
// position of point i as doubles (the refit converts the f32 inputs to f64 first, so undistortPoints keeps doubles).
template <class Pts>
CVM_HD CVM_NOINLINE inline void epnp_n(Heap h, int n, Pts&& pts, const Intrinsics& K, double* R, double* t) {
  using namespace detail;
  const double ifx = 1. / K.fx, ify = 1. / K.fy;
  auto point = [&](int i, double* o3, double* uv) {
    double px[2];
    pts(i, o3, px);
    uv[0] = ((px[0] - K.cx) * ifx) * K.fx + K.cx;   // undistortPoints (CV_64F in, CV_64F out), then epnp's rescale
    uv[1] = ((px[1] - K.cy) * ify) * K.fy + K.cy;
  };
  // ---- choose_control_points
  double c0[3] = {0, 0, 0};
  CVM_NOUNROLL
  for (int i = 0; i < n; ++i) {
    double o3[3], uv[2];
    point(i, o3, uv);
    for (int j = 0; j < 3; ++j) c0[j] += o3[j];
  }
  for (int j = 0; j < 3; ++j) c0[j] /= n;
  {
    double ptp[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
    CVM_NOUNROLL
    for (int k = 0; k < n; ++k) {
      double o3[3], uv[2];
      point(k, o3, uv);
      for (int a = 0; a < 3; ++a)
        for (int b = a; b < 3; ++b) ptp[3 * a + b] += (o3[a] - c0[a]) * (o3[b] - c0[b]);
    }
    ptp[3] = ptp[1]; ptp[6] = ptp[2]; ptp[7] = ptp[5];
    Heap At = h.at(0), W = h.at(9), Vt = h.at(12);
    for (int r = 0; r < 3; ++r)
      for (int c = 0; c < 3; ++c) At[c * 3 + r] = ptp[3 * r + c];
    jacobi_svd<3, 3, true>(At, W, Vt);
    for (int j = 0; j < 3; ++j) h[H_CWS + j] = c0[j];
    for (int i = 1; i < 4; ++i) {
      const double k = sqrt(W[i - 1] / n);
      for (int j = 0; j < 3; ++j) h[H_CWS + 3 * i + j] = c0[j] + k * At[(i - 1) * 3 + j];
    }
  }
  // ---- compute_barycentric_coordinates: cvInvert(CC, CV_SVD); the inverse and cws[0] stay in registers
  double ci[9], cw0[3];
  {
    Heap At = h.at(0), W = h.at(9), Vt = h.at(12);
    for (int i = 0; i < 3; ++i)
      for (int j = 1; j < 4; ++j) At[(j - 1) * 3 + i] = h[H_CWS + 3 * j + i] - h[H_CWS + i];
    jacobi_svd<3, 3, true>(At, W, Vt);
    for (int c = 0; c < 3; ++c) {
      const double e[3] = {c == 0 ? 1. : 0., c == 1 ? 1. : 0., c == 2 ? 1. : 0.};
      double x[3];
      svd_backsubst<3, 3>(At, W, Vt, e, x);
      ci[c] = x[0]; ci[3 + c] = x[1]; ci[6 + c] = x[2];
    }
    for (int j = 0; j < 3; ++j) cw0[j] = h[H_CWS + j];
  }
  auto alphas = [&](const double* o3, double* a) {
    for (int j = 0; j < 3; ++j)
      a[1 + j] = ci[3 * j] * (o3[0] - cw0[0]) + ci[3 * j + 1] * (o3[1] - cw0[1]) + ci[3 * j + 2] * (o3[2] - cw0[2]);
    a[0] = 1.0 - a[1] - a[2] - a[3];
  };
  double rho[6];
  {
    double c[4][3];
    for (int i = 0; i < 4; ++i)
      for (int j = 0; j < 3; ++j) c[i][j] = h[H_CWS + 3 * i + j];
    const int pa[6] = {0, 0, 0, 1, 1, 2}, pb[6] = {1, 2, 3, 2, 3, 3};
    for (int i = 0; i < 6; ++i) {
      const double d0 = c[pa[i]][0] - c[pb[i]][0], d1 = c[pa[i]][1] - c[pb[i]][1], d2 = c[pa[i]][2] - c[pb[i]][2];
      rho[i] = d0 * d0 + d1 * d1 + d2 * d2;
    }
  }
  // ---- M^T M over the 2n x 12 M of fill_M, accumulated point by point (same additions per entry, same order)
  {
    Heap At = h.at(H_AT);
    for (int a = 0; a < 12; ++a)
      for (int b = a; b < 12; ++b) At[a * 12 + b] = 0;
    CVM_NOUNROLL
    for (int i = 0; i < n; ++i) {
      double o3[3], uv[2], al[4];
      point(i, o3, uv);
      alphas(o3, al);
      double m1[12], m2[12];
      CVM_UNROLL
      for (int k4 = 0; k4 < 4; ++k4) {
        const double aa = al[k4];
        m1[3 * k4] = aa * K.fx; m1[3 * k4 + 1] = 0.0; m1[3 * k4 + 2] = aa * (K.cx - uv[0]);
        m2[3 * k4] = 0.0; m2[3 * k4 + 1] = aa * K.fy; m2[3 * k4 + 2] = aa * (K.cy - uv[1]);
      }
      CVM_UNROLL
      for (int a = 0; a < 12; ++a) {
        CVM_UNROLL
        for (int b = a; b < 12; ++b) {
          double s2 = At[a * 12 + b];
          s2 += m1[a] * m1[b];
          s2 += m2[a] * m2[b];
          At[a * 12 + b] = s2;
        }
      }
    }
    for (int a = 0; a < 12; ++a)
      for (int b = a + 1; b < 12; ++b) At[b * 12 + a] = At[a * 12 + b];
    jacobi_svd<12, 12, false>(At, h.at(H_W), h.at(H_W));
  }
  // compute_ccs / compute_pcs / solve_for_sign / estimate_R_and_t / reprojection_error over the n points
  auto pose = [&](const double* betas, double* Rc, double* tc) -> double {
    double ccs[4][3] = {{0, 0, 0}, {0, 0, 0}, {0, 0, 0}, {0, 0, 0}};
    for (int i = 0; i < 4; ++i) {
      const int row = H_UT + (3 - i) * 12;
      for (int j = 0; j < 4; ++j)
        for (int k = 0; k < 3; ++k) ccs[j][k] += betas[i] * h[row + 3 * j + k];
    }
    auto pc_of = [&](const double* o3, double* pc) {
      double al[4];
      alphas(o3, al);
      for (int j = 0; j < 3; ++j) pc[j] = al[0] * ccs[0][j] + al[1] * ccs[1][j] + al[2] * ccs[2][j] + al[3] * ccs[3][j];
    };
    bool flip;
    {
      double o3[3], uv[2], pc[3];
      point(0, o3, uv);
      pc_of(o3, pc);
      flip = pc[2] < 0.0;
    }
    double pc0[3] = {0, 0, 0}, pw0[3] = {0, 0, 0};
    CVM_NOUNROLL
    for (int i = 0; i < n; ++i) {
      double o3[3], uv[2], pc[3];
      point(i, o3, uv);
      pc_of(o3, pc);
      for (int j = 0; j < 3; ++j) {
        pc0[j] += flip ? -pc[j] : pc[j];
        pw0[j] += o3[j];
      }
    }
    for (int j = 0; j < 3; ++j) {
      pc0[j] /= n;
      pw0[j] /= n;
    }
    double abt[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
    CVM_NOUNROLL
    for (int i = 0; i < n; ++i) {
      double o3[3], uv[2], pc[3];
      point(i, o3, uv);
      pc_of(o3, pc);
      for (int j = 0; j < 3; ++j) {
        const double pj = flip ? -pc[j] : pc[j];
        abt[3 * j] += (pj - pc0[j]) * (o3[0] - pw0[0]);
        abt[3 * j + 1] += (pj - pc0[j]) * (o3[1] - pw0[1]);
        abt[3 * j + 2] += (pj - pc0[j]) * (o3[2] - pw0[2]);
      }
    }
    Heap At = h.at(H_SAT), W = h.at(H_SW), Vt = h.at(H_SVT);
    for (int r = 0; r < 3; ++r)
      for (int c = 0; c < 3; ++c) At[c * 3 + r] = abt[3 * r + c];
    jacobi_svd<3, 3, true>(At, W, Vt);
    for (int i = 0; i < 3; ++i)
      for (int j = 0; j < 3; ++j)
        Rc[3 * i + j] = At[0 * 3 + i] * Vt[0 * 3 + j] + At[1 * 3 + i] * Vt[1 * 3 + j] + At[2 * 3 + i] * Vt[2 * 3 + j];
    const double det = Rc[0] * Rc[4] * Rc[8] + Rc[1] * Rc[5] * Rc[6] + Rc[2] * Rc[3] * Rc[7] - Rc[2] * Rc[4] * Rc[6] -
                       Rc[1] * Rc[3] * Rc[8] - Rc[0] * Rc[5] * Rc[7];
    if (det < 0) {
      Rc[6] = -Rc[6];
      Rc[7] = -Rc[7];
      Rc[8] = -Rc[8];
    }
    tc[0] = pc0[0] - dot3(Rc, pw0);
    tc[1] = pc0[1] - dot3(Rc + 3, pw0);
    tc[2] = pc0[2] - dot3(Rc + 6, pw0);
    double sum2 = 0;
    CVM_NOUNROLL
    for (int i = 0; i < n; ++i) {
      double pw[3], uv[2];
      point(i, pw, uv);
      const double Xc = dot3(Rc, pw) + tc[0];
      const double Yc = dot3(Rc + 3, pw) + tc[1];
      const double inv_Zc = 1.0 / (dot3(Rc + 6, pw) + tc[2]);
      const double ue = K.cx + K.fx * Xc * inv_Zc;
      const double ve = K.cy + K.fy * Yc * inv_Zc;
      sum2 += sqrt((uv[0] - ue) * (uv[0] - ue) + (uv[1] - ve) * (uv[1] - ve));
    }
    return sum2 / n;
  };
  solve_betas_and_pose(h, rho, pose, R, t);
}

}  // namespace cvm
