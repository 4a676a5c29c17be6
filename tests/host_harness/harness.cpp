// TEST-ONLY host build of the product's __host__ __device__ math headers (csrc/cvmath.cuh,
// epnp.cuh, pnp_geom.cuh) so that the operation order can be checked bit-for-bit against cv2 on a
// machine without a GPU.  Never linked into libsepb200.so; the product path is the CUDA kernel.
#include "../../multi_robot_slam_separators_b200/csrc/epnp.cuh"
#include "../../multi_robot_slam_separators_b200/csrc/pnp_geom.cuh"
#include "../../multi_robot_slam_separators_b200/csrc/p3p.cuh"
#include "../../multi_robot_slam_separators_b200/csrc/ba_math.cuh"

extern "C" {

// the product instantiates the Jacobi SVD per shape; returns -1 for a shape the path never uses
int hh_svd(double* At, int m, int n, double* W, double* Vt) {
  const cvm::Heap a{At}, w{W}, v{Vt};
  if (m == 12 && n == 12) cvm::jacobi_svd<12, 12, true>(a, w, v);
  else if (m == 3 && n == 3) cvm::jacobi_svd<3, 3, true>(a, w, v);
  else if (m == 6 && n == 3) cvm::jacobi_svd<6, 3, true>(a, w, v);
  else if (m == 6 && n == 4) cvm::jacobi_svd<6, 4, true>(a, w, v);
  else if (m == 6 && n == 5) cvm::jacobi_svd<6, 5, true>(a, w, v);
  else if (m == 6 && n == 6) cvm::jacobi_svd<6, 6, true>(a, w, v);
  else return -1;
  return 0;
}

void hh_epnp5(const float* obj, const float* img, const double* K4, double* R, double* t, double* rvec, double* R2) {
  double heap[cvm::kEpnpHeap];
  cvm::Intrinsics K{K4[0], K4[1], K4[2], K4[3]};
  cvm::epnp5(cvm::Heap{heap}, obj, img, K, R, t);
  cvm::rodrigues_mat_to_vec(cvm::Heap{heap}, R, rvec);
  cvm::rodrigues_vec_to_mat(rvec, R2, nullptr);
}

void hh_reproj_err(const double* rvec, const double* t, const double* K4, const float* obj, const float* img, int n,
                   float* err) {
  double R[9];
  cvm::rodrigues_vec_to_mat(rvec, R, nullptr);
  cvm::Intrinsics K{K4[0], K4[1], K4[2], K4[3]};
  for (int i = 0; i < n; ++i) err[i] = cvm::reproj_err_f32(R, t, K, obj + 3 * i, img + 2 * i);
}

void hh_rodrigues_jac(const double* rvec, double* R, double* J) { cvm::rodrigues_vec_to_mat(rvec, R, J); }

// cv::solvePnP(SOLVEPNP_P3P) on four correspondences; returns 0 when there is no solution
int hh_p3p4(const float* obj, const float* img, const double* K4, double* R, double* t, double* rvec) {
  double heap[cvm::kEpnpHeap];
  cvm::Intrinsics K{K4[0], K4[1], K4[2], K4[3]};
  if (!cvm::p3p4(obj, img, K, R, t)) return 0;
  cvm::rodrigues_mat_to_vec(cvm::Heap{heap}, R, rvec);
  return 1;
}

// cv::solvePnP(SOLVEPNP_EPNP) on n f64 correspondences (the refit of the P3P-flag path)
void hh_epnp_n(const double* obj, const double* img, int n, const double* K4, double* R, double* t, double* rvec) {
  double heap[cvm::kEpnpHeap];
  cvm::Intrinsics K{K4[0], K4[1], K4[2], K4[3]};
  cvm::epnp_n(cvm::Heap{heap}, n, [&](int i, double* o3, double* px) {
    for (int k = 0; k < 3; ++k) o3[k] = obj[3 * i + k];
    px[0] = img[2 * i];
    px[1] = img[2 * i + 1];
  }, K, R, t);
  cvm::rodrigues_mat_to_vec(cvm::Heap{heap}, R, rvec);
}

int hh_update_iters(double p, double ep, int mp, int mx) { return cvm::ransac_update_num_iters(p, ep, mp, mx); }

// ---- 2-view bundle adjustment arithmetic (ba_math.cuh) against oracle/ba.py ------------------------------------------
static ba::BaKernelParams ba_params(const double* K4, double baseline, double var, double delta) {
  ba::BaKernelParams P{};
  P.fx = K4[0]; P.fy = K4[1]; P.cx = K4[2]; P.cy = K4[3];
  P.baseline = baseline; P.var = var; P.delta = delta;
  return P;
}
static ba::Cam ba_cam(const double* q4, const double* t3) {
  ba::Cam c;
  for (int i = 0; i < 4; ++i) c.p.q[i] = q4[i];
  for (int i = 0; i < 3; ++i) c.p.t[i] = t3[i];
  ba::cam_finish(c);
  return c;
}
// one point's block of the normal equations: V (6, symmetric), bl (3), W (18), camera block A (21) and bc (6), chi2
double hh_ba_point_system(const double* K4, double baseline, double var, double delta, const double* q1, const double* t1,
                          const double* q2, const double* t2, const double* X, const float* meas6, unsigned flags, double* V,
                          double* bl, double* W, double* A, double* bc) {
  const ba::BaKernelParams P = ba_params(K4, baseline, var, delta);
  const ba::Cam c1 = ba_cam(q1, t1), c2 = ba_cam(q2, t2);
  ba::PtSys s;
  for (int i = 0; i < 21; ++i) A[i] = 0;
  for (int i = 0; i < 6; ++i) bc[i] = 0;
  const double chi = ba::point_system(P, c1, c2, X, meas6, flags, s, A, bc);
  for (int i = 0; i < 6; ++i) V[i] = s.V[i];
  for (int i = 0; i < 3; ++i) bl[i] = s.bl[i];
  for (int i = 0; i < 18; ++i) W[i] = s.W[i];
  return chi;
}
void hh_ba_se3(const double* qa, const double* ta, const double* qb, const double* tb, const double* upd6, double* mul7,
               double* inv7, double* upd7, double* link6) {
  ba::Se3 a, b;
  for (int i = 0; i < 4; ++i) { a.q[i] = qa[i]; b.q[i] = qb[i]; }
  for (int i = 0; i < 3; ++i) { a.t[i] = ta[i]; b.t[i] = tb[i]; }
  const ba::Se3 m = ba::se3_mul(a, b), iv = ba::se3_inv(a), u = ba::se3_update(a, upd6);
  for (int i = 0; i < 4; ++i) { mul7[i] = m.q[i]; inv7[i] = iv.q[i]; upd7[i] = u.q[i]; }
  for (int i = 0; i < 3; ++i) { mul7[4 + i] = m.t[i]; inv7[4 + i] = iv.t[i]; upd7[4 + i] = u.t[i]; }
  ba::link_error(iv, iv, b, link6);   // measurement^-1 = cam1^-1 = a^-1: any fixed pair exercises the composition
}
int hh_ba_cholesky6(const double* S21, double lam, const double* rhs, double* x) { return ba::cholesky6(S21, lam, rhs, x) ? 1 : 0; }
void hh_ba_quat(const double* R9, double* q4, double* R9_back) {
  ba::quat_from_matrix(R9, q4);
  ba::quat_normalize_rotation(q4);
  ba::quat_to_matrix(q4, R9_back);
}
}
