// TEST-ONLY host build of the product's __host__ __device__ math headers (csrc/cvmath.cuh,
// epnp.cuh, pnp_geom.cuh) so that the operation order can be checked bit-for-bit against cv2 on a
// machine without a GPU.  Never linked into libsepb200.so; the product path is the CUDA kernel.
#include "../../multi_robot_slam_separators_b200/csrc/epnp.cuh"
#include "../../multi_robot_slam_separators_b200/csrc/pnp_geom.cuh"
#include "../../multi_robot_slam_separators_b200/csrc/p3p.cuh"

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
}
