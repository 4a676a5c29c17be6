// TEST-ONLY host build of the product's __host__ __device__ math headers (csrc/cvmath.cuh,
// epnp.cuh, pnp_geom.cuh) so that the operation order can be checked bit-for-bit against cv2 on a
// machine without a GPU.  Never linked into libsepb200.so; the product path is the CUDA kernel.
#include "../../multi_robot_slam_separators_b200/csrc/epnp.cuh"
#include "../../multi_robot_slam_separators_b200/csrc/pnp_geom.cuh"

extern "C" {

void hh_svd(double* At, int m, int n, double* W, double* Vt) {
  cvm::jacobi_svd(cvm::Heap{At, 1}, cvm::Heap{W, 1}, cvm::Heap{Vt, 1}, true, m, n);
}

void hh_epnp5(const float* obj, const float* img, const double* K4, double* R, double* t, double* rvec, double* R2) {
  double heap[cvm::kEpnpHeap];
  cvm::Intrinsics K{K4[0], K4[1], K4[2], K4[3]};
  cvm::epnp5(cvm::Heap{heap, 1}, obj, img, K, R, t);
  cvm::rodrigues_mat_to_vec(cvm::Heap{heap, 1}, R, rvec);
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

int hh_update_iters(double p, double ep, int mp, int mx) { return cvm::ransac_update_num_iters(p, ep, mp, mx); }
}
