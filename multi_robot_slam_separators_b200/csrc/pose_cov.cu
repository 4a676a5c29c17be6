// Covariance of a verified pair, as util3d::estimateMotion3DTo2D (RTAB-Map, reached from
// R/src/myRegistrationVis.cpp:1077-1090 with words3B = the To frame's 3-D points) fills it and
// Registration::computeTransformationModFromFeats clamps it (R/src/myRegistration.cpp:279-295):
//   cov = I6; over the inliers whose To 3-D point is finite: squared distance between the From point and
//   T * (To point), and the angle between the two rays from T's origin (pcl::getAngle3D), all in f32;
//   both lists sorted, variance = 2.1981 * element[size / 4] (lower quartile: "ignore very far features");
//   cov[0:3,0:3] *= variance_lin, cov[3:6,3:6] *= variance_ang; diagonal entries <= 1e-9 become 1e-9.
// No usable point (or no transform): cov stays I6.  One CTA per pair; the order statistic is found by
// rank counting in shared memory (n <= max_feat values).
#include "common.cuh"

namespace {

constexpr int kThreadsC = 128;

__global__ void __launch_bounds__(kThreadsC)
pose_cov_kernel(FramesView from, FramesView to, const int32_t* __restrict__ pairs, int n_pairs,
                const int32_t* __restrict__ status, const int32_t* __restrict__ n_matches,
                const uint16_t* __restrict__ match_from, const uint16_t* __restrict__ match_to,
                const uint32_t* __restrict__ inlier_mask, int mask_words, const float* __restrict__ transform,
                double* __restrict__ cov /*[P][36]*/) {
  extern __shared__ float s_val[];   // [2][max_feat]: squared distances, angles
  __shared__ int s_n;
  __shared__ float s_q[2];
  const int pair = blockIdx.x, tid = threadIdx.x;
  if (pair >= n_pairs) return;
  double* C = cov + 36 * (size_t)pair;
  for (int i = tid; i < 36; i += kThreadsC) C[i] = (i % 7 == 0) ? 1.0 : 0.0;
  if (status[pair] != 1) return;
  const int fi = pairs[2 * pair], ti = pairs[2 * pair + 1];
  const int M = n_matches[pair];
  const float* T = transform + 12 * (size_t)pair;
  const float* p3f = from.kp3d + (size_t)fi * from.max_feat * 3;
  const float* p3t = to.kp3d + (size_t)ti * to.max_feat * 3;
  const uint16_t* mf = match_from + (size_t)pair * from.max_feat;
  const uint16_t* mt = match_to + (size_t)pair * from.max_feat;
  const uint32_t* mask = inlier_mask + (size_t)pair * mask_words;
  float* s_d = s_val;
  float* s_a = s_val + from.max_feat;
  if (tid == 0) s_n = 0;
  __syncthreads();
  for (int m = tid; m < M; m += kThreadsC) {
    if (!((mask[m >> 5] >> (m & 31)) & 1u)) continue;
    const float* b = p3t + 3 * mt[m];
    if (!(isfinite(b[0]) && isfinite(b[1]) && isfinite(b[2]))) continue;
    const float* a = p3f + 3 * mf[m];
    // util3d::transformPoint (f32, no contraction)
    const float nx = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(T[0], b[0]), __fmul_rn(T[1], b[1])), __fmul_rn(T[2], b[2])), T[3]);
    const float ny = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(T[4], b[0]), __fmul_rn(T[5], b[1])), __fmul_rn(T[6], b[2])), T[7]);
    const float nz = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(T[8], b[0]), __fmul_rn(T[9], b[1])), __fmul_rn(T[10], b[2])), T[11]);
    const float dx = a[0] - nx, dy = a[1] - ny, dz = a[2] - nz;
    const float d2 = __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));
    // pcl::getAngle3D(v1, v2): acos of the clamped dot product of the normalised rays from T's origin
    const float v1x = a[0] - T[3], v1y = a[1] - T[7], v1z = a[2] - T[11];
    const float v2x = nx - T[3], v2y = ny - T[7], v2z = nz - T[11];
    const float n1 = sqrtf(__fadd_rn(__fadd_rn(__fmul_rn(v1x, v1x), __fmul_rn(v1y, v1y)), __fmul_rn(v1z, v1z)));
    const float n2 = sqrtf(__fadd_rn(__fadd_rn(__fmul_rn(v2x, v2x), __fmul_rn(v2y, v2y)), __fmul_rn(v2z, v2z)));
    float rad = __fadd_rn(__fadd_rn(__fmul_rn(__fdiv_rn(v1x, n1), __fdiv_rn(v2x, n2)), __fmul_rn(__fdiv_rn(v1y, n1), __fdiv_rn(v2y, n2))),
                          __fmul_rn(__fdiv_rn(v1z, n1), __fdiv_rn(v2z, n2)));
    rad = rad < -1.f ? -1.f : (rad > 1.f ? 1.f : rad);
    const int slot = atomicAdd(&s_n, 1);
    s_d[slot] = d2;
    s_a[slot] = acosf(rad);
  }
  __syncthreads();
  const int n = s_n;
  if (n == 0) return;   // "Not enough close points to compute covariance": identity stands
  // element [n >> 2] of each sorted list = the value whose rank (ties broken by slot) is n >> 2
  const int want = n >> 2;
  for (int which = 0; which < 2; ++which) {
    const float* v = which ? s_a : s_d;
    for (int i = tid; i < n; i += kThreadsC) {
      const float x = v[i];
      int rank = 0;
      for (int j = 0; j < n; ++j) {
        const float y = v[j];
        rank += (y < x || (y == x && j < i)) ? 1 : 0;
      }
      if (rank == want) s_q[which] = x;   // exactly one writer (NaN values would break the order; inputs are finite)
    }
  }
  __syncthreads();
  if (tid == 0) {
    const double vl = 2.1981 * (double)s_q[0], va = 2.1981 * (double)s_q[1];
    const double eps = 1e-9;   // COVARIANCE_EPSILON (myRegistration.cpp:36)
    for (int i = 0; i < 3; ++i) {
      C[7 * i] = vl <= eps ? eps : vl;
      C[7 * (i + 3)] = va <= eps ? eps : va;
    }
  }
}

}  // namespace

int sepb_launch_pose_cov(sepb_ctx* ctx, FramesView from, FramesView to, const int32_t* d_pairs, int n_pairs,
                         const int32_t* d_status, const int32_t* d_n_matches, const uint16_t* d_match_from,
                         const uint16_t* d_match_to, const uint32_t* d_inlier_mask, int mask_words,
                         const float* d_transform, double* d_cov) {
  if (n_pairs == 0) return SEPB_OK;
  const size_t smem = (size_t)from.max_feat * 8;
  pose_cov_kernel<<<n_pairs, kThreadsC, smem, ctx->stream>>>(from, to, d_pairs, n_pairs, d_status, d_n_matches, d_match_from,
                                                           d_match_to, d_inlier_mask, mask_words, d_transform, d_cov);
  ctx->launches++;
  SEPB_CUDA(ctx, cudaGetLastError());
  return SEPB_OK;
}
