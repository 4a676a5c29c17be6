// 2-view bundle adjustment of a verified pair: the block R/src/myRegistrationVis.cpp:1192-1345 (Vis/BundleAdjustment = 1),
// i.e. RTAB-Map's OptimizerG2O::optimizeBA(rootId = 1, {1: identity, 2: transforms[0]}, one link, the inliers' 3-D
// points and their pixels in both frames) on top of g2o -- neither of them vendored by the reference, so this kernel
// follows the published algorithm as oracle/ba.py restates it (parity unpinned to the reference's binaries):
//   * camera 1 (From, pose = localTransform) fixed, camera 2 (To, pose = T * localTransform) free (translation +
//     quaternion-vector increment, SBACam::update), one free 3-D point per inlier starting at the From frame's point;
//   * per inlier and camera a stereo edge (u, v, u - baseline * fx / depth) when the key point has a finite positive
//     depth in that camera, else a mono edge (u, v); information 1 / pixel_variance; Huber kernel;
//   * one EdgeSBACam between the cameras (measurement local^-1 T local, information = covariance^-1, numeric Jacobian);
//   * Levenberg-Marquardt as g2o runs it (lambda_0 = 1e-5 max|diag H|, gain ratio over dx.(lambda dx + b) + 1e-3,
//     lambda *= max(1/3, min(1 - (2 rho - 1)^3, 2/3)) or *= ni, ni *= 2, ten trials), Schur complement over the points;
//   * 5 iterations, edges with chi2 > delta leave (level 1) and flag their inlier, `iterations` more, flag again;
//   * result: T = camera 2 * localTransform^-1, the flagged inliers dropped, null transform when fewer than
//     min_inliers stay (:1340-1345) or the optimisation diverged.
// One CTA per verified pair; point estimates, measurements and edge levels live in shared memory, every pass over the
// points is thread-strided with a fixed-order block reduction (deterministic), the 6 x 6 camera system is solved by
// thread 0.  FP64 throughout; the depth / disparity arithmetic is f32 like rtabmap::Transform and the reference's casts.
#include "common.cuh"
#include "ba_math.cuh"

namespace {

using namespace ba;

#ifndef SEPB_BA_MIN_BLOCKS
#define SEPB_BA_MIN_BLOCKS 1   // resident CTAs per SM the register allocation aims at (section 4.2 of DESIGN.md)
#endif
#ifndef SEPB_BA_THREADS
#define SEPB_BA_THREADS 128
#endif
constexpr int kThreadsB = SEPB_BA_THREADS;
constexpr int kWarpsB = kThreadsB / 32;
constexpr double kHuge = 1.7976931348623157e308;

// ---- block reductions (fixed order: lanes by shuffle tree, warps 0..3 in sequence) --------------------------------
template <int K>
__device__ void block_sum(double* v, double* s_red /*[kWarpsB][K]*/) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int k = 0; k < K; ++k) {
    double x = v[k];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) x += __shfl_down_sync(0xffffffffu, x, o);
    if (lane == 0) s_red[warp * K + k] = x;
  }
  __syncthreads();
#pragma unroll
  for (int k = 0; k < K; ++k) {
    double x = s_red[k];
#pragma unroll
    for (int w = 1; w < kWarpsB; ++w) x += s_red[w * K + k];
    v[k] = x;
  }
  __syncthreads();
}

__device__ double block_max(double x, double* s_red) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) x = fmax(x, __shfl_down_sync(0xffffffffu, x, o));
  if (lane == 0) s_red[warp] = x;
  __syncthreads();
  x = s_red[0];
#pragma unroll
  for (int w = 1; w < kWarpsB; ++w) x = fmax(x, s_red[w]);
  __syncthreads();
  return x;
}

// ---- the kernel -----------------------------------------------------------------------------------------------------
struct Ctl {
  Cam c1, c2, cn;              // fixed camera, camera 2, camera 2 after the trial step
  Se3 link_inv, cam1_inv;
  double info[6];              // diagonal of covariance^-1 (EdgeSBACam information)
  double Al[21], bcl[6], chil; // the link edge's share of the camera block at c2
  double bc[6];                // raw camera gradient of the trial (points + link)
  double xp[6];
  double lam, ni, cur_chi, rho, chi2, lchi_n;   // lchi_n: the link edge's chi2 at the trial pose cn
  int qmax, ok, again, stop, cur, fail, iters;
};

// The link edge's share of the camera block at c2, by the whole CTA: numeric Jacobian as BaseBinaryEdge::linearizeOplus
// does it (central differences, step 1e-9) -- threads 0..12 evaluate the error at c2 and at its twelve perturbations,
// threads 0..27 each assemble one entry of J^T Omega J (21), -J^T Omega e (6) and e^T Omega e.  Ends with a barrier.
__device__ void link_terms(Ctl& c, double* s_le /*[13][6]*/) {
  const int tid = threadIdx.x;
  if (tid < 13) {
    double e[6];
    if (tid == 0) {
      link_error(c.link_inv, c.cam1_inv, c.c2.p, e);
    } else {
      double u[6] = {0, 0, 0, 0, 0, 0};
      u[(tid - 1) >> 1] = (tid & 1) ? 1e-9 : -1e-9;     // 1, 2: +/- step of dimension 0; 3, 4: dimension 1; ...
      link_error(c.link_inv, c.cam1_inv, se3_update(c.c2.p, u), e);
    }
    for (int r = 0; r < 6; ++r) s_le[6 * tid + r] = e[r];
  }
  __syncthreads();
  if (tid < 28) {
    // J[r][d] = (e_plus[d][r] - e_minus[d][r]) / 2e-9
    auto J = [&](int r, int d) { return (s_le[6 * (1 + 2 * d) + r] - s_le[6 * (2 + 2 * d) + r]) * (1.0 / 2e-9); };
    if (tid < 21) {
      int i = 0, k = tid;
      while (k >= 6 - i) { k -= 6 - i; ++i; }
      const int j = i + k;
      double sum = 0;
      for (int r = 0; r < 6; ++r) sum += J(r, i) * c.info[r] * J(r, j);
      c.Al[tid] = sum;
    } else if (tid < 27) {
      const int i = tid - 21;
      double sum = 0;
      for (int r = 0; r < 6; ++r) sum += J(r, i) * c.info[r] * s_le[r];
      c.bcl[i] = -sum;
    } else {
      double sum = 0;
      for (int r = 0; r < 6; ++r) sum += s_le[r] * c.info[r] * s_le[r];
      c.chil = sum;
    }
  }
  __syncthreads();
}

__device__ double link_chi(const Ctl& c, const Se3& cam2) {
  double e[6], s = 0;
  link_error(c.link_inv, c.cam1_inv, cam2, e);
  for (int r = 0; r < 6; ++r) s += e[r] * c.info[r] * e[r];
  return s;
}

// util3d::transformPoint(p, localTransform^-1).z in f32
__device__ __forceinline__ float depth_in_camera(const BaKernelParams& P, const float* p) {
  const float r0 = (float)P.LR[2], r1 = (float)P.LR[5], r2 = (float)P.LR[8];   // third row of R^T = third column of R
  const float tz = -__fadd_rn(__fadd_rn(__fmul_rn(r0, (float)P.Lt[0]), __fmul_rn(r1, (float)P.Lt[1])), __fmul_rn(r2, (float)P.Lt[2]));
  return __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(r0, p[0]), __fmul_rn(r1, p[1])), __fmul_rn(r2, p[2])), tz);
}

__global__ void __launch_bounds__(kThreadsB, SEPB_BA_MIN_BLOCKS)
bundle_adjust_kernel(FramesView from, FramesView to, const int32_t* __restrict__ pairs, int n_pairs,
                     const int32_t* __restrict__ n_matches, const uint16_t* __restrict__ match_from,
                     const uint16_t* __restrict__ match_to, const double* __restrict__ cov, BaKernelParams P,
                     int32_t* __restrict__ status, int32_t* __restrict__ n_inliers, double* __restrict__ pose,
                     float* __restrict__ transform, uint32_t* __restrict__ inlier_mask, int mask_words,
                     double* __restrict__ stats) {
  extern __shared__ __align__(16) unsigned char s_raw[];
  __shared__ Ctl ctl;
  __shared__ double s_red[kWarpsB * 34];
  __shared__ double s_le[13 * 6];
  __shared__ int s_off[64 + 1];
  const int pair = blockIdx.x, tid = threadIdx.x;
  if (pair >= n_pairs || status[pair] != 1) return;
  const int F = from.max_feat;
  double* s_X = reinterpret_cast<double*>(s_raw);                       // [2][F][3]
  float* s_meas = reinterpret_cast<float*>(s_X + 6 * (size_t)F);        // [F][6]
  uint16_t* s_idx = reinterpret_cast<uint16_t*>(s_meas + 6 * (size_t)F);   // [F] match-list position of the inlier
  uint8_t* s_flag = reinterpret_cast<uint8_t*>(s_idx + F);              // [F] bit0,1 edge levels; bit2,3 stereo
  const int fi = pairs[2 * pair], ti = pairs[2 * pair + 1];
  const int M = n_matches[pair];
  const uint16_t* mf = match_from + (size_t)pair * F;
  const uint16_t* mt = match_to + (size_t)pair * F;
  uint32_t* mask = inlier_mask + (size_t)pair * mask_words;
  const int words = (M + 31) >> 5;

  // inliers in match-list order
  if (tid == 0) {
    int n = 0;
    for (int w = 0; w < words; ++w) {
      s_off[w] = n;
      uint32_t b = mask[w];
      if (w == words - 1 && (M & 31)) b &= (1u << (M & 31)) - 1u;
      n += __popc(b);
    }
    s_off[words] = n;
  }
  __syncthreads();
  const int N = s_off[words];
  if (N == 0) return;
  for (int w = tid; w < words; w += kThreadsB) {
    uint32_t b = mask[w];
    if (w == words - 1 && (M & 31)) b &= (1u << (M & 31)) - 1u;
    int o = s_off[w];
    while (b) {
      const int bit = __ffs(b) - 1;
      b &= b - 1;
      s_idx[o++] = (uint16_t)(32 * w + bit);
    }
  }
  __syncthreads();
  const float* k2f = from.kp2d + (size_t)fi * F * 2;
  const float* k3f = from.kp3d + (size_t)fi * F * 3;
  const float* k2t = to.kp2d + (size_t)ti * to.max_feat * 2;
  const float* k3t = to.kp3d + (size_t)ti * to.max_feat * 3;
  for (int i = tid; i < N; i += kThreadsB) {
    const int m = s_idx[i];
    const int a = mf[m], b = mt[m];
    unsigned fl = 0;
    for (int c = 0; c < 2; ++c) {
      const float* p2 = c ? k2t + 2 * b : k2f + 2 * a;
      const float* p3 = c ? k3t + 3 * b : k3f + 3 * a;
      const float depth = depth_in_camera(P, p3);
      const bool st = isfinite(depth) && depth > 0.f && P.baseline > 0;
      float* ms = s_meas + 6 * i + 3 * c;
      ms[0] = p2[0];
      ms[1] = p2[1];
      ms[2] = 0.f;
      if (st) {
        const float disparity = (float)(P.baseline * P.fx / (double)depth);
        ms[2] = __fsub_rn(p2[0], disparity);
        fl |= 4u << c;
      }
    }
    s_flag[i] = (uint8_t)fl;
    for (int k = 0; k < 3; ++k) s_X[3 * i + k] = (double)k3f[3 * a + k];
  }
  if (tid == 0) {
    const float* T = transform + 12 * (size_t)pair;
    double R2[9], Rm[9];
    for (int i = 0; i < 3; ++i) {
      ctl.c1.p.t[i] = P.Lt[i];
      ctl.c2.p.t[i] = (double)T[4 * i] * P.Lt[0] + (double)T[4 * i + 1] * P.Lt[1] + (double)T[4 * i + 2] * P.Lt[2] + (double)T[4 * i + 3];
      for (int j = 0; j < 3; ++j)
        R2[3 * i + j] = (double)T[4 * i] * P.LR[j] + (double)T[4 * i + 1] * P.LR[3 + j] + (double)T[4 * i + 2] * P.LR[6 + j];
    }
    quat_from_matrix(P.LR, ctl.c1.p.q);
    quat_normalize_rotation(ctl.c1.p.q);
    quat_from_matrix(R2, ctl.c2.p.q);
    quat_normalize_rotation(ctl.c2.p.q);
    cam_finish(ctl.c1);
    cam_finish(ctl.c2);
    // EdgeSBACam measurement local^-1 * T * local
    Se3 link;
    for (int i = 0; i < 3; ++i) {
      for (int j = 0; j < 3; ++j) Rm[3 * i + j] = P.LR[i] * R2[j] + P.LR[3 + i] * R2[3 + j] + P.LR[6 + i] * R2[6 + j];
      link.t[i] = P.LR[i] * (ctl.c2.p.t[0] - P.Lt[0]) + P.LR[3 + i] * (ctl.c2.p.t[1] - P.Lt[1]) + P.LR[6 + i] * (ctl.c2.p.t[2] - P.Lt[2]);
    }
    quat_from_matrix(Rm, link.q);
    quat_normalize_rotation(link.q);
    ctl.link_inv = se3_inv(link);
    ctl.cam1_inv = se3_inv(ctl.c1.p);
    for (int i = 0; i < 6; ++i) ctl.info[i] = (P.use_cov && cov) ? 1.0 / cov[36 * (size_t)pair + 7 * i] : 1.0;
    ctl.cur = 0;
    ctl.fail = 0;
    ctl.iters = 0;
    ctl.chi2 = 0;
  }
  __syncthreads();

  const bool robust = P.delta > 0;
  const int rounds = robust ? 2 : 1;
#pragma unroll 1
  for (int round = 0; round < rounds; ++round) {
    const int n_iter = (round == 0 && robust) ? 5 : P.iterations;
    // ---- SparseOptimizer::optimize(n_iter) with OptimizationAlgorithmLevenberg
#pragma unroll 1
    for (int it = 0; it < n_iter; ++it) {
      link_terms(ctl, s_le);
      if (tid == 0) {
        ctl.qmax = 0;
        ctl.rho = 0;
      }
      __syncthreads();
      if (it == 0) {
        // computeLambdaInit: 1e-5 * largest |diagonal entry| of the Hessian
        const double* X = s_X + (size_t)ctl.cur * 3 * F;
        double acc[27], vmax = 0;
        for (int k = 0; k < 27; ++k) acc[k] = 0;
        for (int i = tid; i < N; i += kThreadsB) {
          const unsigned fl = s_flag[i];
          if ((fl & 3u) == 3u) continue;
          PtSys s;
          point_system(P, ctl.c1, ctl.c2, X + 3 * i, s_meas + 6 * i, fl, s, acc, acc + 21);
          vmax = fmax(vmax, fmax(fabs(s.V[0]), fmax(fabs(s.V[3]), fabs(s.V[5]))));
        }
        double d[6] = {acc[0], acc[6], acc[11], acc[15], acc[18], acc[20]};
        block_sum<6>(d, s_red);
        vmax = block_max(vmax, s_red);
        if (tid == 0) {
          const int dk[6] = {0, 6, 11, 15, 18, 20};
          for (int k = 0; k < 6; ++k) vmax = fmax(vmax, fabs(d[k] + ctl.Al[dk[k]]));
          ctl.lam = 1e-5 * vmax;
          ctl.ni = 2;
        }
        __syncthreads();
      }
      bool first = true;
#pragma unroll 1
      while (true) {
        const double lam = ctl.lam;
        const double* X = s_X + (size_t)ctl.cur * 3 * F;
        double* Xn = s_X + (size_t)(ctl.cur ^ 1) * 3 * F;
        {
          // build + Schur: acc = A (21) | Schur share of the right-hand side (6) | raw camera gradient (6) | chi2
          double acc[34];
          for (int k = 0; k < 34; ++k) acc[k] = 0;
          for (int i = tid; i < N; i += kThreadsB) {
            const unsigned fl = s_flag[i];
            if ((fl & 3u) == 3u) continue;
            PtSys s;
            acc[33] += point_system(P, ctl.c1, ctl.c2, X + 3 * i, s_meas + 6 * i, fl, s, acc, acc + 27);
            double Vi[6], WV[18];
            inv3_sym(s.V, lam, Vi);
            for (int r = 0; r < 6; ++r) sym3_mul(Vi, s.W + 3 * r, WV + 3 * r);
            int k = 0;
            for (int r = 0; r < 6; ++r) {
              for (int j = r; j < 6; ++j, ++k)
                acc[k] -= WV[3 * r] * s.W[3 * j] + WV[3 * r + 1] * s.W[3 * j + 1] + WV[3 * r + 2] * s.W[3 * j + 2];
              acc[21 + r] -= WV[3 * r] * s.bl[0] + WV[3 * r + 1] * s.bl[1] + WV[3 * r + 2] * s.bl[2];
            }
          }
          block_sum<34>(acc, s_red);
          if (tid == 0) {
            double S[21], rhs[6];
            for (int k = 0; k < 21; ++k) S[k] = acc[k] + ctl.Al[k];
            for (int k = 0; k < 6; ++k) {
              ctl.bc[k] = acc[27 + k] + ctl.bcl[k];
              rhs[k] = ctl.bc[k] + acc[21 + k];
            }
            if (first) ctl.cur_chi = acc[33] + ctl.chil;
            ctl.ok = cholesky6(S, lam, rhs, ctl.xp) ? 1 : 0;
            if (ctl.ok) {
              ctl.cn.p = se3_update(ctl.c2.p, ctl.xp);
              cam_finish(ctl.cn);
            }
          }
          __syncthreads();
        }
        first = false;
        double part[2] = {0, 0};   // dx . (lambda dx + b) over the points, robust chi2 at the trial estimate
        // the link edge at the trial pose: by one thread of another warp, under the pass below, not by thread 0 after it
        if (ctl.ok && tid == kThreadsB - 1) ctl.lchi_n = link_chi(ctl, ctl.cn.p);
        if (ctl.ok) {
          double xp[6];
          for (int k = 0; k < 6; ++k) xp[k] = ctl.xp[k];
          for (int i = tid; i < N; i += kThreadsB) {
            const unsigned fl = s_flag[i];
            if ((fl & 3u) == 3u) {
              for (int k = 0; k < 3; ++k) Xn[3 * i + k] = X[3 * i + k];
              continue;
            }
            PtSys s;
            point_system(P, ctl.c1, ctl.c2, X + 3 * i, s_meas + 6 * i, fl, s, nullptr, nullptr);
            double Vi[6], r[3], xl[3];
            inv3_sym(s.V, lam, Vi);
            for (int k = 0; k < 3; ++k) {
              double v = s.bl[k];
              for (int j = 0; j < 6; ++j) v -= s.W[3 * j + k] * xp[j];
              r[k] = v;
            }
            sym3_mul(Vi, r, xl);
            double xn[3];
            for (int k = 0; k < 3; ++k) {
              xn[k] = X[3 * i + k] + xl[k];
              Xn[3 * i + k] = xn[k];
              part[0] += xl[k] * (lam * xl[k] + s.bl[k]);
            }
            for (int c = 0; c < 2; ++c) {
              if ((fl >> c) & 1u) continue;
              Edge g;
              double w;
              edge_error(P, c ? ctl.cn : ctl.c1, xn, s_meas + 6 * i + 3 * c, (fl >> (2 + c)) & 1u, g, false);
              part[1] += huber(g.chi, P.delta, w);
            }
          }
        }
        block_sum<2>(part, s_red);
        if (tid == 0) {
          double tmp_chi = kHuge, scale = 0;
          if (ctl.ok) {
            tmp_chi = part[1] + ctl.lchi_n;
            scale = part[0];
            for (int k = 0; k < 6; ++k) scale += ctl.xp[k] * (lam * ctl.xp[k] + ctl.bc[k]);
          }
          const double rho = (ctl.cur_chi - tmp_chi) / (scale + 1e-3);
          ctl.rho = rho;
          bool brk = false;
          if (rho > 0 && isfinite(tmp_chi)) {   // good step
            double alpha = 1.0 - (2 * rho - 1) * (2 * rho - 1) * (2 * rho - 1);
            alpha = fmin(alpha, 2.0 / 3.0);
            ctl.lam = lam * fmax(1.0 / 3.0, alpha);
            ctl.ni = 2;
            ctl.cur_chi = tmp_chi;
            ctl.c2 = ctl.cn;
            ctl.cur ^= 1;
          } else {
            ctl.lam = lam * ctl.ni;
            ctl.ni *= 2;
            if (!isfinite(ctl.lam)) brk = true;
          }
          if (!brk) ctl.qmax++;
          ctl.again = (!brk && rho < 0 && ctl.qmax < 10) ? 1 : 0;
        }
        __syncthreads();
        if (!ctl.again) break;
      }
      if (tid == 0) {
        ctl.iters++;
        ctl.stop = (ctl.qmax == 10 || ctl.rho == 0 || !isfinite(ctl.lam)) ? 1 : 0;
      }
      __syncthreads();
      if (ctl.stop) break;
    }
    // ---- computeActiveErrors / activeRobustChi2, then the edges whose chi2 exceeds the kernel's delta leave
    {
      const double* X = s_X + (size_t)ctl.cur * 3 * F;
      double chi[1] = {0};
      for (int i = tid; i < N; i += kThreadsB) {
        unsigned fl = s_flag[i];
        for (int c = 0; c < 2; ++c) {
          if ((fl >> c) & 1u) continue;
          Edge g;
          double w;
          edge_error(P, c ? ctl.c2 : ctl.c1, X + 3 * i, s_meas + 6 * i + 3 * c, (fl >> (2 + c)) & 1u, g, false);
          chi[0] += huber(g.chi, P.delta, w);
          if (robust && g.chi > P.delta) fl |= 1u << c;
        }
        s_flag[i] = (uint8_t)fl;
      }
      block_sum<1>(chi, s_red);
      if (tid == 0) {
        const double chi2 = chi[0] + link_chi(ctl, ctl.c2.p);
        ctl.chi2 = chi2;
        if (isnan(chi2) || (round == 0 && (chi2 > 1e12 || !isfinite(chi2)))) ctl.fail = 1;
      }
      __syncthreads();
      if (ctl.fail) break;
    }
  }
  if (tid == 0 && ctl.chi2 > 1e12) ctl.fail = 1;
  __syncthreads();

  // ---- results: flagged inliers leave the mask; camera 2 * localTransform^-1 is the transform
  double cnt[1] = {0};
  if (!ctl.fail)
    for (int i = tid; i < N; i += kThreadsB)
      if (s_flag[i] & 3u) {
        const int m = s_idx[i];
        atomicAnd(&mask[m >> 5], ~(1u << (m & 31)));
        cnt[0] += 1;
      }
  block_sum<1>(cnt, s_red);
  if (tid == 0) {
    const int kept = N - (int)cnt[0];
    float* T = transform + 12 * (size_t)pair;
    if (stats) {
      stats[4 * pair] = ctl.iters;
      stats[4 * pair + 1] = ctl.chi2;
      stats[4 * pair + 2] = cnt[0];
      stats[4 * pair + 3] = ctl.lam;
    }
    if (ctl.fail || kept < P.min_inliers) {
      status[pair] = ctl.fail ? 4 : 3;
      n_inliers[pair] = kept;
      for (int i = 0; i < 12; ++i) T[i] = 0.f;
      return;
    }
    n_inliers[pair] = kept;
    const Cam& c = ctl.c2;
    for (int i = 0; i < 3; ++i) {
      double r[3];
      for (int j = 0; j < 3; ++j) {
        r[j] = c.R[3 * i] * P.LR[3 * j] + c.R[3 * i + 1] * P.LR[3 * j + 1] + c.R[3 * i + 2] * P.LR[3 * j + 2];   // R_c L^T
        T[4 * i + j] = (float)r[j];
      }
      T[4 * i + 3] = (float)(c.p.t[i] - (r[0] * P.Lt[0] + r[1] * P.Lt[1] + r[2] * P.Lt[2]));
    }
    // pose = (rvec, tvec) of world -> camera 2: rotation conj(q), translation -R^T t
    double v[3] = {-c.p.q[0], -c.p.q[1], -c.p.q[2]}, w = c.p.q[3];
    if (w < 0) { w = -w; v[0] = -v[0]; v[1] = -v[1]; v[2] = -v[2]; }
    const double nv = sqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]);
    const double k = nv > 1e-12 ? 2.0 * atan2(nv, w) / nv : 2.0;
    for (int i = 0; i < 3; ++i) {
      pose[6 * pair + i] = k * v[i];
      pose[6 * pair + 3 + i] = -c.rt[i];
    }
  }
}

}  // namespace

int sepb_launch_bundle_adjust(sepb_ctx* ctx, FramesView from, FramesView to, const int32_t* d_pairs, int n_pairs,
                              const int32_t* d_n_matches, const uint16_t* d_match_from, const uint16_t* d_match_to,
                              const double* d_cov, const sepb_pnp_params* pprm, const sepb_ba_params* bprm,
                              int32_t* d_status, int32_t* d_n_inliers, double* d_pose, float* d_transform,
                              uint32_t* d_inlier_mask, int mask_words, double* d_stats) {
  if (n_pairs == 0) return SEPB_OK;
  SEPB_REQUIRE(ctx, from.max_feat <= 2048, SEPB_E_SHAPE);   // inlier list offsets: 64 mask words
  BaKernelParams P;
  P.fx = pprm->fx; P.fy = pprm->fy; P.cx = pprm->cx; P.cy = pprm->cy;
  P.baseline = bprm->baseline;
  P.var = bprm->pixel_variance;
  P.delta = bprm->robust_kernel_delta;
  for (int i = 0; i < 9; ++i) P.LR[i] = pprm->local_R[i];
  for (int i = 0; i < 3; ++i) P.Lt[i] = pprm->local_t[i];
  P.iterations = bprm->iterations;
  P.min_inliers = pprm->min_inliers;
  P.use_cov = bprm->variance_ignored ? 0 : 1;
  const size_t smem = (size_t)from.max_feat * (48 + 24 + 2 + 1);
  if (smem > 48 * 1024 && smem > ctx->cfg_ba_smem) {
    SEPB_CUDA(ctx, cudaFuncSetAttribute(bundle_adjust_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    ctx->cfg_ba_smem = smem;
  }
  bundle_adjust_kernel<<<n_pairs, kThreadsB, smem, ctx->stream>>>(from, to, d_pairs, n_pairs, d_n_matches, d_match_from,
                                                                 d_match_to, d_cov, P, d_status, d_n_inliers, d_pose,
                                                                 d_transform, d_inlier_mask, mask_words, d_stats);
  ctx->launches++;
  SEPB_CUDA(ctx, cudaGetLastError());
  return SEPB_OK;
}
