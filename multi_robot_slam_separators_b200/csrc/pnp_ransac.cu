// K3: batched PnP-RANSAC + ITERATIVE refit, one warp per keyframe pair at a time (persistent work-queue kernel).
//
// Replaces util3d::estimateMotion3DTo2D -> cv::solvePnPRansac (useExtrinsicGuess, 5-point EPnP
// minimal sets, confidence 0.99) called at R/src/myRegistrationVis.cpp:1077-1090.
//
// cv::solvePnPRansac is a sequential loop, but its hypotheses do not depend on loop state: the
// RNG stream is consumed unconditionally, so hypothesis h is a pure function of (match count, h).
// A warp therefore SOLVES 32 consecutive hypotheses at once (one lane each; EPnP's summation order
// must be replayed serially, see cvmath.cuh, so a hypothesis cannot be split across lanes), then
// SCORES them one after the other with all 32 lanes striding over the correspondences (warp per
// hypothesis), applying OpenCV's `goodCount > max(best, 4)` rule and RANSACUpdateNumIters after each
// one, which reproduces the sequential result exactly and keeps most of its early exit.  For the same
// reason whole chunks of a pair can be computed by other warps (see "K3 scheduling" below).
// Compiled with -fmad=false (see build.py): no multiply-add contraction, like the cv2 build.
#include <cstdlib>
#include <algorithm>

#include "common.cuh"

// developer build (SEPB_K3_PROFILE=1 python -m ...build): cycle counters per phase, read by sepb_debug_k3_prof
#ifdef SEPB_K3_PROFILE
__device__ unsigned long long g_k3_prof[8];
#define K3_T0(v) const long long v = clock64()
#define K3_ACC(slot, v) if ((threadIdx.x & 31) == 0) atomicAdd(&g_k3_prof[slot], (unsigned long long)(clock64() - (v)))
#else
#define K3_T0(v)
#define K3_ACC(slot, v)
#endif

#include "epnp.cuh"
#include "pnp_geom.cuh"
#include "p3p.cuh"

namespace {

constexpr int kWarp = 32;
constexpr int kSolveLanes = cvm::kHeapStride;   // hypotheses solved at once by a warp (one lane each)
constexpr int kMaskWords = SEPB_MAX_FEATURES / 32;
#ifndef SEPB_SCORE_ILP
#define SEPB_SCORE_ILP 1
#endif
constexpr int kScoreIlp = SEPB_SCORE_ILP;   // correspondences per lane scored together (developer A/B)

struct PnpKernelParams {
  cvm::Intrinsics K;
  int iterations;
  float thr2;          // (float)(reprojError^2)
  double confidence;
  int min_inliers;
  double rvec0[3], tvec0[3];   // guess, camera frame (== (guess * local)^-1)
  double local_R[9], local_t[3];
  int table_h;                 // hypotheses per row of the sample table
  int minimal_model;           // OpenCV 3.2 semantics: return the winning minimal-set model instead of the refit
  int refine_iterations;       // RTAB-Map refineModel rounds after the refit (ITERATIVE only)
  float refine_sigma, reproj_error;
};

// getSubset index stream: table[count][h][mp] (mp = model points: 5, or 4 with the P3P flags); thread per count value
__global__ void sample_table_kernel(uint16_t* table, int max_count, int H, int mp) {
  const int count = blockIdx.x * blockDim.x + threadIdx.x;
  if (count > max_count) return;
  uint16_t* row = table + (size_t)count * H * mp;
  if (count < mp) {
    for (int i = 0; i < H * mp; ++i) row[i] = 0;
    return;
  }
  cvm::Rng rng(0xFFFFFFFFFFFFFFFFULL);
  for (int h = 0; h < H; ++h) {
    int idx[5];
    for (int i = 0; i < mp; ++i) {
      int v;
      bool dup;
      do {
        v = (int)(rng.next() % (unsigned)count);
        dup = false;
        for (int j = 0; j < i; ++j) dup |= idx[j] == v;
      } while (dup);
      idx[i] = v;
      row[h * mp + i] = (uint16_t)v;
    }
  }
}

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xFFFFFFFFu, v, o);
  return v;
}

// The part of the scratch heap above what the refit's 6x6 solve touches (78 strided doubles) doubles as a staging
// area for the pair's correspondences while no EPnP solve is running: scoring and the refit then read them with
// conflict-free LDS instead of two dependent gathers through L2 (four resident pairs overflow the small L1).
constexpr int kStageOffBytes = 80 * cvm::kHeapStride * 8;
constexpr int kStageCap = (cvm::kEpnpHeap * cvm::kHeapStride * 8 - kStageOffBytes) / 20;   // correspondences

struct Corr {  // correspondence accessor of one pair
  const float* kp3d;   // From frame 3-D points
  const float* kp2d;   // To frame pixels
  const uint16_t* mf;
  const uint16_t* mt;
  const float* st;     // staged copy (SoA, stride kStageCap) or nullptr when the pair has too many correspondences
  __device__ __forceinline__ void load_global(int m, float* o3, float* i2) const {
    const int f = mf[m], t = mt[m];
    o3[0] = kp3d[3 * f]; o3[1] = kp3d[3 * f + 1]; o3[2] = kp3d[3 * f + 2];
    i2[0] = kp2d[2 * t]; i2[1] = kp2d[2 * t + 1];
  }
  __device__ __forceinline__ void load(int m, float* o3, float* i2) const {
    if (st) {
      o3[0] = st[m]; o3[1] = st[kStageCap + m]; o3[2] = st[2 * kStageCap + m];
      i2[0] = st[3 * kStageCap + m]; i2[1] = st[4 * kStageCap + m];
    } else {
      load_global(m, o3, i2);
    }
  }
  // (re)fill the staging area; the whole warp calls it whenever the heap is not in use by a solve
  __device__ __forceinline__ void stage(float* dst, int M) const {
    if (!st) return;
    for (int m = threadIdx.x & 31; m < M; m += 32) {
      float o3[3], i2[2];
      load_global(m, o3, i2);
      dst[m] = o3[0]; dst[kStageCap + m] = o3[1]; dst[2 * kStageCap + m] = o3[2];
      dst[3 * kStageCap + m] = i2[0]; dst[4 * kStageCap + m] = i2[1];
    }
    __syncwarp();
  }
};

// residuals / normal equations of the ITERATIVE refit over the inlier set, warp-cooperative
__device__ __noinline__ void lm_accumulate(const Corr& cr, const uint32_t* mask, int M, const cvm::Intrinsics& K, const double* param,
                              bool want_jac, double* JtJ /*21 upper*/, double* JtE /*6*/, double& err2) {
  const int lane = threadIdx.x & 31;
  double R[9], dRdr[27];
  cvm::rodrigues_vec_to_mat(param, R, want_jac ? dRdr : nullptr);
  double aJ[21], aE[6], e2 = 0;
#pragma unroll
  for (int i = 0; i < 21; ++i) aJ[i] = 0;
#pragma unroll
  for (int i = 0; i < 6; ++i) aE[i] = 0;
  for (int base = 0; base < M; base += 32) {
    const int m = base + lane;
    if (m >= M || !((mask[base >> 5] >> lane) & 1u)) continue;
    float o3[3], i2[2];
    cr.load(m, o3, i2);
    const double X = o3[0], Y = o3[1], Z = o3[2];
    double x = R[0] * X + R[1] * Y + R[2] * Z + param[3];
    double y = R[3] * X + R[4] * Y + R[5] * Z + param[4];
    double z = R[6] * X + R[7] * Y + R[8] * Z + param[5];
    z = z ? 1. / z : 1;
    x *= z;
    y *= z;
    const double ex = (x * K.fx + K.cx) - (double)i2[0], ey = (y * K.fy + K.cy) - (double)i2[1];
    e2 += ex * ex + ey * ey;
    if (want_jac) {
      double jx[6], jy[6];
      jx[3] = K.fx * z; jx[4] = 0; jx[5] = K.fx * (-x * z);
      jy[3] = 0; jy[4] = K.fy * z; jy[5] = K.fy * (-y * z);
#pragma unroll
      for (int j = 0; j < 3; ++j) {
        const double* d = dRdr + 9 * j;
        const double dx0 = X * d[0] + Y * d[1] + Z * d[2];
        const double dy0 = X * d[3] + Y * d[4] + Z * d[5];
        const double dz0 = X * d[6] + Y * d[7] + Z * d[8];
        jx[j] = K.fx * (z * (dx0 - x * dz0));
        jy[j] = K.fy * (z * (dy0 - y * dz0));
      }
      int k = 0;
#pragma unroll
      for (int a = 0; a < 6; ++a) {
        aE[a] += jx[a] * ex + jy[a] * ey;
#pragma unroll
        for (int b = a; b < 6; ++b) aJ[k++] += jx[a] * jx[b] + jy[a] * jy[b];
      }
    }
  }
  err2 = warp_sum(e2);
  if (want_jac) {
#pragma unroll
    for (int i = 0; i < 21; ++i) JtJ[i] = warp_sum(aJ[i]);
#pragma unroll
    for (int i = 0; i < 6; ++i) JtE[i] = warp_sum(aE[i]);
  }
}

#ifndef SEPB_LM_WARP_SVD
#define SEPB_LM_WARP_SVD 1   // 0 (developer A/B): every lane runs the serial 6 x 6 SVD on its own copy
#endif
#if !SEPB_LM_WARP_SVD
// CvLevMarq::step(): (JtJ with diag *= 1 + lambda) x = JtErr by SVD; param = prev - x
__device__ __noinline__ void lm_step(cvm::Heap h, const double* JtJ, const double* JtE, int lambda_lg10, const double* prev,
                        double* param) {
  const double lambda = exp(lambda_lg10 * log(10.));
  cvm::Heap At = h.at(0), W = h.at(36), Vt = h.at(42);
  int k = 0;
  for (int a = 0; a < 6; ++a)
    for (int b = a; b < 6; ++b) {
      const double v = JtJ[k++];
      At[a * 6 + b] = v;
      At[b * 6 + a] = v;
    }
  for (int a = 0; a < 6; ++a) At[a * 6 + a] = At[a * 6 + a] * (1. + lambda);
  cvm::jacobi_svd<6, 6, true>(At, W, Vt);
  double x[6];
  cvm::svd_backsubst<6, 6>(At, W, Vt, JtE, x);
  for (int i = 0; i < 6; ++i) param[i] = prev[i] - x[i];
}

#endif
// CvLevMarq::step() for the whole warp: every lane holds the same normal equations and needs the new parameters.
// h0 = lane 0's scratch slice: ONE copy of the 6 x 6 problem, its Jacobi sweeps spread over lanes (jacobi_svd6_warp).
__device__ __forceinline__ void lm_step_warp(cvm::Heap h, cvm::Heap h0, const double* JtJ, const double* JtE, int lambda_lg10,
                                             const double* prev, double* param) {
#if SEPB_LM_WARP_SVD
  (void)h;
  const double lambda = exp(lambda_lg10 * log(10.));
  cvm::Heap At = h0.at(0), W = h0.at(36), Vt = h0.at(42);
  if ((threadIdx.x & 31) == 0) {
    int k = 0;
    for (int a = 0; a < 6; ++a)
      for (int b = a; b < 6; ++b) {
        const double v = JtJ[k++];
        At[a * 6 + b] = v;
        At[b * 6 + a] = v;
      }
    for (int a = 0; a < 6; ++a) At[a * 6 + a] = At[a * 6 + a] * (1. + lambda);
  }
  __syncwarp();
  cvm::jacobi_svd6_warp(At, W, Vt);
  if ((threadIdx.x & 31) == 0) {
    double x[6];
    cvm::svd_backsubst<6, 6>(At, W, Vt, JtE, x);
    for (int i = 0; i < 6; ++i) param[i] = prev[i] - x[i];
  }
  __syncwarp();
#else
  (void)h0;
  lm_step(h, JtJ, JtE, lambda_lg10, prev, param);
  __syncwarp();
#endif
#pragma unroll
  for (int i = 0; i < 6; ++i) param[i] = __shfl_sync(0xFFFFFFFFu, param[i], 0);
}

// solvePnP(points under `mask`, SOLVEPNP_ITERATIVE, useExtrinsicGuess) == CvLevMarq started at `param`; whole warp,
// every lane ends with the same refined [rvec | tvec] in `param`
__device__ __noinline__ void lm_refit(const Corr& cr, const uint32_t* mask, int M, const cvm::Intrinsics& K, cvm::Heap heap,
                                      cvm::Heap h0, double* param) {
  double prev[6], JtJ[21], JtE[6];
  for (int i = 0; i < 6; ++i) prev[i] = param[i];
  int lambda_lg10 = -3, iters = 0;
  double err2 = 0, prev_err_norm = 0;
  const int max_iter = 20;
  const double eps = 1.1920928955078125e-07;  // FLT_EPSILON
  lm_accumulate(cr, mask, M, K, param, true, JtJ, JtE, err2);  // STARTED -> CALC_J
  for (;;) {
    // CALC_J
    for (int i = 0; i < 6; ++i) prev[i] = param[i];
    lm_step_warp(heap, h0, JtJ, JtE, lambda_lg10, prev, param);
    if (iters == 0) prev_err_norm = sqrt(err2);
    double dJ[21], dE[6];
    lm_accumulate(cr, mask, M, K, param, false, dJ, dE, err2);
    // CHECK_ERR
    double err_norm = sqrt(err2);
    while (err_norm > prev_err_norm && ++lambda_lg10 <= 16) {
      lm_step_warp(heap, h0, JtJ, JtE, lambda_lg10, prev, param);
      lm_accumulate(cr, mask, M, K, param, false, dJ, dE, err2);
      err_norm = sqrt(err2);
    }
    lambda_lg10 = max(lambda_lg10 - 1, -16);
    double dn = 0, pn = 0;
    for (int i = 0; i < 6; ++i) {
      dn += (param[i] - prev[i]) * (param[i] - prev[i]);
      pn += prev[i] * prev[i];
    }
    if (++iters >= max_iter || sqrt(dn) / (sqrt(pn) + cvm::kDblEps) < eps) break;
    prev_err_norm = err_norm;
    lm_accumulate(cr, mask, M, K, param, true, JtJ, JtE, err2);
  }
}

// RTAB-Map util3d::solvePnPRansac, `refineIterations > 0` (a clone of PCL's SampleConsensus::refineModel; SURVEY.md A.3
// step 3, reached from R/src/myRegistrationVis.cpp:1085): refit on the current inliers from the current pose, re-select
// the inliers of ALL points with the current threshold (strict <, on the f32 pixel distance), shrink the threshold to
// min(reprojError, sigma * std of all errors), until the set stops changing, oscillates, or the round limit is hit.
// prev / other: the two mask buffers (prev holds the RANSAC consensus set on entry).  On return *final points at the
// buffer that refineModel leaves in `inliers` (the set the last fit ran on) and the pose is in `param`.
__device__ __noinline__ int refine_model(const Corr& cr, int M, const PnpKernelParams& prm, cvm::Heap heap, cvm::Heap h0,
                                         uint32_t* prev, uint32_t* other, int prev_count, double* param,
                                         const uint32_t** final_mask) {
  const int lane = threadIdx.x & 31;
  float thr = prm.reproj_error;
  int it = 0, n_sizes = 0, sizes[4] = {0, 0, 0, 0};
  bool changed = false;
  uint32_t* cur_new = other;
  int new_count = 0;
  for (;;) {
    lm_refit(cr, prev, M, prm.K, heap, h0, param);
    sizes[0] = sizes[1]; sizes[1] = sizes[2]; sizes[2] = sizes[3]; sizes[3] = prev_count;
    ++n_sizes;
    double R[9];
    cvm::rodrigues_vec_to_mat(param, R, nullptr);
    // errors of all points: new inlier set, and their sum in index order (uMean is a sequential f32 sum)
    float sum = 0.f;
    new_count = 0;
    for (int base = 0; base < M; base += 32) {
      const int m = base + lane;
      float o3[3], i2[2];
      cr.load(min(m, M - 1), o3, i2);
      const float e = sqrtf(cvm::reproj_err_f32(R, param + 3, prm.K, o3, i2));
      const uint32_t b = __ballot_sync(0xFFFFFFFFu, e < thr && m < M);
      if (lane == 0) cur_new[base >> 5] = b;
      new_count += __popc(b);
      for (int k = 0; k < 32 && base + k < M; ++k) sum += __shfl_sync(0xFFFFFFFFu, e, k);
    }
    __syncwarp();
    bool to_cond = false;
    if (new_count < prm.min_inliers) {
      ++it;
      if (it >= prm.refine_iterations) break;
      to_cond = true;
    }
    if (!to_cond) {
      const float mean = sum / (float)(unsigned)M;
      float sum2 = 0.f;
      for (int base = 0; base < M; base += 32) {
        const int m = base + lane;
        float o3[3], i2[2];
        cr.load(min(m, M - 1), o3, i2);
        const float e = sqrtf(cvm::reproj_err_f32(R, param + 3, prm.K, o3, i2));
        const float d2 = (e - mean) * (e - mean);
        for (int k = 0; k < 32 && base + k < M; ++k) sum2 += __shfl_sync(0xFFFFFFFFu, d2, k);
      }
      const float variance = M > 1 ? sum2 / (float)(unsigned)(M - 1) : 0.f;
      thr = fminf(prm.reproj_error, prm.refine_sigma * sqrtf(variance));
      changed = false;
      {   // std::swap(prev_inliers, new_inliers)
        uint32_t* t = prev; prev = cur_new; cur_new = t;
        const int c = prev_count; prev_count = new_count; new_count = c;
      }
      if (new_count != prev_count) {
        if (n_sizes >= 4 && sizes[3] == sizes[1] && sizes[2] == sizes[0]) break;   // oscillating between two sizes
        changed = true;
      } else {
        for (int w = 0; w < (M + 31) / 32; ++w) changed = changed || prev[w] != cur_new[w];
      }
    }
    if (!(changed && ++it < prm.refine_iterations)) break;
  }
  *final_mask = cur_new;
  return new_count;
}

// ---- K3 scheduling ---------------------------------------------------------------------------------------------
// A pair's cost is its number of 32-hypothesis chunks (1..10 at 300 iterations), unknown until RANSAC runs, and one
// chunk is ~0.7 ms of serial f64 work for a warp: a pair that needs every chunk is a 7 ms dependency chain, longer than
// the rest of the batch.  But hypothesis h is a pure function of (match count, h) and so is its inlier count, hence
// the chunks of a pair can be computed by DIFFERENT warps at the same time and only the cheap scan over the counts
// (OpenCV's `>` rule + RANSACUpdateNumIters, in hypothesis order) has to be sequential.
//
// pnp_persist_kernel: one CTA per SM, kPersistWarps warps, every warp a worker:
//   * owner:  pops pairs from a list (pnp_prepare_kernel compacts the pairs that run RANSAC, the others get their null
//             result there), runs the chunks of its pair one after the other -- solve 32 hypotheses (lane each), score
//             them in order (warp each) with the live early exit -- then the refit;
//   * helper: once the list is empty a worker looks for the owner with the most iterations left, claims that pair's
//             next chunk (atomic counter per owner), solves + scores it and leaves {32 inlier counts, 32 poses} in the
//             owner's slot array.  The owner claims chunks from the same counter; when it finds chunks taken by
//             helpers it computes the one it got into a slot too, then consumes the slots in hypothesis order (the
//             same scan), and re-scores only the winning hypothesis to get its mask.
// A helper exists only after the list ran empty, so an owner being helped is on its last pair and its slots are
// written at most once per launch.  Results are bit-identical to the one-warp-per-pair kernel (pnp_ransac_kernel).
constexpr int kPersistWarps = 5;      // workers per SM (shared-memory bound: 44.5 KB heap each)
constexpr int kMaxHelpChunks = 64;    // slots per worker; more chunks than that (iterations > 2048): no helping

struct PnpArgs {
  FramesView from, to;
  const int32_t* pairs;
  int n_pairs;
  const int32_t* n_matches;
  const uint16_t* match_from;
  const uint16_t* match_to;
  const int32_t* n_unique_to;
  const uint16_t* table;
  const int32_t* custom_samples;
  int custom_h;
  const double* guess;          // [P][6] or null
  const int32_t* only_if_ok;    // second call of estimate_transformation: pairs whose first call failed keep its result
  PnpKernelParams prm;
  int32_t* status;
  int32_t* n_inliers;
  double* pose;                 // [P][6]
  float* transform;             // [P][12]
  uint32_t* inlier_mask;
  int mask_words;
  int32_t* best_hyp;
  int32_t* iters_run;
};

struct PnpQueue {          // device memory, head (ctrl, wk, slot_ready) zeroed before every launch
  int32_t* ctrl;           // [0] list cursor, [1] pairs listed, [2] pairs finished
  int32_t* wk;             // [n_workers][4]: pair, next chunk to claim, published iteration bound, accepting help
  int32_t* slot_ready;     // [n_workers][max_chunks]
  int32_t* list;           // [n_pairs] pairs that run RANSAC
  int32_t* slot_good;      // [n_workers][max_chunks][32] inlier counts
  double* slot_rt;         // [n_workers][max_chunks][12][32] R (9) | t (3), hypothesis-minor
  double* pre_rt;          // [n_workers][12][32] poses of the NEXT pair's first hypotheses, solved by the spare lanes of a
                           // worker's last chunk (null: off)
  int n_workers, max_chunks;   // max_chunks == 0: helping disabled
};

__device__ __forceinline__ int ld_volatile(const int32_t* p) { return *reinterpret_cast<const volatile int32_t*>(p); }
__device__ __forceinline__ void st_volatile(int32_t* p, int v) { *reinterpret_cast<volatile int32_t*>(p) = v; }
// one lane reads, every lane gets the same value (control flow that depends on it stays warp-uniform)
__device__ __forceinline__ int ld_volatile_warp(const int32_t* p) {
  int v = 0;
  if ((threadIdx.x & 31) == 0) v = ld_volatile(p);
  return __shfl_sync(0xFFFFFFFFu, v, 0);
}

struct PairIn {
  int pair, fi, ti, M, nf, nu;
  bool valid;
  Corr cr;
  double g[6];
};

__device__ __forceinline__ PairIn load_pair(const PnpArgs& a, int pair, float* stage_buf) {
  PairIn in;
  in.pair = pair;
  in.fi = a.pairs[2 * pair];
  in.ti = a.pairs[2 * pair + 1];
  in.valid = in.fi >= 0 && in.fi < a.from.size && in.ti >= 0 && in.ti < a.to.size;
  if (!in.valid) in.fi = in.ti = 0;
  in.M = in.valid ? a.n_matches[pair] : 0;
  in.nf = in.valid ? a.from.count[in.fi] : 0;
  in.nu = a.n_unique_to ? a.n_unique_to[pair] : in.M;
  in.cr = Corr{a.from.kp3d + (size_t)in.fi * a.from.max_feat * 3, a.to.kp2d + (size_t)in.ti * a.to.max_feat * 2,
               a.match_from + (size_t)pair * a.from.max_feat, a.match_to + (size_t)pair * a.from.max_feat,
               in.M <= kStageCap ? stage_buf : nullptr};
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    in.g[i] = a.guess ? a.guess[6 * pair + i] : a.prm.rvec0[i];
    in.g[3 + i] = a.guess ? a.guess[6 * pair + 3 + i] : a.prm.tvec0[i];
  }
  return in;
}

// myRegistrationVis.cpp:1069-1070 and estimateMotion3DTo2D: >= minInliers on every side (and a minimal set)
__device__ __forceinline__ bool pair_runs(const PnpArgs& a, const PairIn& in) {
  return !(in.M < a.prm.min_inliers || in.nf < a.prm.min_inliers || in.nu < a.prm.min_inliers || in.M < 4);
}

// result of one pair (one thread)
__device__ __noinline__ void write_result(const PnpArgs& a, int pair, int st, int ninl, const double* p6, int bh, int it) {
  a.status[pair] = st;
  a.n_inliers[pair] = ninl;
  if (a.best_hyp) a.best_hyp[pair] = bh;
  if (a.iters_run) a.iters_run[pair] = it;
  for (int i = 0; i < 6; ++i) a.pose[6 * pair + i] = p6[i];
  float* T = a.transform + 12 * (size_t)pair;
  if (st == 1) {
    // T = (local * [R|t])^-1  (maps To-base coordinates into the From base frame)
    const PnpKernelParams& prm = a.prm;
    double R[9], RP[9], tP[3];
    cvm::rodrigues_vec_to_mat(p6, R, nullptr);
    for (int i = 0; i < 3; ++i) {
      for (int j = 0; j < 3; ++j)
        RP[3 * i + j] = prm.local_R[3 * i] * R[j] + prm.local_R[3 * i + 1] * R[3 + j] + prm.local_R[3 * i + 2] * R[6 + j];
      tP[i] = prm.local_R[3 * i] * p6[3] + prm.local_R[3 * i + 1] * p6[4] + prm.local_R[3 * i + 2] * p6[5] + prm.local_t[i];
    }
    for (int i = 0; i < 3; ++i) {
      for (int j = 0; j < 3; ++j) T[4 * i + j] = (float)RP[3 * j + i];
      T[4 * i + 3] = (float)(-(RP[i] * tP[0] + RP[3 + i] * tP[1] + RP[6 + i] * tP[2]));
    }
  } else {
    for (int i = 0; i < 12; ++i) T[i] = 0.f;  // null transform (Transform::isNull())
  }
}

// thread per pair: pairs that run RANSAC go to the work list, the others get their null result here
__global__ void __launch_bounds__(256) pnp_prepare_kernel(PnpArgs a, PnpQueue q) {
  const int pair = blockIdx.x * blockDim.x + threadIdx.x;
  if (pair >= a.n_pairs) return;
  // second call (estimate_transformation's call with guess = first result): pairs whose first call failed keep it
  if (a.only_if_ok && a.only_if_ok[pair] != 1) return;
  const PairIn in = load_pair(a, pair, nullptr);
  if (pair_runs(a, in)) {
    q.list[atomicAdd(&q.ctrl[1], 1)] = pair;
    return;
  }
  if (a.inlier_mask)
    for (int w = 0; w < a.mask_words; ++w) a.inlier_mask[(size_t)pair * a.mask_words + w] = 0;
  write_result(a, pair, 0, 0, in.g, -1, 0);
}

struct WarpScratch {
  double* heap_base;    // this warp's strided scratch heap
  float* stage_buf;     // correspondence staging area inside it
  uint32_t* s_mask;     // [2][kMaskWords] ping-pong inlier masks
};

// ---- solve: lane <-> hypothesis.  h_cap: hypotheses at or past it are never consumed (their lanes redo h_cap - 1).
// MP = model points: 5 -> EPnP, 4 -> P3P (a set without a solution gets a NaN translation: it scores no inlier, which
// is what the skipped iteration of RANSACPointSetRegistrator::run amounts to).  rv_out (optional): the model as rvec.
// nx (optional): the pair this worker runs next.  The lanes past h_cap -- which would only repeat hypothesis h_cap - 1 --
// solve ITS hypotheses 0 .. instead (lane chunk0 + lane - h_cap -> hypothesis lane - (h_cap - chunk0)); a hypothesis is a
// pure function of (pair, index), so who solves it and when changes nothing.  Their poses go to `stash` ([12][32]).
template <int MP>
__device__ __forceinline__ void solve_chunk(const PnpArgs& a, const PairIn& in, const WarpScratch& ws, int chunk0, int h_cap,
                                            double* Rh, double* th, double* rv_out = nullptr, const PairIn* nx = nullptr,
                                            double* stash = nullptr) {
  const int lane = threadIdx.x & 31;
  const cvm::Heap heap{ws.heap_base + lane};
  int h = chunk0 + lane;
  const int useful = h_cap - chunk0;          // lanes [0, useful) belong to `in`
  const bool spare = nx && lane >= useful;
  if (h >= h_cap) h = h_cap - 1;
  float o[3 * MP], im[2 * MP];
  if (spare) {
    const int hn = lane - useful;
    for (int i = 0; i < MP; ++i) {
      const int s = a.custom_samples ? a.custom_samples[((size_t)nx->pair * a.custom_h + hn) * MP + i]
                                     : a.table[((size_t)nx->M * a.prm.table_h + hn) * MP + i];
      nx->cr.load_global(min(max(s, 0), nx->M - 1), o + 3 * i, im + 2 * i);
    }
  } else {
    for (int i = 0; i < MP; ++i) {
      const int s = a.custom_samples ? a.custom_samples[((size_t)in.pair * a.custom_h + h) * MP + i]
                                     : a.table[((size_t)in.M * a.prm.table_h + h) * MP + i];
      in.cr.load(min(max(s, 0), in.M - 1), o + 3 * i, im + 2 * i);
    }
  }
  __syncwarp();   // every lane has read its samples from the staging area before a solve overwrites it
  double R0[9], rv[3];
  K3_T0(t_e);
  bool solved = true;
  if (MP == 5) {
    cvm::epnp5(heap, o, im, a.prm.K, R0, th);
  } else {
    solved = cvm::p3p4(o, im, a.prm.K, R0, th);
    if (!solved) {
      for (int i = 0; i < 9; ++i) R0[i] = (i % 4 == 0) ? 1.0 : 0.0;
    }
  }
  K3_ACC(0, t_e);
  K3_T0(t_r);
  cvm::rodrigues_mat_to_vec(heap, R0, rv);   // runKernel hands the model on as rvec
  cvm::rodrigues_vec_to_mat(rv, Rh, nullptr);  // projectPoints turns it back into a matrix
  K3_ACC(1, t_r);
  if (!solved) th[0] = th[1] = th[2] = __longlong_as_double(0x7FF8000000000000LL);
  if (rv_out)
    for (int i = 0; i < 3; ++i) rv_out[i] = rv[i];
  if (spare) {
    const int hn = lane - useful;
#pragma unroll
    for (int i = 0; i < 9; ++i) stash[i * 32 + hn] = Rh[i];
#pragma unroll
    for (int i = 0; i < 3; ++i) stash[(9 + i) * 32 + hn] = th[i];
  }
  __syncwarp();
  in.cr.stage(ws.stage_buf, in.M);   // the solve overwrote the staging area
}

// ---- score: warp <-> hypothesis.  Inlier mask of pose (R, t) into cand[], returns the count (PnPRansacCallback::computeError)
__device__ __forceinline__ int score_one(const PnpArgs& a, const PairIn& in, const double* R, const double* t, uint32_t* cand) {
  const int lane = threadIdx.x & 31;
  int good = 0;
  // kScoreIlp independent correspondences per lane in flight: the projection is a dependent f64 chain with a divide
  for (int base = 0; base < in.M; base += 32 * kScoreIlp) {
    bool inl[kScoreIlp];
#pragma unroll
    for (int u = 0; u < kScoreIlp; ++u) {
      const int m = base + 32 * u + lane;
      float o3[3], i2[2];
      in.cr.load(min(m, in.M - 1), o3, i2);
      inl[u] = (cvm::reproj_err_f32(R, t, a.prm.K, o3, i2) <= a.prm.thr2) && m < in.M;
    }
#pragma unroll
    for (int u = 0; u < kScoreIlp; ++u) {
      const uint32_t b = __ballot_sync(0xFFFFFFFFu, inl[u]);
      if (base + 32 * u < in.M) {
        if (lane == 0) cand[(base >> 5) + u] = b;
        good += __popc(b);
      }
    }
  }
  return good;
}

// a whole chunk for somebody else's scan: solve, score every hypothesis, leave counts and poses in the owner's slot
template <int MP>
__device__ __noinline__ void chunk_to_slot(const PnpArgs& a, const PnpQueue& q, const PairIn& in, const WarpScratch& ws,
                                           int owner, int cc, int h_off, uint32_t* cand) {
  const int lane = threadIdx.x & 31;
  double Rh[9], th[3];
  // h_off: the pair's first h_off hypotheses were solved ahead of time by its owner's previous pair; chunk cc of the
  // pair covers hypotheses h_off + 32 cc ..
  solve_chunk<MP>(a, in, ws, h_off + cc * kSolveLanes, a.prm.iterations, Rh, th);
  K3_T0(t_s);
  int mygood = 0;
  for (int k = 0; k < kSolveLanes; ++k) {
    // the owner's published bound never undercuts its live iteration count: what lies past it is never consumed
    if (h_off + cc * kSolveLanes + k >= ld_volatile_warp(q.wk + 4 * owner + 2)) break;
    double R[9], t[3];
#pragma unroll
    for (int i = 0; i < 9; ++i) R[i] = __shfl_sync(0xFFFFFFFFu, Rh[i], k);
#pragma unroll
    for (int i = 0; i < 3; ++i) t[i] = __shfl_sync(0xFFFFFFFFu, th[i], k);
    const int good = score_one(a, in, R, t, cand);
    if (lane == k) mygood = good;
  }
  K3_ACC(2, t_s);
#ifdef SEPB_K3_PROFILE
  if (lane == 0) atomicAdd(&g_k3_prof[4], 1ull);
#endif
  const size_t slot = (size_t)owner * q.max_chunks + cc;
  q.slot_good[slot * 32 + lane] = mygood;
#pragma unroll
  for (int i = 0; i < 9; ++i) q.slot_rt[(slot * 12 + i) * 32 + lane] = Rh[i];
#pragma unroll
  for (int i = 0; i < 3; ++i) q.slot_rt[(slot * 12 + 9 + i) * 32 + lane] = th[i];
  __threadfence();
  __syncwarp();
  if (lane == 0) st_volatile(q.slot_ready + slot, 1);
}

// One keyframe pair, one warp.  worker >= 0: persistent kernel (chunks are claimed from the owner's counter and may
// arrive from helpers); worker < 0: the warp runs every chunk itself.
// pre: this pair's hypotheses [0, pre) are already solved (poses in the worker's stash).  next_pair / next_pre (persistent
// kernel): when the pair's LAST chunk has spare lanes, the worker pops its next pair early and those lanes solve its first
// hypotheses (*next_pair: -1 nothing popped, -2 the list is exhausted, else the pair; *next_pre: how many were solved).
// The spare lanes exist because a chunk costs the same whatever the number of useful lanes; the average pair wastes
// half a chunk of ~2.9 that way.
template <int MP>
__device__ __forceinline__ void run_pair(const PnpArgs& a, const PnpQueue& q, int worker, int pair, const WarpScratch& ws,
                                         int pre = 0, int* next_pair = nullptr, int* next_pre = nullptr) {
  const PnpKernelParams& prm = a.prm;
  const int lane = threadIdx.x & 31;
  const cvm::Heap heap{ws.heap_base + lane};
  uint32_t* s_mask = ws.s_mask;
  const PairIn in = load_pair(a, pair, ws.stage_buf);
  const int M = in.M;
  const Corr& cr = in.cr;
  uint32_t* out_mask = a.inlier_mask ? a.inlier_mask + (size_t)pair * a.mask_words : nullptr;
  if (out_mask)
    for (int w = lane; w < a.mask_words; w += 32) out_mask[w] = 0;

  if (M == 4) {  // count == 4: cv::solvePnPRansac turns to ONE P3P solve over the four, whatever the flags; all inliers
    float o[12], im[8];
    for (int i = 0; i < 4; ++i) cr.load_global(i, o + 3 * i, im + 2 * i);
    double R[9], p6[6];
    for (int i = 0; i < 6; ++i) p6[i] = in.g[i];
    const bool ok = cvm::p3p4(o, im, prm.K, R, p6 + 3);
    if (ok) cvm::rodrigues_mat_to_vec(heap, R, p6);
    if (lane == 0) {
      if (!ok) {
        for (int i = 0; i < 3; ++i) p6[3 + i] = in.g[3 + i];
        write_result(a, pair, 2, 0, p6, -1, 1);
      } else {
        if (out_mask) out_mask[0] = 0xFu;
        write_result(a, pair, 4 >= prm.min_inliers ? 1 : 3, 4, p6, 0, 1);
      }
    }
    return;
  }
  if (MP == 5 && M == 5) {  // count == modelPoints: one EPnP over all five, no refit, all inliers
    float o[15], im[10];
    for (int i = 0; i < 5; ++i) cr.load_global(i, o + 3 * i, im + 2 * i);
    double R[9], p6[6] = {0, 0, 0, 0, 0, 0};
    cvm::epnp5(heap, o, im, prm.K, R, p6 + 3);
    cvm::rodrigues_mat_to_vec(heap, R, p6);
    if (lane == 0) {
      if (out_mask) out_mask[0] = 0x1Fu;
      write_result(a, pair, 5 >= prm.min_inliers ? 1 : 3, 5, p6, 0, 1);
    }
    return;
  }

  int niters = prm.iterations > 1 ? prm.iterations : 1;
  int best_good = 0, best_h = -1, cur = 0, hk = 0;
  bool best_in_smem = true;    // the mask of the best hypothesis is in s_mask[cur] (else: re-score it from its slot)
  const bool queued = worker >= 0 && q.max_chunks > 0;
  int32_t* wk = queued ? q.wk + 4 * worker : nullptr;
  const int h_off = pre;   // chunk c of this pair covers hypotheses h_off + 32 c ..
  double* stash = (worker >= 0 && q.pre_rt) ? q.pre_rt + (size_t)worker * 12 * 32 : nullptr;
  if (queued) {
    if (lane == 0) {
      st_volatile(wk + 0, pair);
      st_volatile(wk + 1, 0);
      st_volatile(wk + 2, niters);
      __threadfence();
      st_volatile(wk + 3, 1 | (h_off << 1));   // accepting help; helpers need the hypothesis offset of chunk 0
    }
    __syncwarp();
  }
  cr.stage(ws.stage_buf, M);
  if (pre > 0) {
    // ---- the hypotheses solved ahead of time: score them first, in order, like any chunk
    double Rh[9], th[3];
#pragma unroll
    for (int i = 0; i < 9; ++i) Rh[i] = lane < pre ? stash[i * 32 + lane] : 0.0;
#pragma unroll
    for (int i = 0; i < 3; ++i) th[i] = lane < pre ? stash[(9 + i) * 32 + lane] : 0.0;
    for (int k = 0; k < pre; ++k) {
      hk = k;
      if (hk >= niters) break;
      double R[9], t[3];
#pragma unroll
      for (int i = 0; i < 9; ++i) R[i] = __shfl_sync(0xFFFFFFFFu, Rh[i], k);
#pragma unroll
      for (int i = 0; i < 3; ++i) t[i] = __shfl_sync(0xFFFFFFFFu, th[i], k);
      const int good = score_one(a, in, R, t, s_mask + (cur ^ 1) * kMaskWords);
      if (good > max(best_good, MP - 1)) {
        cur ^= 1;
        best_good = good;
        best_h = hk;
        best_in_smem = true;
        niters = cvm::ransac_update_num_iters(prm.confidence, (double)(M - good) / M, MP, niters);
      }
      hk = hk + 1;
    }
    if (queued && lane == 0) st_volatile(wk + 2, niters);
  }
  int c_next = 0;   // chunks below it have been consumed
  while (h_off + c_next * kSolveLanes < niters) {
    int my = c_next;
    if (queued) {
      if (lane == 0) my = atomicAdd(wk + 1, 1);
      my = __shfl_sync(0xFFFFFFFFu, my, 0);
    }
    if (my == c_next) {
      // ---- nobody is helping: solve, then score in RANSAC order with the live early exit
      const int chunk = h_off + my * kSolveLanes;
      double Rh[9], th[3];
      // last chunk with spare lanes: pop the next pair now and let them solve its first hypotheses.  This only happens
      // while the list is not exhausted, i.e. before any helper exists, and this pair needs no further chunk, so the
      // "helped owners are on their last pair" rule of the slots still holds.
      PairIn nx;
      const PairIn* nxp = nullptr;
      const int useful = niters - chunk;
      if (stash && next_pair && *next_pair == -1 && useful < kSolveLanes) {
        int slot = 0;
        if (lane == 0) slot = atomicAdd(&q.ctrl[0], 1);
        slot = __shfl_sync(0xFFFFFFFFu, slot, 0);
        if (slot >= q.ctrl[1]) {
          *next_pair = -2;
        } else {
          *next_pair = q.list[slot];
          nx = load_pair(a, *next_pair, nullptr);
          if (nx.M > MP && prm.iterations >= kSolveLanes) {   // a pair that runs the RANSAC loop (not the 4 / 5 point cases)
            nxp = &nx;
            *next_pre = kSolveLanes - useful;
          }
          if (queued && lane == 0) st_volatile(wk + 3, 0);   // this is the last chunk: no help wanted any more
        }
      }
      solve_chunk<MP>(a, in, ws, chunk, niters, Rh, th, nullptr, nxp, stash);
      K3_T0(t_s);
      for (int k = 0; k < kSolveLanes; ++k) {
        hk = chunk + k;
        if (hk >= niters) break;
        double R[9], t[3];
#pragma unroll
        for (int i = 0; i < 9; ++i) R[i] = __shfl_sync(0xFFFFFFFFu, Rh[i], k);
#pragma unroll
        for (int i = 0; i < 3; ++i) t[i] = __shfl_sync(0xFFFFFFFFu, th[i], k);
        const int good = score_one(a, in, R, t, s_mask + (cur ^ 1) * kMaskWords);
        if (good > max(best_good, MP - 1)) {
          cur ^= 1;
          best_good = good;
          best_h = hk;
          best_in_smem = true;
          niters = cvm::ransac_update_num_iters(prm.confidence, (double)(M - good) / M, MP, niters);
        }
        hk = hk + 1;
      }
      K3_ACC(2, t_s);
#ifdef SEPB_K3_PROFILE
      if (lane == 0) atomicAdd(&g_k3_prof[4], 1ull);
#endif
      c_next = my + 1;
    } else {
      // ---- helpers hold chunks c_next .. my-1: do chunk `my` the way they do, then scan everything in order
      if (h_off + my * kSolveLanes < niters) chunk_to_slot<MP>(a, q, in, ws, worker, my, h_off, s_mask + (cur ^ 1) * kMaskWords);
      for (int cc = c_next; cc <= my && h_off + cc * kSolveLanes < niters; ++cc) {
        const size_t slot = (size_t)worker * q.max_chunks + cc;
        while (ld_volatile_warp(q.slot_ready + slot) == 0) __nanosleep(200);
        __threadfence();
        const int gk = *reinterpret_cast<const volatile int32_t*>(q.slot_good + slot * 32 + lane);
        for (int k = 0; k < kSolveLanes; ++k) {
          hk = h_off + cc * kSolveLanes + k;
          if (hk >= niters) break;
          const int good = __shfl_sync(0xFFFFFFFFu, gk, k);
          if (good > max(best_good, MP - 1)) {
            best_good = good;
            best_h = hk;
            best_in_smem = false;
            niters = cvm::ransac_update_num_iters(prm.confidence, (double)(M - good) / M, MP, niters);
          }
          hk = hk + 1;
        }
      }
      c_next = my + 1;
    }
    if (queued && lane == 0) st_volatile(wk + 2, niters);
  }
  if (queued) {
    __syncwarp();
    if (lane == 0) {
      st_volatile(wk + 3, 0);   // no more chunks wanted
      __threadfence();          // ... and nobody may see this pair's flag once the list cursor has moved on
    }
  }
  __syncwarp();
  K3_T0(t_lm);
  if (best_h < 0) {
    if (lane == 0) write_result(a, pair, 2, 0, in.g, -1, hk);
    return;
  }
  if (!best_in_smem) {   // the winner came out of a slot: its mask is one more scoring pass
    const size_t slot = (size_t)worker * q.max_chunks + (best_h - h_off) / kSolveLanes;
    const int k = (best_h - h_off) % kSolveLanes;
    double R[9], t[3];
#pragma unroll
    for (int i = 0; i < 9; ++i) R[i] = *reinterpret_cast<const volatile double*>(q.slot_rt + (slot * 12 + i) * 32 + k);
#pragma unroll
    for (int i = 0; i < 3; ++i) t[i] = *reinterpret_cast<const volatile double*>(q.slot_rt + (slot * 12 + 9 + i) * 32 + k);
    score_one(a, in, R, t, s_mask + cur * kMaskWords);
    __syncwarp();
  }
  const uint32_t* best = s_mask + cur * kMaskWords;
  if (out_mask)
    for (int w = lane; w < (M + 31) / 32; w += 32) out_mask[w] = best[w];

  double param[6];
  int final_good = best_good;
  if (prm.minimal_model) {
    // OpenCV 3.2 (the reference image's version): solvePnPRansac hands back the winning minimal-set model itself
    double Rh[9], th[3], rv[3];
    solve_chunk<MP>(a, in, ws, best_h, best_h + 1, Rh, th, rv);   // every lane redoes hypothesis best_h
    for (int i = 0; i < 3; ++i) {
      param[i] = rv[i];
      param[3 + i] = th[i];
    }
  } else if (MP == 4) {
    // ---- refit, P3P / AP3P flags: solvePnP(inliers as f64, SOLVEPNP_EPNP) -- EPnP over the whole consensus set (the
    // guess is not used).  Every lane runs the same serial solve on its own scratch; points come through the mask.
    int w = 0;
    uint32_t bits = 0;
    double R[9];
    cvm::epnp_n(heap, best_good, [&](int i, double* o3, double* px) {
      if (i == 0) {
        w = 0;
        bits = best[0];
      }
      while (!bits) bits = best[++w];
      const int m = w * 32 + __ffs(bits) - 1;
      bits &= bits - 1;
      float of[3], i2[2];
      cr.load_global(m, of, i2);
      o3[0] = of[0]; o3[1] = of[1]; o3[2] = of[2];
      px[0] = i2[0]; px[1] = i2[1];
    }, prm.K, R, param + 3);
    cvm::rodrigues_mat_to_vec(heap, R, param);
  } else {
    // ---- refit: solvePnP(inliers, ITERATIVE, useExtrinsicGuess) == CvLevMarq from the caller's guess
    for (int i = 0; i < 6; ++i) param[i] = in.g[i];
    lm_refit(cr, best, M, prm.K, heap, cvm::Heap{ws.heap_base}, param);
    if (prm.refine_iterations > 0 && best_good >= prm.min_inliers) {
      const uint32_t* fin = best;
      final_good = refine_model(cr, M, prm, heap, cvm::Heap{ws.heap_base}, s_mask + cur * kMaskWords,
                                s_mask + (cur ^ 1) * kMaskWords, best_good, param, &fin);
      __syncwarp();
      if (out_mask)
        for (int w = lane; w < (M + 31) / 32; w += 32) out_mask[w] = fin[w];
    }
  }
  K3_ACC(3, t_lm);
  if (lane == 0) write_result(a, pair, final_good >= prm.min_inliers ? 1 : 3, final_good, param, best_h, hk);
}

constexpr size_t kWarpBytes = (size_t)cvm::kEpnpHeap * kSolveLanes * 8 + 2 * kMaskWords * 4;

__device__ __forceinline__ WarpScratch warp_scratch(uint8_t* smem_raw) {
  uint8_t* wsm = smem_raw + (threadIdx.x >> 5) * kWarpBytes;   // this warp's heap + masks
  return WarpScratch{reinterpret_cast<double*>(wsm), reinterpret_cast<float*>(wsm + kStageOffBytes),
                     reinterpret_cast<uint32_t*>(wsm + (size_t)cvm::kEpnpHeap * kSolveLanes * 8)};
}

// reference scheduling (SEPB_PNP_SCHED_PER_PAIR): one one-warp CTA per listed pair, every chunk by that warp
template <int MP>
__global__ void __launch_bounds__(kWarp, 5) pnp_ransac_kernel(PnpArgs a, PnpQueue q) {
  extern __shared__ __align__(16) uint8_t smem_raw[];
  if ((int)blockIdx.x >= q.ctrl[1]) return;
  run_pair<MP>(a, q, -1, q.list[blockIdx.x], warp_scratch(smem_raw));
}

template <int MP>
__global__ void __launch_bounds__(kWarp * kPersistWarps, 1) pnp_persist_kernel(PnpArgs a, PnpQueue q) {
  extern __shared__ __align__(16) uint8_t smem_raw[];
  const WarpScratch ws = warp_scratch(smem_raw);
  const int lane = threadIdx.x & 31;
  const int worker = blockIdx.x * kPersistWarps + (threadIdx.x >> 5);
  const int n_listed = q.ctrl[1];
  // ---- owner: pairs off the list until it is empty
  int nxt = -1, npre = 0;   // pair popped early by the previous run_pair and how many of its hypotheses are solved
  for (;;) {
    int pair, pre = 0;
    if (nxt >= 0) {
      pair = nxt;
      pre = npre;
    } else if (nxt == -2) {
      break;
    } else {
      int slot = 0;
      if (lane == 0) slot = atomicAdd(&q.ctrl[0], 1);
      slot = __shfl_sync(0xFFFFFFFFu, slot, 0);
      if (slot >= n_listed) break;
      pair = q.list[slot];
    }
    nxt = -1;
    npre = 0;
    run_pair<MP>(a, q, worker, pair, ws, pre, &nxt, &npre);
    __syncwarp();
    if (lane == 0) {
      __threadfence();
      atomicAdd(&q.ctrl[2], 1);
    }
  }
  if (q.max_chunks == 0) return;
  // ---- helper: chunks of the owners still running, most iterations left first.  The list is empty for good, so every
  //      owner seen from here on is on its last pair (its worker state is never reset again).
  __threadfence();
  // Claimable work only shrinks once the list is empty (bounds fall, claim counters rise); the one exception is an
  // owner that popped the last pairs a moment ago and has not published yet (a few microseconds).  So a helper that
  // finds nothing looks again after 1, 2, 4, 8 us and then leaves: an idle warp must not take issue slots from the
  // owners on its SM.
  unsigned backoff = 1000;
  for (;;) {
    int best_rem = 0, best_w = -1;
    // a racy look at every worker's state (one 16-byte load each); whatever is picked is re-checked after the claim
#pragma unroll 1
    for (int i = lane; i < q.n_workers; i += 32) {
      int4 w4;   // pair, next chunk, bound, accepting help
      asm volatile("ld.volatile.global.v4.s32 {%0, %1, %2, %3}, [%4];"
                   : "=r"(w4.x), "=r"(w4.y), "=r"(w4.z), "=r"(w4.w)
                   : "l"(q.wk + 4 * i));
      const int rem = w4.w ? w4.z - ((w4.w >> 1) + w4.y * kSolveLanes) : 0;
      if (rem > best_rem) {
        best_rem = rem;
        best_w = i;
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const int r2 = __shfl_xor_sync(0xFFFFFFFFu, best_rem, o), w2 = __shfl_xor_sync(0xFFFFFFFFu, best_w, o);
      if (r2 > best_rem || (r2 == best_rem && w2 > best_w)) {
        best_rem = r2;
        best_w = w2;
      }
    }
    if (best_w < 0) {
      if (ld_volatile_warp(q.ctrl + 2) >= n_listed) break;   // every pair has its result
      if (backoff > 8000u) break;
      __nanosleep(backoff);
      backoff *= 2;
      continue;
    }
    backoff = 1000;
    int cc = 0;
    if (lane == 0) cc = atomicAdd(q.wk + 4 * best_w + 1, 1);
    cc = __shfl_sync(0xFFFFFFFFu, cc, 0);
    // the owner only waits for chunks below its live iteration count (<= the bound read here) while it accepts help
    const int w3 = ld_volatile_warp(q.wk + 4 * best_w + 3);
    const int h_off = w3 >> 1;
    if (cc >= q.max_chunks || !w3 || h_off + cc * kSolveLanes >= ld_volatile_warp(q.wk + 4 * best_w + 2)) continue;
    __threadfence();
    const int pair = ld_volatile_warp(q.wk + 4 * best_w);
    const PairIn in = load_pair(a, pair, ws.stage_buf);
    in.cr.stage(ws.stage_buf, in.M);
    chunk_to_slot<MP>(a, q, in, ws, best_w, cc, h_off, ws.s_mask);
  }
}

}  // namespace

// developer hook: {epnp5, rodrigues, scoring, refit} cycles summed over warps, chunk-solves; zeros unless built with
// SEPB_K3_PROFILE.  reset != 0 clears the counters.
extern "C" SEPB_API int sepb_debug_k3_prof(unsigned long long* out8, int reset) {
#ifdef SEPB_K3_PROFILE
  cudaDeviceSynchronize();
  if (out8) cudaMemcpyFromSymbol(out8, g_k3_prof, sizeof(unsigned long long) * 8);
  if (reset) {
    unsigned long long z[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    cudaMemcpyToSymbol(g_k3_prof, z, sizeof z);
  }
  return 1;
#else
  if (out8) for (int i = 0; i < 8; ++i) out8[i] = 0;
  (void)reset;
  return 0;
#endif
}

int sepb_launch_sample_table(sepb_ctx* ctx, uint16_t* d_table, int max_count, int H, int model_points) {
  const int threads = 128;
  sample_table_kernel<<<(max_count + 1 + threads - 1) / threads, threads, 0, ctx->stream>>>(d_table, max_count, H, model_points);
  ctx->launches++;
  SEPB_CUDA(ctx, cudaGetLastError());
  return SEPB_OK;
}

int sepb_launch_pnp_ransac(sepb_ctx* ctx, FramesView from, FramesView to, const int32_t* d_pairs, int n_pairs,
                           const int32_t* d_n_matches, const uint16_t* d_match_from, const uint16_t* d_match_to,
                           const int32_t* d_n_unique_to, const uint16_t* d_table, int table_h,
                           const int32_t* d_custom_samples, int custom_h, const double* d_guess,
                           const int32_t* d_only_if_ok, const sepb_pnp_params* p, int32_t* d_status, int32_t* d_n_inliers, double* d_pose,
                           float* d_transform, uint32_t* d_inlier_mask, int mask_words, int32_t* d_best_hyp,
                           int32_t* d_iters_run) {
  if (n_pairs == 0) return SEPB_OK;
  PnpArgs a;
  a.from = from;
  a.to = to;
  a.pairs = d_pairs;
  a.n_pairs = n_pairs;
  a.n_matches = d_n_matches;
  a.match_from = d_match_from;
  a.match_to = d_match_to;
  a.n_unique_to = d_n_unique_to;
  a.table = d_table;
  a.custom_samples = d_custom_samples;
  a.custom_h = custom_h;
  a.guess = d_guess;
  a.only_if_ok = d_only_if_ok;
  PnpKernelParams& k = a.prm;
  k.K = cvm::Intrinsics{p->fx, p->fy, p->cx, p->cy};
  k.iterations = p->iterations;
  k.thr2 = (float)(p->reproj_error * p->reproj_error);
  k.confidence = p->confidence;
  k.min_inliers = p->min_inliers;
  for (int i = 0; i < 3; ++i) {
    k.rvec0[i] = p->guess_rvec[i];
    k.tvec0[i] = p->guess_tvec[i];
    k.local_t[i] = p->local_t[i];
  }
  for (int i = 0; i < 9; ++i) k.local_R[i] = p->local_R[i];
  k.table_h = table_h;
  k.minimal_model = p->minimal_model;
  k.refine_iterations = p->refine_iterations;
  k.refine_sigma = (float)p->refine_sigma;
  k.reproj_error = (float)p->reproj_error;
  const bool four = p->flags != SEPB_PNP_ITERATIVE;
  a.status = d_status;
  a.n_inliers = d_n_inliers;
  a.pose = d_pose;
  a.transform = d_transform;
  a.inlier_mask = d_inlier_mask;
  a.mask_words = mask_words;
  a.best_hyp = d_best_hyp;
  a.iters_run = d_iters_run;

  size_t smem = kWarpBytes;
  if (const char* e = sepb_dev_env("SEPB_K3_PAD_SMEM")) smem += (size_t)atoi(e);   // developer switch: fewer resident CTAs per SM
  if (!ctx->cfg_pnp) {
    const int big = (int)(smem + 200000 > 232448 ? 232448 : smem + 200000);
    SEPB_CUDA(ctx, cudaFuncSetAttribute(pnp_ransac_kernel<5>, cudaFuncAttributeMaxDynamicSharedMemorySize, big));
    SEPB_CUDA(ctx, cudaFuncSetAttribute(pnp_ransac_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, big));
    SEPB_CUDA(ctx, cudaFuncSetAttribute(pnp_persist_kernel<5>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(kPersistWarps * kWarpBytes)));
    SEPB_CUDA(ctx, cudaFuncSetAttribute(pnp_persist_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(kPersistWarps * kWarpBytes)));
    ctx->cfg_pnp = true;
  }
  // queue memory (grow-only): zeroed head {ctrl, wk, slot_ready} | list | slot_good | slot_rt
  const bool persistent = ctx->opt_pnp_sched == SEPB_PNP_SCHED_PERSISTENT;
  const int n_workers = persistent ? ctx->sm_count * kPersistWarps : 0;
  const int chunks = (p->iterations + kSolveLanes - 1) / kSolveLanes;
  int max_chunks = persistent && chunks > 1 && chunks <= kMaxHelpChunks ? chunks : 0;
  if (const char* e = sepb_dev_env("SEPB_K3_NO_HELP")) max_chunks = atoi(e) ? 0 : max_chunks;   // developer A/B
  const size_t n_head = 4 + (size_t)4 * n_workers + (size_t)n_workers * max_chunks;
  const size_t n_good = (size_t)n_workers * max_chunks * 32;
  size_t n_ints = n_head + (size_t)n_pairs + n_good;
  n_ints = (n_ints + 3) & ~(size_t)3;
  const bool presolve = persistent && max_chunks > 0 && !sepb_dev_env("SEPB_K3_NO_PRESOLVE");
  const size_t n_slot_rt = (size_t)n_workers * max_chunks * 12 * 32;
  const size_t want = n_ints + (n_slot_rt + (presolve ? (size_t)n_workers * 12 * 32 : 0)) * 2;   // doubles counted as two ints
  if (want > ctx->k3_order_cap) {
    if (ctx->k3_order) {
      cudaStreamSynchronize(ctx->stream);
      cudaFree(ctx->k3_order);
      ctx->k3_order = nullptr;
      ctx->k3_order_cap = 0;
    }
    const size_t cap = want + want / 4;
    SEPB_CUDA(ctx, cudaMalloc(&ctx->k3_order, cap * sizeof(int32_t)));
    ctx->k3_order_cap = cap;
  }
  PnpQueue q;
  q.ctrl = ctx->k3_order;
  q.wk = q.ctrl + 4;
  q.slot_ready = q.wk + (size_t)4 * n_workers;
  q.list = q.ctrl + n_head;
  q.slot_good = q.list + n_pairs;
  q.slot_rt = reinterpret_cast<double*>(ctx->k3_order + n_ints);
  q.pre_rt = presolve ? q.slot_rt + n_slot_rt : nullptr;
  q.n_workers = n_workers;
  q.max_chunks = max_chunks;
  SEPB_CUDA(ctx, cudaMemsetAsync(q.ctrl, 0, n_head * sizeof(int32_t), ctx->stream));
  pnp_prepare_kernel<<<(n_pairs + 255) / 256, 256, 0, ctx->stream>>>(a, q);
  ctx->launches++;
  if (persistent) {
    // fewer CTAs than SMs when the batch cannot feed them all even with every chunk of every pair in flight
    const long long units = (long long)n_pairs * (max_chunks > 0 ? max_chunks : 1);
    const int ctas = (int)std::min<long long>(ctx->sm_count, (units + kPersistWarps - 1) / kPersistWarps);
    q.n_workers = ctas * kPersistWarps;
    if (four) pnp_persist_kernel<4><<<ctas, kWarp * kPersistWarps, kPersistWarps * kWarpBytes, ctx->stream>>>(a, q);
    else pnp_persist_kernel<5><<<ctas, kWarp * kPersistWarps, kPersistWarps * kWarpBytes, ctx->stream>>>(a, q);
  } else {
    if (four) pnp_ransac_kernel<4><<<n_pairs, kWarp, smem, ctx->stream>>>(a, q);
    else pnp_ransac_kernel<5><<<n_pairs, kWarp, smem, ctx->stream>>>(a, q);
  }
  ctx->launches++;
  SEPB_CUDA(ctx, cudaGetLastError());
  return SEPB_OK;
}
