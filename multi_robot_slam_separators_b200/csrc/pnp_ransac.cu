// K3: batched PnP-RANSAC + ITERATIVE refit, one keyframe pair per CTA (one warp).
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
// one, which reproduces the sequential result exactly and keeps most of its early exit.
// Compiled with -fmad=false (see build.py): no multiply-add contraction, like the cv2 build.
#include <cstdlib>

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

namespace {

constexpr int kWarp = 32;
constexpr int kSolveLanes = cvm::kHeapStride;   // hypotheses solved at once by a warp (one lane each)
constexpr int kMaskWords = SEPB_MAX_FEATURES / 32;
#ifndef SEPB_SCORE_ILP
#define SEPB_SCORE_ILP 1   // measured: 1 -> 26.57 ms, 2 -> 26.98 ms, 4 -> 29.8 ms per verify of 50 k pairs (staged LDS are cheap)
#endif
constexpr int kScoreIlp = SEPB_SCORE_ILP;                    // correspondences per lane scored together

struct PnpKernelParams {
  cvm::Intrinsics K;
  int iterations;
  float thr2;          // (float)(reprojError^2)
  double confidence;
  int min_inliers;
  double rvec0[3], tvec0[3];   // guess, camera frame (== (guess * local)^-1)
  double local_R[9], local_t[3];
  int table_h;                 // hypotheses per row of the sample table
};

// getSubset index stream: table[count][h][5]; thread per count value
__global__ void sample_table_kernel(uint16_t* table, int max_count, int H) {
  const int count = blockIdx.x * blockDim.x + threadIdx.x;
  if (count > max_count) return;
  uint16_t* row = table + (size_t)count * H * 5;
  if (count < 5) {
    for (int i = 0; i < H * 5; ++i) row[i] = 0;
    return;
  }
  cvm::Rng rng(0xFFFFFFFFFFFFFFFFULL);
  for (int h = 0; h < H; ++h) {
    int idx[5];
    for (int i = 0; i < 5; ++i) {
      int v;
      bool dup;
      do {
        v = (int)(rng.next() % (unsigned)count);
        dup = false;
        for (int j = 0; j < i; ++j) dup |= idx[j] == v;
      } while (dup);
      idx[i] = v;
      row[h * 5 + i] = (uint16_t)v;
    }
  }
}

// Two-phase launch of K3.  A pair's cost is its number of 32-hypothesis chunks (1..10 at 300 iterations), unknown
// until RANSAC runs, and the hardware hands CTAs out in blockIdx order: in one launch a 7-chunk pair that happens to
// start last keeps an SM busy for 5 ms after the rest has drained (measured: 16 ms against 11.6 ms of work).  So
// phase 0 runs chunk 0 of every pair (uniform cost; pairs that are done after it finish there), saves
// {niters, best count, best hypothesis, mask} of the others, and phase 1 resumes them in DESCENDING order of the
// iteration bound left after chunk 0 (longest first).  Counting sort in one CTA; order inside a bin is arbitrary
// because every pair writes only its own outputs.
constexpr int kOrderThreads = 1024;
constexpr int kStateInts = 4;   // per-pair state header: niters (0 = nothing left to do), best_good, best_h, hk
__global__ void __launch_bounds__(kOrderThreads)
pnp_order_kernel(const int32_t* __restrict__ state, int state_stride, int n_pairs, int32_t* __restrict__ order,
                 const int32_t* __restrict__ n_matches, const int32_t* __restrict__ only_if_ok) {
  __shared__ int hist[SEPB_MAX_FEATURES + 2];   // bin max(0, 4096 - niters); last bin = pairs with nothing left
  __shared__ int warp_tot[32];
  constexpr int kBins = SEPB_MAX_FEATURES + 2;
  const int tid = threadIdx.x;
  for (int i = tid; i < kBins; i += kOrderThreads) hist[i] = 0;
  __syncthreads();
  auto key_of = [&](int p) {
    if (n_matches) {   // before phase 0: pairs that will run RANSAC first (so that they share CTAs), every pair listed
      const bool run = n_matches[p] >= 5 && (!only_if_ok || only_if_ok[p] == 1);
      return run ? 0 : kBins - 1;
    }
    const int it = state[(size_t)p * state_stride];
    return it > 0 ? max(0, SEPB_MAX_FEATURES - it) : kBins - 1;
  };
  for (int p = tid; p < n_pairs; p += kOrderThreads) atomicAdd(&hist[key_of(p)], 1);
  __syncthreads();
  if (tid == 0) order[n_pairs] = n_matches ? n_pairs : n_pairs - hist[kBins - 1];   // pairs listed
  // exclusive scan of the bins: each thread owns a contiguous run of bins
  constexpr int kPer = (kBins + kOrderThreads - 1) / kOrderThreads;
  int local[kPer], sum = 0;
  for (int i = 0; i < kPer; ++i) {
    const int b = tid * kPer + i;
    local[i] = b < kBins ? hist[b] : 0;
    sum += local[i];
  }
  int incl = sum;
  for (int o = 1; o < 32; o <<= 1) {
    const int v = __shfl_up_sync(0xFFFFFFFFu, incl, o);
    if ((tid & 31) >= o) incl += v;
  }
  if ((tid & 31) == 31) warp_tot[tid >> 5] = incl;
  __syncthreads();
  if (tid < 32) {
    int w = warp_tot[tid];
    for (int o = 1; o < 32; o <<= 1) {
      const int v = __shfl_up_sync(0xFFFFFFFFu, w, o);
      if (tid >= o) w += v;
    }
    warp_tot[tid] = w;
  }
  __syncthreads();
  int base = incl - sum + ((tid >> 5) ? warp_tot[(tid >> 5) - 1] : 0);
  __syncthreads();
  for (int i = 0; i < kPer; ++i) {
    const int b = tid * kPer + i;
    if (b < kBins) hist[b] = base;
    base += local[i];
  }
  __syncthreads();
  for (int p = tid; p < n_pairs; p += kOrderThreads) {
    const int k = key_of(p);
    if (k != kBins - 1 || n_matches) order[atomicAdd(&hist[k], 1)] = p;
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

// every lane needs the new parameters, but only the first kSolveLanes lanes own scratch: lane 0's result is broadcast
__device__ __forceinline__ void lm_step_warp(cvm::Heap h, const double* JtJ, const double* JtE, int lambda_lg10,
                                             const double* prev, double* param) {
  if ((threadIdx.x & 31) < kSolveLanes) lm_step(h, JtJ, JtE, lambda_lg10, prev, param);
  __syncwarp();
#pragma unroll
  for (int i = 0; i < 6; ++i) param[i] = __shfl_sync(0xFFFFFFFFu, param[i], 0);
}

// One warp per keyframe pair.  WPC warps (pairs) share a CTA: the warps of a CTA start together and, as long as their
// work is alike (phase 0: chunk 0 of every pair), run the same code at about the same time, which is what the
// instruction caches want -- ncu shows the GPC-level cache at 40-50 % of its request rate for this 330 KB kernel.
template <int WPC>
__global__ void __launch_bounds__(kWarp * WPC, 5 / WPC)
pnp_ransac_kernel(FramesView from, FramesView to, const int32_t* __restrict__ pairs, int n_pairs,
                  const int32_t* __restrict__ n_matches, const uint16_t* __restrict__ match_from,
                  const uint16_t* __restrict__ match_to, const int32_t* __restrict__ n_unique_to,
                  const uint16_t* __restrict__ table, const int32_t* __restrict__ custom_samples, int custom_h,
                  const double* __restrict__ guess /*[P][6] or null*/, const int32_t* __restrict__ only_if_ok,
                  PnpKernelParams prm,
                  int32_t* __restrict__ status, int32_t* __restrict__ n_inliers, double* __restrict__ pose /*[P][6]*/,
                  float* __restrict__ transform /*[P][12]*/, uint32_t* __restrict__ inlier_mask, int mask_words,
                  int32_t* __restrict__ best_hyp, int32_t* __restrict__ iters_run, const int32_t* __restrict__ order,
                  int phase /*-1: whole run, 0: chunk 0 (+ finish if done), 1: resume*/, int32_t* __restrict__ state,
                  int state_stride) {
  extern __shared__ __align__(16) uint8_t smem_raw[];
  constexpr size_t kWarpBytes = (size_t)cvm::kEpnpHeap * kSolveLanes * 8 + 2 * kMaskWords * 4;
  uint8_t* wsm = smem_raw + (threadIdx.x >> 5) * kWarpBytes;   // this warp's heap + masks
  double* heap_base = reinterpret_cast<double*>(wsm);
  uint32_t* s_mask = reinterpret_cast<uint32_t*>(wsm + (size_t)cvm::kEpnpHeap * kSolveLanes * 8);  // 2 * kMaskWords
  const int lane = threadIdx.x & 31;
  const int slot = blockIdx.x * WPC + (threadIdx.x >> 5);
  if (slot >= (order ? order[n_pairs] : n_pairs)) return;
  const int pair = order ? order[slot] : slot;
  int32_t* st = state ? state + (size_t)pair * state_stride : nullptr;
  if (phase == 0 && lane == 0) st[0] = 0;   // nothing to resume unless the chunk loop says so
  // second pass (estimate_transformation's call with guess = first result): pairs whose first pass failed keep it
  if (only_if_ok && only_if_ok[pair] != 1) return;
  const cvm::Heap heap{heap_base + lane};
  int fi = pairs[2 * pair], ti = pairs[2 * pair + 1];
  const bool valid_pair = fi >= 0 && fi < from.size && ti >= 0 && ti < to.size;
  if (!valid_pair) fi = ti = 0;
  const int M = valid_pair ? n_matches[pair] : 0;
  const int nf = valid_pair ? from.count[fi] : 0;
  const int nu = n_unique_to ? n_unique_to[pair] : M;
  float* stage_buf = reinterpret_cast<float*>(wsm + kStageOffBytes);
  Corr cr{from.kp3d + (size_t)fi * from.max_feat * 3, to.kp2d + (size_t)ti * to.max_feat * 2,
          match_from + (size_t)pair * from.max_feat, match_to + (size_t)pair * from.max_feat,
          M <= kStageCap ? stage_buf : nullptr};
  double g[6];
  for (int i = 0; i < 3; ++i) {
    g[i] = guess ? guess[6 * pair + i] : prm.rvec0[i];
    g[3 + i] = guess ? guess[6 * pair + 3 + i] : prm.tvec0[i];
  }
  uint32_t* out_mask = inlier_mask ? inlier_mask + (size_t)pair * mask_words : nullptr;
  if (out_mask && phase != 1)
    for (int w = lane; w < mask_words; w += 32) out_mask[w] = 0;

  auto finish = [&](int st, int ninl, const double* p6, int bh, int it) {
    if (lane == 0) {
      status[pair] = st;
      n_inliers[pair] = ninl;
      if (best_hyp) best_hyp[pair] = bh;
      if (iters_run) iters_run[pair] = it;
      for (int i = 0; i < 6; ++i) pose[6 * pair + i] = p6[i];
      float* T = transform + 12 * (size_t)pair;
      if (st == 1) {
        // T = (local * [R|t])^-1  (maps To-base coordinates into the From base frame)
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
  };

  // myRegistrationVis.cpp:1069-1070 and estimateMotion3DTo2D: need >= minInliers on every side
  if (M < prm.min_inliers || nf < prm.min_inliers || nu < prm.min_inliers || M < 5) {
    finish(0, 0, g, -1, 0);
    return;
  }

  if (M == 5) {  // count == modelPoints: one EPnP over all five, no refit, all inliers
    float o[15], im[10];
    for (int i = 0; i < 5; ++i) cr.load_global(i, o + 3 * i, im + 2 * i);
    double R[9], p6[6] = {0, 0, 0, 0, 0, 0};
    if (lane < kSolveLanes) {   // only these lanes own a slice of the scratch heap; lane 0 reports
      cvm::epnp5(heap, o, im, prm.K, R, p6 + 3);
      cvm::rodrigues_mat_to_vec(heap, R, p6);
    }
    if (out_mask && lane == 0) out_mask[0] = 0x1Fu;
    finish(5 >= prm.min_inliers ? 1 : 3, 5, p6, 0, 1);
    return;
  }

  int niters = prm.iterations > 1 ? prm.iterations : 1;
  int best_good = 0, best_h = -1, cur = 0, hk = 0, chunk0 = 0;
  if (phase == 1) {   // resume after chunk 0
    niters = st[0];
    best_good = st[1];
    best_h = st[2];
    hk = st[3];
    chunk0 = kSolveLanes;
    for (int w = lane; w < (M + 31) / 32; w += 32) s_mask[w] = (uint32_t)st[kStateInts + w];
    __syncwarp();
  }
  cr.stage(stage_buf, M);
  for (int chunk = chunk0; chunk < niters; chunk += kSolveLanes) {
    // ---- solve: lane <-> hypothesis (the first kSolveLanes lanes; more resident warps hide the f64 latency
    //      better than wider warps, and shorter chunks waste fewer hypotheses past the adaptive stop)
    int h = chunk + lane;
    if (h >= niters) h = niters - 1;
    double Rh[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0}, th[3] = {0, 0, 0};
    if (lane < kSolveLanes) {
      float o[15], im[10];
      for (int i = 0; i < 5; ++i) {
        const int s = custom_samples ? custom_samples[((size_t)pair * custom_h + h) * 5 + i]
                                     : table[((size_t)M * prm.table_h + h) * 5 + i];
        cr.load(min(max(s, 0), M - 1), o + 3 * i, im + 2 * i);
      }
      __syncwarp();   // every lane has read its samples from the staging area before a solve overwrites it
      double R0[9], rv[3];
      K3_T0(t_e);
      cvm::epnp5(heap, o, im, prm.K, R0, th);
      K3_ACC(0, t_e);
      K3_T0(t_r);
      cvm::rodrigues_mat_to_vec(heap, R0, rv);   // runKernel hands the model on as rvec
      cvm::rodrigues_vec_to_mat(rv, Rh, nullptr);  // projectPoints turns it back into a matrix
      K3_ACC(1, t_r);
    }
    __syncwarp();
    cr.stage(stage_buf, M);   // the solve overwrote the staging area
    K3_T0(t_s);
    // ---- score: warp <-> hypothesis, in RANSAC order
    for (int k = 0; k < kSolveLanes; ++k) {
      hk = chunk + k;
      if (hk >= niters) break;
      double R[9], t[3];
#pragma unroll
      for (int i = 0; i < 9; ++i) R[i] = __shfl_sync(0xFFFFFFFFu, Rh[i], k);
#pragma unroll
      for (int i = 0; i < 3; ++i) t[i] = __shfl_sync(0xFFFFFFFFu, th[i], k);
      uint32_t* cand = s_mask + (cur ^ 1) * kMaskWords;
      int good = 0;
      // kScoreIlp independent correspondences per lane in flight: the projection is a long dependent f64 chain
      // behind two dependent loads, and a warp has only itself to hide that latency
      for (int base = 0; base < M; base += 32 * kScoreIlp) {
        bool inl[kScoreIlp];
#pragma unroll
        for (int u = 0; u < kScoreIlp; ++u) {
          const int m = base + 32 * u + lane;
          float o3[3], i2[2];
          cr.load(min(m, M - 1), o3, i2);
          inl[u] = (cvm::reproj_err_f32(R, t, prm.K, o3, i2) <= prm.thr2) && m < M;
        }
#pragma unroll
        for (int u = 0; u < kScoreIlp; ++u) {
          const uint32_t b = __ballot_sync(0xFFFFFFFFu, inl[u]);
          if (base + 32 * u < M) {
            if (lane == 0) cand[(base >> 5) + u] = b;
            good += __popc(b);
          }
        }
      }
      if (good > max(best_good, 4)) {
        cur ^= 1;
        best_good = good;
        best_h = hk;
        niters = cvm::ransac_update_num_iters(prm.confidence, (double)(M - good) / M, 5, niters);
      }
      hk = hk + 1;
    }
    K3_ACC(2, t_s);
#ifdef SEPB_K3_PROFILE
    if (lane == 0) atomicAdd(&g_k3_prof[4], 1ull);
#endif
    if (phase == 0 && chunk + kSolveLanes < niters) {   // more chunks to run: hand the pair over to phase 1
      __syncwarp();
      if (lane == 0) {
        st[0] = niters;
        st[1] = best_good;
        st[2] = best_h;
        st[3] = hk;
      }
      const uint32_t* bm = s_mask + cur * kMaskWords;
      for (int w = lane; w < (M + 31) / 32; w += 32) st[kStateInts + w] = (int32_t)bm[w];
      return;
    }
  }
  __syncwarp();
  K3_T0(t_lm);
  if (best_h < 0) {
    finish(2, 0, g, -1, hk);
    return;
  }
  const uint32_t* best = s_mask + cur * kMaskWords;
  if (out_mask)
    for (int w = lane; w < (M + 31) / 32; w += 32) out_mask[w] = best[w];

  // ---- refit: solvePnP(inliers, ITERATIVE, useExtrinsicGuess) == CvLevMarq from the caller's guess
  double param[6], prev[6], JtJ[21], JtE[6];
  for (int i = 0; i < 6; ++i) param[i] = prev[i] = g[i];
  int lambda_lg10 = -3, iters = 0;
  double err2 = 0, prev_err_norm = 0;
  const int max_iter = 20;
  const double eps = 1.1920928955078125e-07;  // FLT_EPSILON
  lm_accumulate(cr, best, M, prm.K, param, true, JtJ, JtE, err2);  // STARTED -> CALC_J
  for (;;) {
    // CALC_J
    for (int i = 0; i < 6; ++i) prev[i] = param[i];
    lm_step_warp(heap, JtJ, JtE, lambda_lg10, prev, param);
    if (iters == 0) prev_err_norm = sqrt(err2);
    double dJ[21], dE[6];
    lm_accumulate(cr, best, M, prm.K, param, false, dJ, dE, err2);
    // CHECK_ERR
    double err_norm = sqrt(err2);
    while (err_norm > prev_err_norm && ++lambda_lg10 <= 16) {
      lm_step_warp(heap, JtJ, JtE, lambda_lg10, prev, param);
      lm_accumulate(cr, best, M, prm.K, param, false, dJ, dE, err2);
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
    lm_accumulate(cr, best, M, prm.K, param, true, JtJ, JtE, err2);
  }
  K3_ACC(3, t_lm);
  finish(best_good >= prm.min_inliers ? 1 : 3, best_good, param, best_h, hk);
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

int sepb_launch_sample_table(sepb_ctx* ctx, uint16_t* d_table, int max_count, int H) {
  const int threads = 128;
  sample_table_kernel<<<(max_count + 1 + threads - 1) / threads, threads, 0, ctx->stream>>>(d_table, max_count, H);
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
  PnpKernelParams k;
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
  size_t smem = (size_t)cvm::kEpnpHeap * kSolveLanes * 8 + 2 * kMaskWords * 4;
  if (const char* e = sepb_dev_env("SEPB_K3_PAD_SMEM")) smem += (size_t)atoi(e);   // developer switch: fewer resident CTAs per SM
  bool& configured = ctx->cfg_pnp;
  if (!configured) {
    SEPB_CUDA(ctx, cudaFuncSetAttribute(pnp_ransac_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(smem + 200000 > 232448 ? 232448 : smem + 200000)));
    SEPB_CUDA(ctx, cudaFuncSetAttribute(pnp_ransac_kernel<5>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(5 * smem)));
    configured = true;
  }
  int wpc0 = 5;   // warps (pairs) per CTA in phase 0; the other launches use one-warp CTAs (variable work per pair)
  if (const char* e = sepb_dev_env("SEPB_K3_WPC")) wpc0 = atoi(e);
  int wpc1 = 1;
  if (const char* e = sepb_dev_env("SEPB_K3_WPC1")) wpc1 = atoi(e);
  auto launch = [&](int phase, const int32_t* d_order, int32_t* d_state, int stride) {
#define SEPB_K3_LAUNCH(W)                                                                                              \
  pnp_ransac_kernel<W><<<(n_pairs + W - 1) / W, kWarp * W, smem * W, ctx->stream>>>(                                   \
      from, to, d_pairs, n_pairs, d_n_matches, d_match_from, d_match_to, d_n_unique_to, d_table, d_custom_samples,     \
      custom_h, d_guess, d_only_if_ok, k, d_status, d_n_inliers, d_pose, d_transform, d_inlier_mask, mask_words,      \
      d_best_hyp, d_iters_run, d_order, phase, d_state, stride)
    const int w = phase == 0 ? wpc0 : phase == 1 ? wpc1 : 1;
    if (w == 5) {
      SEPB_K3_LAUNCH(5);
    } else {
      SEPB_K3_LAUNCH(1);
    }
#undef SEPB_K3_LAUNCH
    ctx->launches++;
  };
  // below a few waves of CTAs there is no tail to trim: one launch.  Otherwise chunk 0, sort, resume (pnp_order_kernel)
  const int two_phase_min = ctx->opt_two_phase_min < 0 ? 4 * ctx->sm_count : ctx->opt_two_phase_min;
  if (n_pairs <= two_phase_min || p->iterations <= kSolveLanes) {
    launch(-1, nullptr, nullptr, 0);
    SEPB_CUDA(ctx, cudaGetLastError());
    return SEPB_OK;
  }
  const int stride = kStateInts + (from.max_feat + 31) / 32;
  const size_t want = (size_t)n_pairs * (stride + 1) + 1;   // state + order + resume count
  if (want > ctx->k3_order_cap) {
    if (ctx->k3_order) {
      cudaStreamSynchronize(ctx->stream);
      cudaFree(ctx->k3_order);
      ctx->k3_order = nullptr;
      ctx->k3_order_cap = 0;
    }
    const size_t cap = want + want / 2;
    SEPB_CUDA(ctx, cudaMalloc(&ctx->k3_order, cap * sizeof(int32_t)));
    ctx->k3_order_cap = cap;
  }
  int32_t* d_state = ctx->k3_order;
  int32_t* d_order = ctx->k3_order + (size_t)n_pairs * stride;
  if (wpc0 > 1) {   // group the pairs that will run RANSAC into the same CTAs
    pnp_order_kernel<<<1, kOrderThreads, 0, ctx->stream>>>(nullptr, 0, n_pairs, d_order, d_n_matches, d_only_if_ok);
    ctx->launches++;
    launch(0, d_order, d_state, stride);
  } else {
    launch(0, nullptr, d_state, stride);
  }
  pnp_order_kernel<<<1, kOrderThreads, 0, ctx->stream>>>(d_state, stride, n_pairs, d_order, nullptr, nullptr);
  ctx->launches++;
  launch(1, d_order, d_state, stride);
  SEPB_CUDA(ctx, cudaGetLastError());
  return SEPB_OK;
}
