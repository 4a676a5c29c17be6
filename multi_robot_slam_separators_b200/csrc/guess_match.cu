// K2G: pass-2 correspondences -- matching inside a window around the projected guess.
//
// Replaces the guess-set branch of RegistrationVis::computeTransformationFromFeatsImpl
// (R/src/myRegistrationVis.cpp:477-825, default sub-branch :667-818) that
// StereoCamGeometricTools::estimateTransformation reaches on its SECOND call with guess = result of the
// first (R/src/stereoCamGeometricTools.cpp:153-164):
//   * guessCameraRef = (guess * localTransform)^-1, rvec = Rodrigues(R)                       (:486-493)
//   * cv::projectPoints of every From 3-D point, kept iff inside the image and z > 0            (:494-512)
//   * To keypoints within Vis/CorGuessWinSize pixels of the projection (the reference asks FLANN's approximate
//     randomised kd-tree, :672-680; this is the exact search, squared distance dx*dx + dy*dy in f32)
//   * >= 2 in the window: keep the ones with the From feature's octave; >= 2 left -> 2-NN Hamming among them
//     (ties -> nearer to the projection, then lower index: the order of a sorted result set), accept iff
//     d1 < nndr * d2 (strict, :744); exactly 1 left -> accept.  1 in the window -> accept iff octaves agree.
//   * a To feature goes to the first From feature (ascending index) that claims it            (:776-787)
// Output: the same correspondence list K2 emits (ascending From index), plus the guess (rvec, tvec) K3 is
// seeded with.  One CTA per pair; pairs whose first pass failed (status != 1) are skipped untouched:
// the reference then runs the global matcher again with a null guess and gets the first result back.
// Compiled with -fmad=false (projection arithmetic is replayed like pnp_geom.cuh's).
#include "common.cuh"
#include "pnp_geom.cuh"

namespace {

constexpr int kThreadsG = 256;

struct GuessParams {
  cvm::Intrinsics K;
  double local_R[9], local_t[3];
  float width_m1, height_m1;   // uIsInBounds(x, 0, width - 1): 0 <= x < width - 1
  float r2;                    // window radius squared (f32)
  float nndr;
};

__device__ __forceinline__ int octave_of(int o) {
  o &= 255;
  return o < 128 ? o : (-128 | o);
}

__global__ void __launch_bounds__(kThreadsG)
guess_match_kernel(FramesView from, FramesView to, const int32_t* __restrict__ pairs, int n_pairs,
                   const int32_t* __restrict__ status1, const float* __restrict__ transform1 /*[P][12]*/, GuessParams prm,
                   int32_t* __restrict__ n_matches, uint16_t* __restrict__ match_from, uint16_t* __restrict__ match_to,
                   int32_t* __restrict__ n_unique_to, double* __restrict__ guess_out /*[P][6]*/) {
  extern __shared__ __align__(16) uint8_t smem_raw[];
  const int Ft = to.max_feat;
  uint4* s_desc = reinterpret_cast<uint4*>(smem_raw);                   // [Ft][2] To descriptors
  float2* s_kp = reinterpret_cast<float2*>(s_desc + 2 * Ft);            // [Ft] To keypoints
  int32_t* s_owner = reinterpret_cast<int32_t*>(s_kp + Ft);             // [Ft] smallest From index claiming the To feature
  int8_t* s_oct = reinterpret_cast<int8_t*>(s_owner + Ft);              // [Ft]
  __shared__ double s_heap[21 * cvm::kHeapStride];
  __shared__ double s_R[9], s_t[3];
  __shared__ int s_scan[kThreadsG / 32];

  const int pair = blockIdx.x, tid = threadIdx.x;
  if (pair >= n_pairs) return;
  if (status1 && status1[pair] != 1) return;   // first pass failed: its (null) result stands
  const int fi = pairs[2 * pair], ti = pairs[2 * pair + 1];
  if (fi < 0 || fi >= from.size || ti < 0 || ti >= to.size) {
    if (tid == 0) {
      n_matches[pair] = 0;
      if (n_unique_to) n_unique_to[pair] = 0;
    }
    return;
  }
  const int nf = min(from.count[fi], from.max_feat), nt = min(to.count[ti], to.max_feat);

  if (tid == 0) {
    // guessCameraRef = (guess * local)^-1 on f32-rounded entries (rtabmap::Transform is f32), as sepb_pnp_set_guess
    const float* G = transform1 + 12 * (size_t)pair;
    double RP[9], tP[3], Rinv[9], g[6];
    for (int i = 0; i < 3; ++i) {
      for (int j = 0; j < 3; ++j)
        RP[3 * i + j] = (double)(float)((double)G[4 * i] * prm.local_R[j] + (double)G[4 * i + 1] * prm.local_R[3 + j] +
                                        (double)G[4 * i + 2] * prm.local_R[6 + j]);
      tP[i] = (double)(float)((double)G[4 * i] * prm.local_t[0] + (double)G[4 * i + 1] * prm.local_t[1] +
                              (double)G[4 * i + 2] * prm.local_t[2] + (double)G[4 * i + 3]);
    }
    for (int i = 0; i < 3; ++i)
      for (int j = 0; j < 3; ++j) Rinv[3 * i + j] = RP[3 * j + i];
    for (int i = 0; i < 3; ++i)
      g[3 + i] = (double)(float)(-(Rinv[3 * i] * tP[0] + Rinv[3 * i + 1] * tP[1] + Rinv[3 * i + 2] * tP[2]));
    cvm::rodrigues_mat_to_vec(cvm::Heap{s_heap}, Rinv, g);
    double R[9];
    cvm::rodrigues_vec_to_mat(g, R, nullptr);   // cv::projectPoints turns rvec back into a matrix
    for (int i = 0; i < 9; ++i) s_R[i] = R[i];
    for (int i = 0; i < 3; ++i) s_t[i] = g[3 + i];
    if (guess_out)
      for (int i = 0; i < 6; ++i) guess_out[6 * (size_t)pair + i] = g[i];
  }
  // stage the To frame
  const float2* g_kp = reinterpret_cast<const float2*>(to.kp2d + (size_t)ti * to.max_feat * 2);
  const uint4* g_desc = reinterpret_cast<const uint4*>(to.desc + (size_t)ti * to.max_feat * 32);
  const int32_t* g_oct = to.octave + (size_t)ti * to.max_feat;
  for (int j = tid; j < nt; j += kThreadsG) {
    s_kp[j] = g_kp[j];
    s_desc[2 * j] = g_desc[2 * j];
    s_desc[2 * j + 1] = g_desc[2 * j + 1];
    s_owner[j] = INT_MAX;
    s_oct[j] = (int8_t)octave_of(g_oct[j]);
  }
  __syncthreads();

  const float* p3 = from.kp3d + (size_t)fi * from.max_feat * 3;
  const uint4* f_desc = reinterpret_cast<const uint4*>(from.desc + (size_t)fi * from.max_feat * 32);
  const int32_t* f_oct = from.octave + (size_t)fi * from.max_feat;
  // chunked ownership of From features (ascending, <= 32 per thread) so that the compaction below is a plain scan
  const int chunk = (nf + kThreadsG - 1) / kThreadsG;
  const int lo = min(nf, tid * chunk), hi = min(nf, lo + chunk);
  uint32_t cand_mask = 0;           // From features of my chunk that claimed a To feature
  int16_t my_match[32];             // chunk <= 32 (max_feat <= 4096... enforced by the launcher: chunk <= 32)
#pragma unroll 1
  for (int i = lo; i < hi; ++i) {
    my_match[i - lo] = -1;
    const float X = p3[3 * i], Y = p3[3 * i + 1], Z = p3[3 * i + 2];
    if (!(isfinite(X) && isfinite(Y) && isfinite(Z))) continue;
    double u, v;
    cvm::project_point(s_R, s_t, prm.K, (double)X, (double)Y, (double)Z, u, v);
    const double zc = s_R[6] * (double)X + s_R[7] * (double)Y + s_R[8] * (double)Z + s_t[2];
    const float pu = (float)u, pv = (float)v;
    if (!(pu >= 0.f && pu < prm.width_m1 && pv >= 0.f && pv < prm.height_m1 && zc > 0.0)) continue;
    const uint4 da = f_desc[2 * i], db = f_desc[2 * i + 1];
    const int of = octave_of(f_oct[i]);
    int n_f = 0;
    unsigned long long k1 = ~0ull, k2 = ~0ull;   // (hamming << 48) | (d2 bits << 16) | To index
#pragma unroll 4
    for (int j = 0; j < nt; ++j) {
      const float2 q = s_kp[j];
      const float dx = pu - q.x, dy = pv - q.y;
      const float d2 = __fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy));
      if (d2 < prm.r2 && (int)s_oct[j] == of) {
        const uint4 ta = s_desc[2 * j], tb = s_desc[2 * j + 1];
        const int ham = __popc(da.x ^ ta.x) + __popc(da.y ^ ta.y) + __popc(da.z ^ ta.z) + __popc(da.w ^ ta.w) +
                        __popc(db.x ^ tb.x) + __popc(db.y ^ tb.y) + __popc(db.z ^ tb.z) + __popc(db.w ^ tb.w);
        const unsigned long long key = ((unsigned long long)ham << 48) | ((unsigned long long)__float_as_uint(d2) << 16) |
                                       (unsigned long long)j;
        if (key < k1) {
          k2 = k1;
          k1 = key;
        } else if (key < k2) {
          k2 = key;
        }
        ++n_f;
      }
    }
    // n_f == 1: accepted whether the window held one feature (:756-764) or several with one octave match (:751-754)
    int m = -1;
    if (n_f == 1) {
      m = (int)(k1 & 0xFFFFu);
    } else if (n_f >= 2) {
      const float d1 = (float)(int)(k1 >> 48), d2v = (float)(int)(k2 >> 48);
      if (d1 < __fmul_rn(prm.nndr, d2v)) m = (int)(k1 & 0xFFFFu);
    }
    if (m >= 0) {
      my_match[i - lo] = (int16_t)m;
      cand_mask |= 1u << (i - lo);
      atomicMin(&s_owner[m], i);
    }
  }
  __syncthreads();
  // keep (i, m) iff i is the first claimant; ascending compaction
  int local = 0;
  uint32_t keep = 0;
  for (int i = lo; i < hi; ++i)
    if ((cand_mask >> (i - lo)) & 1u)
      if (s_owner[my_match[i - lo]] == i) {
        keep |= 1u << (i - lo);
        ++local;
      }
  const int lane = tid & 31, warp = tid >> 5;
  int incl = local;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int v = __shfl_up_sync(0xFFFFFFFFu, incl, o);
    if (lane >= o) incl += v;
  }
  if (lane == 31) s_scan[warp] = incl;
  __syncthreads();
  int base = 0, total = 0;
#pragma unroll
  for (int w = 0; w < kThreadsG / 32; ++w) {
    const int v = s_scan[w];
    if (w < warp) base += v;
    total += v;
  }
  int pos = base + incl - local;
  const size_t obase = (size_t)pair * from.max_feat;
  for (int i = lo; i < hi; ++i)
    if ((keep >> (i - lo)) & 1u) {
      match_from[obase + pos] = (uint16_t)i;
      match_to[obase + pos] = (uint16_t)my_match[i - lo];
      ++pos;
    }
  if (tid == 0) {
    n_matches[pair] = total;
    if (n_unique_to) n_unique_to[pair] = nt;   // wordsTo keeps every To feature (matched ids + fresh ids, :803-817)
  }
}

}  // namespace

int sepb_launch_guess_match(sepb_ctx* ctx, FramesView from, FramesView to, const int32_t* d_pairs, int n_pairs,
                            const int32_t* d_status1, const float* d_transform1, const sepb_pnp_params* p, int width,
                            int height, float win, float nndr, int32_t* d_n_matches, uint16_t* d_match_from,
                            uint16_t* d_match_to, int32_t* d_n_unique_to, double* d_guess) {
  if (n_pairs == 0) return SEPB_OK;
  if (from.max_feat > kThreadsG * 32 || to.max_feat > 65535)
    return sepb_fail(ctx, SEPB_E_SHAPE, "guess_match: frame too large", __FILE__, __LINE__);
  GuessParams g;
  g.K = cvm::Intrinsics{p->fx, p->fy, p->cx, p->cy};
  for (int i = 0; i < 9; ++i) g.local_R[i] = p->local_R[i];
  for (int i = 0; i < 3; ++i) g.local_t[i] = p->local_t[i];
  g.width_m1 = (float)(width - 1);
  g.height_m1 = (float)(height - 1);
  g.r2 = win * win;
  g.nndr = nndr;
  const size_t smem = (size_t)to.max_feat * (8 + 32 + 4 + 1) + 16;
  if (smem > ctx->cfg_guess_smem) {
    SEPB_CUDA(ctx, cudaFuncSetAttribute(guess_match_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    ctx->cfg_guess_smem = smem;
  }
  guess_match_kernel<<<n_pairs, kThreadsG, smem, ctx->stream>>>(from, to, d_pairs, n_pairs, d_status1, d_transform1, g,
                                                              d_n_matches, d_match_from, d_match_to, d_n_unique_to, d_guess);
  ctx->launches++;
  SEPB_CUDA(ctx, cudaGetLastError());
  return SEPB_OK;
}
