// K2: brute-force 256-bit Hamming 2-NN + NNDR + one-to-one filter + PnP correspondence list.
//
// Replaces R/src/myRegistrationVis.cpp:839-894 (VWDictionary::addNewWords(descriptorsTo, 2) in
// brute-force mode == cv::BFMatcher(NORM_HAMMING).knnMatch(k=2), NNDR, uniqueness filter) and the
// correspondence gathering of util3d::estimateMotion3DTo2D.
//
// One CTA per keyframe pair.  The From ("train") descriptors are staged tile by tile in shared
// memory and read back as 128-bit broadcasts; every thread keeps QPT To ("query") descriptors in
// registers (8 x u32 each) and folds each distance into a packed (dist << 16 | trainIdx) key, so a
// running top-2 is two integer min/max operations and ties resolve to the lowest train index exactly
// like BFMatcher.  Integer work only: bit-exact by construction.
//
// Roofline: 8 POPC per distance (16 lanes/clk/SM) -> INT/XU-pipe bound; algorithmic HBM bytes per
// pair = (F_to + F_from) * 32 + F_from * 12 (descriptors once, From 3-D points once).
#include "common.cuh"
#include "match_tail.cuh"

namespace {

constexpr int kThreads = 256;
constexpr int kQPT = 4;                      // queries per thread
constexpr int kQTile = kThreads * kQPT;      // 1024 queries per pass
constexpr int kTTile = 1024;                 // train descriptors staged per pass (32 KB)

__device__ __forceinline__ uint32_t ham256(const uint32_t (&q)[8], const uint4& a, const uint4& b) {
  uint32_t d0 = __popc(q[0] ^ a.x) + __popc(q[1] ^ a.y);
  uint32_t d1 = __popc(q[2] ^ a.z) + __popc(q[3] ^ a.w);
  uint32_t d2 = __popc(q[4] ^ b.x) + __popc(q[5] ^ b.y);
  uint32_t d3 = __popc(q[6] ^ b.z) + __popc(q[7] ^ b.w);
  return (d0 + d1) + (d2 + d3);
}

__global__ void __launch_bounds__(kThreads)
hamming_match_kernel(FramesView from, FramesView to, const int32_t* __restrict__ pairs, int n_pairs, float nndr,
                     int strict, int32_t* __restrict__ knn_idx, int32_t* __restrict__ knn_dist,
                     int32_t* __restrict__ n_matches, uint16_t* __restrict__ match_from,
                     uint16_t* __restrict__ match_to, int32_t* __restrict__ n_unique_to) {
  extern __shared__ __align__(16) uint8_t smem[];
  uint4* s_train = reinterpret_cast<uint4*>(smem);                                   // kTTile * 2 uint4
  int32_t* s_assigned = reinterpret_cast<int32_t*>(smem + kTTile * 32);              // [to.max_feat]
  uint32_t* s_count = reinterpret_cast<uint32_t*>(s_assigned + to.max_feat);         // [from.max_feat]
  uint16_t* s_owner = reinterpret_cast<uint16_t*>(s_count + from.max_feat);          // [from.max_feat]
  __shared__ int s_scan[kThreads / 32];
  __shared__ int s_dropped;

  const int pair = blockIdx.x;
  if (pair >= n_pairs) return;
  const int fi = pairs[2 * pair], ti = pairs[2 * pair + 1];
  const int tid = threadIdx.x;
  const bool valid_pair = fi >= 0 && fi < from.size && ti >= 0 && ti < to.size;
  const int nf = valid_pair ? min(from.count[fi], from.max_feat) : 0;
  const int nt = valid_pair ? min(to.count[ti], to.max_feat) : 0;
  const uint4* g_train = reinterpret_cast<const uint4*>(from.desc + (size_t)(valid_pair ? fi : 0) * from.max_feat * 32);
  const uint4* g_query = reinterpret_cast<const uint4*>(to.desc + (size_t)(valid_pair ? ti : 0) * to.max_feat * 32);

  for (int i = tid; i < nf; i += kThreads) s_count[i] = 0;
  if (tid == 0) s_dropped = 0;

  for (int qbase = 0; qbase < nt; qbase += kQTile) {
    uint32_t q[kQPT][8];
    uint32_t m1[kQPT], m2[kQPT];
#pragma unroll
    for (int r = 0; r < kQPT; ++r) {
      const int qi = qbase + r * kThreads + tid;
      uint4 a = make_uint4(0, 0, 0, 0), b = a;
      if (qi < nt) {
        a = __ldg(g_query + 2 * qi);
        b = __ldg(g_query + 2 * qi + 1);
      }
      q[r][0] = a.x; q[r][1] = a.y; q[r][2] = a.z; q[r][3] = a.w;
      q[r][4] = b.x; q[r][5] = b.y; q[r][6] = b.z; q[r][7] = b.w;
      m1[r] = 0xFFFFFFFFu;
      m2[r] = 0xFFFFFFFFu;
    }
    for (int tbase = 0; tbase < nf; tbase += kTTile) {
      const int tn = min(kTTile, nf - tbase);
      __syncthreads();  // previous tile fully consumed
      for (int i = tid; i < 2 * tn; i += kThreads) s_train[i] = __ldg(g_train + 2 * tbase + i);
      __syncthreads();
#pragma unroll 4
      for (int j = 0; j < tn; ++j) {
        const uint4 a = s_train[2 * j], b = s_train[2 * j + 1];
        const uint32_t id = (uint32_t)(tbase + j);
#pragma unroll
        for (int r = 0; r < kQPT; ++r) {
          const uint32_t key = (ham256(q[r], a, b) << 16) | id;
          m2[r] = min(m2[r], max(m1[r], key));
          m1[r] = min(m1[r], key);
        }
      }
    }
#pragma unroll
    for (int r = 0; r < kQPT; ++r) {
      const int qi = qbase + r * kThreads + tid;
      if (qi >= nt) continue;
      const int i1 = m1[r] == 0xFFFFFFFFu ? -1 : (int)(m1[r] & 0xFFFFu);
      const int d1 = m1[r] == 0xFFFFFFFFu ? -1 : (int)(m1[r] >> 16);
      const int i2 = m2[r] == 0xFFFFFFFFu ? -1 : (int)(m2[r] & 0xFFFFu);
      const int d2 = m2[r] == 0xFFFFFFFFu ? -1 : (int)(m2[r] >> 16);
      if (knn_idx) {
        const size_t o = ((size_t)pair * to.max_feat + qi) * 2;
        knn_idx[o] = i1;
        knn_idx[o + 1] = i2;
      }
      if (knn_dist) {
        const size_t o = ((size_t)pair * to.max_feat + qi) * 2;
        knn_dist[o] = d1;
        knn_dist[o + 1] = d2;
      }
      const int a = mtail::nndr_assign(i1, d1, i2, d2, nndr, strict);
      s_assigned[qi] = a;
    }
  }
  __syncthreads();
  const float* p3 = from.kp3d_early(valid_pair ? fi : 0);
  mtail::unique_and_compact<kThreads>(tid, nf, nt, p3, s_assigned, s_count, s_owner, s_scan, &s_dropped,
                                      (size_t)pair * from.max_feat, pair, n_matches, match_from, match_to, n_unique_to);
}

}  // namespace

int sepb_launch_hamming_match(sepb_ctx* ctx, FramesView from, FramesView to, const int32_t* d_pairs, int n_pairs,
                              float nndr, int strict, int32_t* d_knn_idx, int32_t* d_knn_dist,
                              int32_t* d_n_matches, uint16_t* d_match_from, uint16_t* d_match_to,
                              int32_t* d_n_unique_to) {
  if (n_pairs == 0) return SEPB_OK;
  const size_t smem = (size_t)kTTile * 32 + (size_t)to.max_feat * 4 + (size_t)from.max_feat * 4 +
                      align_up((size_t)from.max_feat * 2, 16);
  size_t& configured = ctx->cfg_hamming_smem;
  if (smem > configured) {
    SEPB_CUDA(ctx, cudaFuncSetAttribute(hamming_match_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    configured = smem;
  }
  hamming_match_kernel<<<n_pairs, kThreads, smem, ctx->stream>>>(from, to, d_pairs, n_pairs, nndr, strict, d_knn_idx,
                                                                d_knn_dist, d_n_matches, d_match_from, d_match_to,
                                                                d_n_unique_to);
  ctx->launches++;
  SEPB_CUDA(ctx, cudaGetLastError());
  return SEPB_OK;
}
