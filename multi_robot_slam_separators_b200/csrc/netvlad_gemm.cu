// K1: NetVLAD nearest-neighbour search as a tcgen05 / TMA GEMM with a fused distance threshold.
//
// Replaces `cdist(local, received)` + masks + per-row argmin of DataHandler.find_matches
// (R/scripts/data_handler.py:166-188).  The reference compares f64 Euclidean distances of
// non-unit-norm truncated descriptors with a threshold, and the selected candidate ids must be
// bit-exact, so the tensor cores act as a FILTER with a proven error bound and the survivors are
// re-ranked in f64 in cdist's summation order:
//
//   1. netvlad_split_kernel   x (f64) -> hi = bf16(x), lo = bf16(x - hi), stored [row][hi(Dp) | lo(Dp)],
//                             ||x||^2 (f32) and the running maximum of ||d||^2 over the database.
//   2. netvlad_gemm_kernel    persistent, warp-specialised: TMA (SWIZZLE_128B boxes of 64 bf16) ->
//                             shared-memory ring -> tcgen05.mma kind::f16 M=128 x N=256 x K=16 with f32
//                             accumulators in two 256-column TMEM stages.  Per 64-wide block of D three
//                             products are accumulated: hi.hi + hi.lo + lo.hi (the dropped terms are
//                             below 3 * 2^-18 |q||d|).  Epilogue (two groups of four warps, one TMEM
//                             stage each): v = ||d||^2 - 2 acc (column mask folded into ||d||^2 = +inf),
//                             FUSED THRESHOLD v <= (k-th best v of the row so far) + 2E, survivors go
//                             to a short sorted per-row list in shared memory (ignored cells are
//                             checked there).  E = C(D) (||q||^2 + max||d||^2) bounds |v~ - v|.
//   3. netvlad_rerank_kernel  warp per query: union of the row's lists, tau = v_(k) + 2E; proves the
//                             candidate set complete (nothing dropped from a full list at or below
//                             tau), computes the exact f64 distances of the candidates (sequential sum,
//                             no FMA, sqrt -- bit-identical to scipy's cdist) and emits the top-k by
//                             (distance, index).  Rows that cannot be proven complete, or with
//                             non-finite norms, go to a fallback list.
//   4. netvlad_exact_kernel   (netvlad_exact.cu, gather mode) full f64 scan of the fallback rows only.
//
// Completeness: let a_j be the exact squared distances and v~_j the filter values of a row, |v~_j - a_j| <= E.
// At least k columns have a_j <= v~_(k) + E, so every true top-k member j has a_j <= v~_(k) + E and therefore
// v~_j <= v~_(k) + 2E = tau.  A list only ever rejects values above (its own k-th best) + 2E >= tau, or records
// the smallest value it dropped for lack of space; if that is <= tau the row falls back to the exact scan.
#include <cuda.h>
#include <cuda_bf16.h>
#include <math.h>
#include <stdlib.h>

#include <algorithm>

#include "common.cuh"
#include "umma.cuh"

namespace {

constexpr int BM = 128;            // query rows per tile (TMEM lanes)
constexpr int BN = 256;            // database rows per tile (TMEM columns of one accumulator stage)
constexpr int KB = 64;             // bf16 elements per K-block (128 bytes: one swizzle row)
constexpr int kStages = 2;         // ring stages; a stage holds B_hi, B_lo (and A_hi, A_lo when A streams)
constexpr int kABlockBytes = BM * 128;   // 16 KB
constexpr int kBBlockBytes = BN * 128;   // 32 KB
constexpr int KP = 12;             // per-list capacity (k <= 8)
constexpr int kMaxK = 8;
constexpr int kEpiWarps = 8;
constexpr int kSeedTiles = 0;                      // database tiles of the threshold seed pass
constexpr int kThreads = (2 + kEpiWarps) * 32;   // warp 0 TMA, warp 1 MMA, warps 2..9 epilogue
constexpr int kQueue = 256;        // candidates per row the re-rank kernel handles before falling back
constexpr int kMaxLists = 1024 / KP;

struct GemmParams {
  int nq, n, Dp, k;
  int row_blocks, splits, tiles_n, n_items;
  int64_t base;
  float C;                         // error coefficient: E = C * (||q||^2 + max ||d||^2)
  const float* q_norm2;            // [nq]
  const float* d_norm2;            // [n]
  const float* d_maxnorm2;         // scalar
  const uint8_t* row_mask;         // [nq] or null
  const uint8_t* col_mask;         // [n] or null
  const int32_t* ign_start;        // CSR of ignored cells or null
  const int32_t* ign_col;
  int lists;                       // 2 * splits candidate lists per row
  float* row_thr;                  // [nq] smallest threshold reached by any list of the row (shared, racy min)
  float* list_v;                   // [nq][lists][KP]
  int32_t* list_i;                 // [nq][lists][KP]  (local column index, -1 = empty)
  float* list_drop;                // [nq][lists] smallest value dropped for lack of space (+inf = none)
  float* dbg_v;                    // optional [nq][n] filter values (tests)
  int32_t* stats;                  // optional: [3] += list insert events
  int dbg_flags;                   // developer switches (SEPB_K1_DBG env): 2 = skip the epilogue scan, 4 = skip the MMAs
};

// 2E for a row, rounded up; NaN / inf when a norm is not finite
__device__ __forceinline__ float two_e(float C, float nq2, float maxnd2) {
  return __fmul_ru(2.0f * C, __fadd_ru(nq2, maxnd2));
}
__device__ __forceinline__ bool row_filterable(float nq2, float maxnd2) {
  return isfinite(nq2) && isfinite(maxnd2) && (nq2 + maxnd2) > 1e-30f;
}

// ---- 1. split conversion ---------------------------------------------------------------------
__global__ void netvlad_split_kernel(const double* __restrict__ src, int r0, int r1, int D, int Dp,
                                     __nv_bfloat16* __restrict__ dst, float* __restrict__ norm2,
                                     float* __restrict__ maxnorm2, float* __restrict__ row_thr = nullptr,
                                     int32_t* __restrict__ zero_me = nullptr) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  const int r = r0 + warp;
  if (zero_me && blockIdx.x == 0 && threadIdx.x == 0) *zero_me = 0;   // query pass: fallback-row counter
  if (r >= r1) return;
  if (row_thr && lane == 0) row_thr[r] = __uint_as_float(0x7f7f7f7fu);   // query pass: 3.39e38, no threshold known yet
  const double* x = src + (size_t)r * D;
  __nv_bfloat16* o = dst + (size_t)r * 2 * Dp;
  double s = 0.0;
  for (int j = lane; j < Dp; j += 32) {
    double v = j < D ? x[j] : 0.0;
    const __nv_bfloat16 hi = __double2bfloat16(v);
    const double rem = v - (double)__bfloat162float(hi);
    o[j] = hi;
    o[Dp + j] = __double2bfloat16(rem);
    s += v * v;
  }
#pragma unroll
  for (int m = 16; m > 0; m >>= 1) s += __shfl_xor_sync(0xFFFFFFFFu, s, m);
  if (lane == 0) {
    float f = __double2float_ru(s);
    if (!(f == f)) f = INFINITY;   // NaN rows poison the bound -> every row falls back to the exact scan
    norm2[r] = __double2float_rn(s);
    if (maxnorm2) atomicMax(reinterpret_cast<unsigned int*>(maxnorm2), __float_as_uint(f));
  }
}

// ---- 2. GEMM + fused threshold ------------------------------------------------------------------
__global__ void __launch_bounds__(kThreads, 1)
netvlad_gemm_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmD, GemmParams p) {
  extern __shared__ uint8_t smem_raw[];
  // 1024-byte alignment by pointer arithmetic on the shared array (keeps the shared address space: LDS/STS, not LD.E)
  uint8_t* smem = smem_raw + ((1024u - (umma::smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t* sB = smem;                                        // [kStages][2][BN][128]
  uint8_t* sA = sB + kStages * 2 * kBBlockBytes;             // [2 * kStages slots][BM][128]
  float* s_lv = reinterpret_cast<float*>(sA + 2 * kStages * kABlockBytes);   // [2 groups][KP][128]
  int32_t* s_li = reinterpret_cast<int32_t*>(s_lv + 2 * KP * 128);
  float* s_nd = reinterpret_cast<float*>(s_li + 2 * KP * 128);             // [2 groups][2 buffers][BN]
  __shared__ uint64_t bar_full[kStages], bar_empty[kStages], bar_t_full[2], bar_t_empty[2];
  __shared__ uint32_t tmem_slot;

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int nkb = p.Dp / KB;
  const bool a_resident = 2 * nkb <= 2 * kStages;   // all K-blocks of the query tile fit the A slots

  if (tid == 0) {
    for (int s = 0; s < kStages; ++s) {
      umma::mbar_init(&bar_full[s], 1);
      umma::mbar_init(&bar_empty[s], 1);
    }
    for (int s = 0; s < 2; ++s) {
      umma::mbar_init(&bar_t_full[s], 1);
      umma::mbar_init(&bar_t_empty[s], 4);
    }
    umma::fence_mbar_init();
    umma::tma_prefetch_desc(&tmQ);
    umma::tma_prefetch_desc(&tmD);
  }
  if (warp == 1) umma::tmem_alloc(&tmem_slot, 512);
  umma::tc_fence_before();
  __syncthreads();
  umma::tc_fence_after();
  const uint32_t tmem = tmem_slot;

  if (warp == 0) {
    // ---- TMA producer (one elected lane) --------------------------------------------------------
    if (umma::elect_one()) {
      uint32_t uses[kStages];
      for (int s = 0; s < kStages; ++s) uses[s] = 0;
      uint32_t it = 0;
      for (int item = blockIdx.x; item < p.n_items; item += gridDim.x) {
        const int sp = item / p.row_blocks, rb = item % p.row_blocks;
        const int t_lo = (int)((int64_t)p.tiles_n * sp / p.splits), t_hi = (int)((int64_t)p.tiles_n * (sp + 1) / p.splits);
        for (int t = t_lo; t < t_hi; ++t) {
          for (int kb = 0; kb < nkb; ++kb, ++it) {
            const int s = it % kStages;
            const bool load_a = !a_resident || t == t_lo;
            if (a_resident && t == t_lo && kb == 0) {
              // the query tile is about to be overwritten: every MMA issued so far must have retired
              for (int s2 = 0; s2 < kStages; ++s2) umma::mbar_wait(&bar_empty[s2], (uses[s2] & 1) ^ 1);
            } else {
              umma::mbar_wait(&bar_empty[s], (uses[s] & 1) ^ 1);
            }
            ++uses[s];
            umma::mbar_arrive_expect_tx(&bar_full[s], (uint32_t)(2 * kBBlockBytes + (load_a ? 2 * kABlockBytes : 0)));
            uint8_t* b = sB + (size_t)s * 2 * kBBlockBytes;
            umma::tma_load_2d(b, &tmD, &bar_full[s], kb * KB, t * BN);
            umma::tma_load_2d(b + kBBlockBytes, &tmD, &bar_full[s], p.Dp + kb * KB, t * BN);
            if (load_a) {
              uint8_t* a = sA + (size_t)(a_resident ? kb : s) * 2 * kABlockBytes;
              umma::tma_load_2d(a, &tmQ, &bar_full[s], kb * KB, rb * BM);
              umma::tma_load_2d(a + kABlockBytes, &tmQ, &bar_full[s], p.Dp + kb * KB, rb * BM);
            }
          }
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ---- MMA issuer -------------------------------------------------------------------------------
    const uint32_t idesc = umma::instr_desc(1, 1, 1, BM, BN);   // f32 accumulate, bf16 x bf16
    const uint32_t a0 = umma::smem_u32(sA), b0 = umma::smem_u32(sB);
    uint32_t it = 0, tc = 0;
    for (int item = blockIdx.x; item < p.n_items; item += gridDim.x) {
      const int sp = item / p.row_blocks;
      const int t_lo = (int)((int64_t)p.tiles_n * sp / p.splits), t_hi = (int)((int64_t)p.tiles_n * (sp + 1) / p.splits);
      for (int t = t_lo; t < t_hi; ++t, ++tc) {
        const uint32_t st = tc & 1;
        umma::mbar_wait(&bar_t_empty[st], ((tc >> 1) & 1) ^ 1);
        for (int kb = 0; kb < nkb; ++kb, ++it) {
          const int s = it % kStages;
          umma::mbar_wait(&bar_full[s], (it / kStages) & 1);
          umma::tc_fence_after();
          if (umma::elect_one()) {
            const uint32_t d = tmem + st * BN;
            const uint32_t a_hi = a0 + (uint32_t)((a_resident ? kb : s) * 2 * kABlockBytes), a_lo = a_hi + kABlockBytes;
            const uint32_t b_hi = b0 + (uint32_t)(s * 2 * kBBlockBytes), b_lo = b_hi + kBBlockBytes;
            if (!(p.dbg_flags & 4)) {
#pragma unroll
            for (int k = 0; k < 4; ++k)
              umma::mma_f16(d, umma::smem_desc_sw128(a_hi + k * 32), umma::smem_desc_sw128(b_hi + k * 32), idesc,
                            (kb | k) != 0);
#pragma unroll
            for (int k = 0; k < 4; ++k)
              umma::mma_f16(d, umma::smem_desc_sw128(a_hi + k * 32), umma::smem_desc_sw128(b_lo + k * 32), idesc, true);
#pragma unroll
            for (int k = 0; k < 4; ++k)
              umma::mma_f16(d, umma::smem_desc_sw128(a_lo + k * 32), umma::smem_desc_sw128(b_hi + k * 32), idesc, true);
            }
            umma::mma_commit(&bar_empty[s]);
            if (kb == nkb - 1) umma::mma_commit(&bar_t_full[st]);
          }
          __syncwarp();
        }
      }
    }
  } else {
    // ---- epilogue: fused threshold + per-row candidate buffers ---------------------------------------
    const int ew = warp - 2, grp = ew >> 2, lq = warp & 3;   // TMEM lane quarter = warp % 4
    const int t128 = lq * 32 + lane;                          // row inside the query tile
    const int gtid = (ew & 3) * 32 + lane;                    // thread index inside the group (staging)
    float* bv = s_lv + grp * KP * 128 + t128;                 // this thread's candidate buffer: bv[j * 128], bi[j * 128]
    int32_t* bi = s_li + grp * KP * 128 + t128;
    float* nd_base = s_nd + grp * 2 * BN;
    const float maxnd2 = *p.d_maxnorm2;
    const int k = p.k;
    uint32_t tc = 0, my_tiles = 0;
    for (int item = blockIdx.x; item < p.n_items; item += gridDim.x) {
      const int sp = item / p.row_blocks, rb = item % p.row_blocks;
      const int t_lo = (int)((int64_t)p.tiles_n * sp / p.splits), t_hi = (int)((int64_t)p.tiles_n * (sp + 1) / p.splits);
      const int row = rb * BM + t128;
      const bool row_ok = row < p.nq && !(p.row_mask && p.row_mask[row]);
      const float nq2 = row < p.nq ? p.q_norm2[row] : 0.f;
      const bool live = row_ok && row_filterable(nq2, maxnd2);
      const float e2 = live ? two_e(p.C, nq2, maxnd2) : 0.f;
      const int ig0 = (p.ign_start && live) ? p.ign_start[row] : 0, ig1 = (p.ign_start && live) ? p.ign_start[row + 1] : 0;
      volatile float* U = p.row_thr + (row < p.nq ? row : 0);   // best threshold any list of this row has reached
      float l[8];                                                // the (up to) 8 smallest values this list has seen
#pragma unroll
      for (int j = 0; j < 8; ++j) l[j] = INFINITY;
      int cnt = 0, n_ins = 0;
      float thr = live ? INFINITY : -INFINITY, min_drop = INFINITY;

      auto hit = [&](float v, int c) {
        ++n_ins;
        float x = v;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float lo = fminf(x, l[j]);
          x = fmaxf(x, l[j]);
          l[j] = lo;
        }
        float kth = l[7];
        switch (k) {
          case 1: kth = l[0]; break;
          case 2: kth = l[1]; break;
          case 3: kth = l[2]; break;
          case 4: kth = l[3]; break;
          case 5: kth = l[4]; break;
          case 6: kth = l[5]; break;
          case 7: kth = l[6]; break;
          default: break;
        }
        if (kth < INFINITY) thr = fminf(thr, __fadd_ru(kth, e2));
        if (cnt == KP) {   // buffer full: keep only what is still at or below the threshold
          int w = 0;
          for (int j = 0; j < KP; ++j) {
            const float y = bv[j * 128];
            if (y <= thr) {
              if (w != j) {
                bv[w * 128] = y;
                bi[w * 128] = bi[j * 128];
              }
              ++w;
            }
          }
          cnt = w;
        }
        if (v <= thr) {
          if (cnt == KP) {
            min_drop = fminf(min_drop, v);   // more than KP values inside the window: the row cannot be proven
          } else {
            bv[cnt * 128] = v;
            bi[cnt * 128] = c;
            ++cnt;
          }
        }
      };

      for (int t = t_lo; t < t_hi; ++t, ++tc) {
        if ((int)(tc & 1) != grp) continue;   // the other group drains this accumulator stage
        const uint32_t st = tc & 1;
        const float u_now = live ? *U : INFINITY;
        float* nd = nd_base + (my_tiles & 1) * BN;
        ++my_tiles;
        {
          const int c = t * BN + gtid * 2;
          float2 w;
          w.x = (c < p.n && !(p.col_mask && p.col_mask[c])) ? p.d_norm2[c] : INFINITY;
          w.y = (c + 1 < p.n && !(p.col_mask && p.col_mask[c + 1])) ? p.d_norm2[c + 1] : INFINITY;
          *reinterpret_cast<float2*>(nd + gtid * 2) = w;
        }
        asm volatile("bar.sync %0, 128;" ::"r"(1 + grp) : "memory");
        thr = fminf(thr, u_now);
        const float thr_in = thr;
        umma::mbar_wait(&bar_t_full[st], (tc >> 1) & 1);
        umma::tc_fence_after();
        const uint32_t taddr = tmem + ((uint32_t)(lq * 32) << 16) + st * BN;

        // fast pass over 32 columns: branch-free, flags the 4-column groups holding a value at or below the
        // threshold; flagged groups are re-read from TMEM in ONE out-of-line loop (a hit path inlined at every
        // unrolled position made the kernel miss the instruction cache on each hit)
        auto scan = [&](const uint32_t (&r)[32], int ch) {
          uint32_t gm = 0;
#pragma unroll
          for (int j = 0; j < 32; j += 4) {
            const float4 w = *reinterpret_cast<const float4*>(nd + ch * 32 + j);
            const float v0 = fmaf(-2.f, __uint_as_float(r[j]), w.x), v1 = fmaf(-2.f, __uint_as_float(r[j + 1]), w.y);
            const float v2 = fmaf(-2.f, __uint_as_float(r[j + 2]), w.z), v3 = fmaf(-2.f, __uint_as_float(r[j + 3]), w.w);
            const float m = fminf(fminf(v0, v1), fminf(v2, v3));
            gm |= (m <= thr) ? (1u << (j >> 2)) : 0u;
          }
          // Slow path, lane-private: a lane walks its OWN flagged groups and picks the group's four accumulators out of
          // the registers of the branch-free pass with a select tree on the group index (no TMEM re-read, which is a
          // collective load and forced a warp-uniform loop over the flagged groups of any of the 32 rows).
#pragma unroll 1
          while (gm) {
            const int g = __ffs(gm) - 1;
            gm &= gm - 1;
            const bool b0 = g & 1, b1 = g & 2, b2 = g & 4;
            uint32_t x[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
              const uint32_t s0 = b0 ? r[4 + u] : r[u], s1 = b0 ? r[12 + u] : r[8 + u];
              const uint32_t s2 = b0 ? r[20 + u] : r[16 + u], s3 = b0 ? r[28 + u] : r[24 + u];
              const uint32_t t0 = b1 ? s1 : s0, t1 = b1 ? s3 : s2;
              x[u] = b2 ? t1 : t0;
            }
            const float4 w = *reinterpret_cast<const float4*>(nd + ch * 32 + g * 4);
            const float vs[4] = {fmaf(-2.f, __uint_as_float(x[0]), w.x), fmaf(-2.f, __uint_as_float(x[1]), w.y),
                                 fmaf(-2.f, __uint_as_float(x[2]), w.z), fmaf(-2.f, __uint_as_float(x[3]), w.w)};
            uint32_t hm = (vs[0] <= thr && vs[0] < INFINITY ? 1u : 0u) | (vs[1] <= thr && vs[1] < INFINITY ? 2u : 0u) |
                          (vs[2] <= thr && vs[2] < INFINITY ? 4u : 0u) | (vs[3] <= thr && vs[3] < INFINITY ? 8u : 0u);
#pragma unroll 1
            while (hm) {
              const int u = __ffs(hm) - 1;
              hm &= hm - 1;
              const float v = u == 0 ? vs[0] : u == 1 ? vs[1] : u == 2 ? vs[2] : vs[3];
              if (v <= thr) {   // an earlier hit may have tightened the threshold
                const int c = t * BN + ch * 32 + g * 4 + u;
                bool ign = false;
                for (int q = ig0; q < ig1; ++q) ign |= p.ign_col[q] == (int32_t)(p.base + c);
                if (!ign) hit(v, c);
              }
            }
          }
          __syncwarp();
          if (p.dbg_v && row < p.nq) {
            float* o = p.dbg_v + (size_t)row * p.n;
#pragma unroll
            for (int j = 0; j < 32; ++j) {
              const int c = t * BN + ch * 32 + j;
              if (c < p.n) o[c] = fmaf(-2.f, __uint_as_float(r[j]), nd[ch * 32 + j]);
            }
          }
        };

        if (!(p.dbg_flags & 2)) {
          uint32_t ra[32], rb2[32];
          umma::tmem_ld32(taddr, ra);
#pragma unroll 1
          for (int ch = 0; ch < BN / 32; ch += 2) {
            umma::tmem_ld_wait();
            umma::tmem_ld32(taddr + (uint32_t)((ch + 1) * 32), rb2);
            scan(ra, ch);
            umma::tmem_ld_wait();
            if (ch + 2 < BN / 32) umma::tmem_ld32(taddr + (uint32_t)((ch + 2) * 32), ra);
            scan(rb2, ch + 1);
          }
        }
        umma::tc_fence_before();
        __syncwarp();
        if (lane == 0) umma::mbar_arrive(&bar_t_empty[st]);
        if (live && thr < thr_in && thr < *U) *U = thr;   // racy minimum: every value ever written is a valid bound
      }
      // flush this item's candidates (those still inside the window)
      if (row < p.nq) {
        const int list = sp * 2 + grp;
        const size_t o = ((size_t)row * p.lists + list) * KP;
        int w = 0;
        for (int j = 0; j < cnt; ++j) {
          const float y = bv[j * 128];
          if (y <= thr) {
            p.list_v[o + w] = y;
            p.list_i[o + w] = bi[j * 128];
            ++w;
          }
        }
        for (; w < KP; ++w) {
          p.list_v[o + w] = INFINITY;
          p.list_i[o + w] = -1;
        }
        p.list_drop[(size_t)row * p.lists + list] = min_drop;
        if (p.stats && n_ins) atomicAdd(p.stats + 3, n_ins);
      }
    }
  }
  umma::tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    umma::tc_fence_after();
    umma::tmem_dealloc(tmem, 512);
  }
}

// ---- 3. exact re-rank ------------------------------------------------------------------------------
struct RerankParams {
  const double* q;       // [nq][D]
  const double* db;      // [n][D]
  int nq, n, D, k, lists;
  int64_t base;
  float C;
  const float* q_norm2;
  const float* d_maxnorm2;
  const uint8_t* row_mask;
  const float* list_v;
  const int32_t* list_i;
  const float* list_drop;
  int32_t* out_idx;      // [nq][k]
  double* out_dist;
  int32_t* fb_rows;      // fallback row list
  int32_t* fb_count;
  int32_t* stats;        // [4]: candidates re-ranked, rows proven by the filter, -, -
};

constexpr int kRerankWarps = 4;
constexpr int kRrGroup = 8;     // survivors whose rows are staged together
constexpr int kRrChunk = 64;    // dimensions per staged chunk (rows padded to kRrChunk + 1 doubles: conflict-free chains)

__global__ void __launch_bounds__(kRerankWarps * 32) netvlad_rerank_kernel(RerankParams p) {
  __shared__ float s_qv[kRerankWarps][kQueue];
  __shared__ int32_t s_ci[kRerankWarps][kQueue];
  __shared__ double s_cd[kRerankWarps][kQueue];
  __shared__ uint16_t s_act[kRerankWarps][kQueue];
  __shared__ double s_rq[kRerankWarps][kRrChunk];
  __shared__ double s_rr[kRerankWarps][kRrGroup][kRrChunk + 1];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int row = blockIdx.x * kRerankWarps + warp;
  if (row >= p.nq) return;
  int32_t* o_i = p.out_idx + (size_t)row * p.k;
  double* o_d = p.out_dist + (size_t)row * p.k;
  if (p.row_mask && p.row_mask[row]) {
    if (lane < p.k) {
      o_i[lane] = -1;
      o_d[lane] = INFINITY;
    }
    return;
  }
  const float nq2 = p.q_norm2[row], maxnd2 = *p.d_maxnorm2;
  bool fallback = !row_filterable(nq2, maxnd2);
  const int n_ent = p.lists * KP;
  int nc = 0;
  if (!fallback) {
    // gather the row's buffered candidates (contiguous [lists][KP]) into the warp's queue
    const float* lv = p.list_v + (size_t)row * n_ent;
    const int32_t* li = p.list_i + (size_t)row * n_ent;
    for (int e0 = 0; e0 < n_ent; e0 += 32) {
      const int e = e0 + lane;
      int32_t ci = -1;
      float v = INFINITY;
      if (e < n_ent) {
        ci = li[e];
        v = lv[e];
      }
      const bool ok = ci >= 0 && v < INFINITY;
      const uint32_t m = __ballot_sync(0xFFFFFFFFu, ok);
      const int pos = nc + __popc(m & ((1u << lane) - 1));
      if (nc + __popc(m) > kQueue) {
        fallback = true;
        break;
      }
      if (ok) {
        s_qv[warp][pos] = v;
        s_ci[warp][pos] = ci;
      }
      nc += __popc(m);
    }
    __syncwarp();
  }
  if (!fallback) {
    // tau = (k-th smallest filter value of the union) + 2E: k rounds of warp arg-min with removal
    const float e2 = two_e(p.C, nq2, maxnd2);
    uint32_t taken = 0;
    float vk = INFINITY;
    for (int r = 0; r < p.k; ++r) {
      float bvv = INFINITY;
      int be = -1;
      for (int c = lane, sl = 0; c < nc; c += 32, ++sl) {
        if (taken & (1u << sl)) continue;
        const float v = s_qv[warp][c];
        if (v < bvv) {
          bvv = v;
          be = c;
        }
      }
      float wv = bvv;
      int we = be;
#pragma unroll
      for (int m = 16; m > 0; m >>= 1) {
        const float ov = __shfl_xor_sync(0xFFFFFFFFu, wv, m);
        const int oe = __shfl_xor_sync(0xFFFFFFFFu, we, m);
        if (oe >= 0 && (we < 0 || ov < wv || (ov == wv && oe < we))) {
          wv = ov;
          we = oe;
        }
      }
      if (we < 0) {
        vk = INFINITY;   // fewer than k candidates exist at all
        break;
      }
      if (we == be) taken |= 1u << ((we - lane) >> 5);
      vk = wv;
    }
    const float tau = vk < INFINITY ? __fadd_ru(vk, e2) : INFINITY;
    // completeness: nothing at or below tau may have been dropped for lack of buffer space
    bool bad = false;
    for (int l = lane; l < p.lists; l += 32) {
      const float d = p.list_drop[(size_t)row * p.lists + l];
      bad |= d < INFINITY && d <= tau;
    }
    fallback = __any_sync(0xFFFFFFFFu, bad);
    if (!fallback) {
      // exact distances in cdist's order: s = 0; for j: d = u[j] - v[j]; s += d*d; sqrt(s)   (no FMA).  The sum of a
      // candidate is a serial chain owned by one lane, but the ROWS are fetched by the whole warp: the survivors are taken
      // kRrGroup at a time, their rows (and the query's) staged through shared memory in chunks of kRrChunk dimensions
      // with coalesced loads (a lane reading its own row 8 bytes at a time spends its life waiting for L2: measured
      // 77 us for 10 k queries x 128-d, half of the stall samples on the first use of the loaded value)
      const double* qr = p.q + (size_t)row * p.D;
      int n_exact = 0;
      for (int c = lane; c < nc; c += 32) s_cd[warp][c] = INFINITY;
      __syncwarp();
      // compact the survivors (filter value <= tau), in candidate order
      int n_act = 0;
      for (int c0 = 0; c0 < nc; c0 += 32) {
        const int c = c0 + lane;
        const bool act = c < nc && s_qv[warp][c] <= tau;
        const uint32_t m = __ballot_sync(0xFFFFFFFFu, act);
        if (act) s_act[warp][n_act + __popc(m & ((1u << lane) - 1))] = (uint16_t)c;
        n_act += __popc(m);
      }
      __syncwarp();
      n_exact = lane == 0 ? n_act : 0;   // (summed over the lanes for the statistics below)
      for (int g0 = 0; g0 < n_act; g0 += kRrGroup) {
        const int gn = min(kRrGroup, n_act - g0);
        const int myc = lane < gn ? s_act[warp][g0 + lane] : 0;
        const double* rowp[kRrGroup];
#pragma unroll
        for (int k = 0; k < kRrGroup; ++k)
          rowp[k] = p.db + (size_t)s_ci[warp][s_act[warp][g0 + min(k, gn - 1)]] * p.D;
        // software pipeline: the loads of chunk i + 1 are in flight while the chains walk chunk i
        constexpr int kT = kRrChunk / 32;
        double cur[kRrGroup][kT], qcur[kT];
        auto fetch = [&](int d0) {
          const int dn = min(kRrChunk, p.D - d0);
#pragma unroll
          for (int t = 0; t < kT; ++t) {
            const int j = lane + 32 * t;
            qcur[t] = j < dn ? qr[d0 + j] : 0.0;
#pragma unroll
            for (int k = 0; k < kRrGroup; ++k) cur[k][t] = (k < gn && j < dn) ? rowp[k][d0 + j] : 0.0;
          }
        };
        fetch(0);
        double sacc = 0.0;
        for (int d0 = 0; d0 < p.D; d0 += kRrChunk) {
          const int dn = min(kRrChunk, p.D - d0);
#pragma unroll
          for (int t = 0; t < kT; ++t) {
            s_rq[warp][lane + 32 * t] = qcur[t];
#pragma unroll
            for (int k = 0; k < kRrGroup; ++k) s_rr[warp][k][lane + 32 * t] = cur[k][t];
          }
          __syncwarp();
          if (d0 + kRrChunk < p.D) fetch(d0 + kRrChunk);
          if (lane < gn) {
            const double* rr = s_rr[warp][lane];
            for (int j = 0; j < dn; ++j) {
              const double d = __dsub_rn(s_rq[warp][j], rr[j]);
              sacc = __dadd_rn(sacc, __dmul_rn(d, d));
            }
          }
          __syncwarp();
        }
        if (lane < gn) {
          double sv = __dsqrt_rn(sacc);
          if (!(sv == sv) || isinf(sv)) sv = INFINITY;   // NaN / inf are never candidates (netvlad_exact.cu)
          s_cd[warp][myc] = sv;
        }
      }
      __syncwarp();
      // top-k by (distance, index): k rounds of warp arg-min with removal
      for (int r = 0; r < p.k; ++r) {
        double bd = INFINITY;
        int32_t bi = INT32_MAX;
        int bc = -1;
        for (int c = lane; c < nc; c += 32) {
          const double d = s_cd[warp][c];
          const int32_t i = s_ci[warp][c];
          if (d < INFINITY && (d < bd || (d == bd && i < bi))) {
            bd = d;
            bi = i;
            bc = c;
          }
        }
        double wd = bd;
        int32_t wi = bi;
#pragma unroll
        for (int m = 16; m > 0; m >>= 1) {
          const double od = __shfl_xor_sync(0xFFFFFFFFu, wd, m);
          const int32_t oi = __shfl_xor_sync(0xFFFFFFFFu, wi, m);
          if (od < wd || (od == wd && oi < wi)) {
            wd = od;
            wi = oi;
          }
        }
        if (bc >= 0 && wi == bi && wd == bd) s_cd[warp][bc] = INFINITY;   // column indices are unique within a row
        if (lane == 0) {
          const bool none = wi == INT32_MAX;
          o_i[r] = none ? -1 : (int32_t)(p.base + wi);
          o_d[r] = none ? INFINITY : wd;
        }
        __syncwarp();
      }
      if (p.stats) {
#pragma unroll
        for (int m = 16; m > 0; m >>= 1) n_exact += __shfl_xor_sync(0xFFFFFFFFu, n_exact, m);
        if (lane == 0) {
          atomicAdd(p.stats + 0, n_exact);
          atomicAdd(p.stats + 1, 1);
        }
      }
      return;
    }
  }
  if (lane == 0) p.fb_rows[atomicAdd(p.fb_count, 1)] = row;
}

typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                             const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                             CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeFn encode_fn() {
  static EncodeFn fn = [] {
    EncodeFn f = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", (void**)&f, cudaEnableDefault, &q) != cudaSuccess) f = nullptr;
    return f;
  }();
  return fn;
}

// [rows][2*Dp] bf16, boxes of 64 elements x box_rows rows, 128-byte swizzle, out-of-range rows read as zero
int make_tmap(sepb_ctx* ctx, CUtensorMap* m, const void* base, int rows, int Dp, int box_rows) {
  EncodeFn enc = encode_fn();
  if (!enc) return sepb_fail(ctx, SEPB_E_STATE, "cuTensorMapEncodeTiled unavailable", __FILE__, __LINE__);
  cuuint64_t dims[2] = {(cuuint64_t)(2 * Dp), (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)(2 * Dp) * 2};
  cuuint32_t box[2] = {(cuuint32_t)KB, (cuuint32_t)box_rows};
  cuuint32_t es[2] = {1, 1};
  const CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, es,
                         CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                         CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return sepb_fail(ctx, SEPB_E_STATE, "cuTensorMapEncodeTiled failed", __FILE__, __LINE__);
  return SEPB_OK;
}

size_t gemm_smem_bytes() {
  return (size_t)kStages * 2 * kBBlockBytes + (size_t)2 * kStages * kABlockBytes + (size_t)2 * KP * 128 * 8 +
         (size_t)2 * 2 * BN * 4 + 1024;
}

// error coefficient of the filter value v~ = ||d||^2 - 2 (hi.hi + hi.lo + lo.hi) relative to (||q||^2 + ||d||^2):
// dropped split terms 3 * 2^-18 |q||d|, one f32 accumulator rounding per 16-wide MMA (3 Dp / 16 of them, 2^-23
// each, relative to sum |q_i d_i| <= |q||d|), f32 norms and the final fma; |q||d| <= (|q|^2 + |d|^2) / 2.
// 1.5x safety on top.  tests/test_k1_netvlad_gpu.py measures the real error against this bound.
float filter_coeff(int Dp) {
  const double dot = 3.0 * ldexp(1.0, -18) + (3.0 * Dp / 16.0) * ldexp(1.0, -23);
  return (float)(1.5 * (dot + 3.0 * ldexp(1.0, -24)));
}

// work decomposition: items = row blocks x database splits, round-robin over a persistent grid
void choose_splits(int sm, int row_blocks, int tiles_n, int* splits, int* grid) {
  const int max_s = std::max(1, std::min(tiles_n, kMaxLists / 2));
  double best = 1e300;
  int bs = 1;
  for (int s = 1; s <= max_s; ++s) {
    const int items = row_blocks * s, g = std::min(sm, items);
    const double rounds = (double)((items + g - 1) / g);
    const double cost = rounds * ((double)((tiles_n + s - 1) / s) + 0.75) + 0.002 * s;
    if (cost < best) {
      best = cost;
      bs = s;
    }
  }
  *splits = bs;
  *grid = std::min(sm, row_blocks * bs);
}

}  // namespace

int sepb_netvlad_tensor_supported(int dim, int k) { return k <= kMaxK && dim >= 1 && dim <= 16384; }

// d_dbg_v: optional [nq][n] f32 filter values; d_stats: optional int32[4]
int sepb_netvlad_tensor_topk(sepb_ctx* ctx, const sepb_db* cdb, const double* d_q, int nq, int k,
                             const uint8_t* d_row_mask, const int32_t* d_ign_start, const int32_t* d_ign_col,
                             int32_t* d_out_idx, double* d_out_dist, float* d_dbg_v, int32_t* d_stats) {
  sepb_db* db = const_cast<sepb_db*>(cdb);   // lazily filled operand cache
  const int n = db->size, D = db->dim;
  const int Dp = (D + KB - 1) / KB * KB;
  if (n == 0) {
    // no columns: every row is empty
    return sepb_launch_netvlad_exact(ctx, d_q, nq, db->rows, 0, D, k, db->base, d_row_mask, nullptr, d_ign_start,
                                     d_ign_col, 1, nullptr, nullptr, d_out_idx, d_out_dist);
  }
  // ---- database operands (converted once per appended row) ------------------------------------------
  if (!db->split) {
    db->dim_pad = Dp;
    SEPB_CUDA(ctx, cudaMalloc(&db->split, (size_t)db->capacity * 2 * Dp * 2));
    SEPB_CUDA(ctx, cudaMalloc(&db->norm2, (size_t)db->capacity * 4));
    SEPB_CUDA(ctx, cudaMalloc(&db->maxnorm2, 256));
    db->split_rows = 0;
  }
  if (db->split_rows == 0) SEPB_CUDA(ctx, cudaMemsetAsync(db->maxnorm2, 0, 4, ctx->stream));
  if (db->split_rows < n) {
    const int rows = n - db->split_rows;
    netvlad_split_kernel<<<(rows + 7) / 8, 256, 0, ctx->stream>>>(db->rows, db->split_rows, n, D, Dp,
                                                                  reinterpret_cast<__nv_bfloat16*>(db->split), db->norm2,
                                                                  db->maxnorm2);
    ctx->launches++;
    SEPB_CUDA(ctx, cudaGetLastError());
    db->split_rows = n;
  }
  if (db->tmap_rows != n) {
    int rc = make_tmap(ctx, reinterpret_cast<CUtensorMap*>(db->tmap), db->split, n, Dp, BN);
    if (rc) return rc;
    db->tmap_rows = n;
  }
  // ---- work decomposition + scratch ---------------------------------------------------------------------
  const int row_blocks = (nq + BM - 1) / BM, tiles_n = (n + BN - 1) / BN;
  int splits, grid;
  choose_splits(ctx->sm_count, row_blocks, tiles_n, &splits, &grid);
  if (const char* e = sepb_dev_env("SEPB_K1_SPLITS")) {   // developer override
    splits = std::max(1, std::min(atoi(e), std::min(tiles_n, kMaxLists / 2)));
    grid = std::min(ctx->sm_count, row_blocks * splits);
  }
  const int lists = 2 * splits;
  const int fb_splits = std::max(1, std::min(64, (n + 31) / 32));
  size_t off = 0;
  auto take = [&](size_t bytes) {
    const size_t o = off;
    off += align_up(bytes, 256);
    return o;
  };
  const size_t o_qs = take((size_t)nq * 2 * Dp * 2), o_qn = take((size_t)nq * 4), o_lv = take((size_t)lists * nq * KP * 4),
               o_li = take((size_t)lists * nq * KP * 4), o_ld = take((size_t)lists * nq * 4), o_fb = take((size_t)nq * 4), o_u = take((size_t)nq * 4),
               o_cnt = take(256), o_pi = take((size_t)fb_splits * nq * k * 4), o_pd = take((size_t)fb_splits * nq * k * 8);
  uint8_t* s = (uint8_t*)sepb_scratch(ctx, off);
  if (!s) return (int)cudaErrorMemoryAllocation;
  __nv_bfloat16* q_split = reinterpret_cast<__nv_bfloat16*>(s + o_qs);
  float* q_norm2 = reinterpret_cast<float*>(s + o_qn);
  int32_t* fb_rows = reinterpret_cast<int32_t*>(s + o_fb);
  int32_t* fb_count = reinterpret_cast<int32_t*>(s + o_cnt);
  // the query pass of the split kernel also resets the fallback counter and the shared row thresholds
  netvlad_split_kernel<<<(nq + 7) / 8, 256, 0, ctx->stream>>>(d_q, 0, nq, D, Dp, q_split, q_norm2, nullptr,
                                                             reinterpret_cast<float*>(s + o_u), fb_count);
  ctx->launches++;
  SEPB_CUDA(ctx, cudaGetLastError());
  CUtensorMap tmQ;
  {
    int rc = make_tmap(ctx, &tmQ, q_split, nq, Dp, BM);
    if (rc) return rc;
  }
  const float C = filter_coeff(Dp);
  GemmParams gp{};
  gp.nq = nq; gp.n = n; gp.Dp = Dp; gp.k = k;
  gp.row_blocks = row_blocks; gp.splits = splits; gp.tiles_n = tiles_n; gp.n_items = row_blocks * splits;
  gp.base = db->base; gp.C = C;
  gp.q_norm2 = q_norm2; gp.d_norm2 = db->norm2; gp.d_maxnorm2 = db->maxnorm2;
  gp.row_mask = d_row_mask; gp.col_mask = db->has_mask ? db->col_mask : nullptr;
  gp.ign_start = d_ign_start; gp.ign_col = d_ign_col;
  gp.lists = lists; gp.row_thr = reinterpret_cast<float*>(s + o_u);
  gp.list_v = reinterpret_cast<float*>(s + o_lv); gp.list_i = reinterpret_cast<int32_t*>(s + o_li);
  gp.list_drop = reinterpret_cast<float*>(s + o_ld); gp.dbg_v = d_dbg_v; gp.stats = d_stats;
  {
    const char* e = sepb_dev_env("SEPB_K1_DBG");
    gp.dbg_flags = e ? atoi(e) : 0;
  }
  const size_t smem = gemm_smem_bytes();
  if (!ctx->cfg_gemm) {
    SEPB_CUDA(ctx, cudaFuncSetAttribute(netvlad_gemm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    ctx->cfg_gemm = true;
  }
  // Seed pass: the same kernel over the first kSeedTiles database tiles only (one split, its lists are overwritten by
  // the main pass).  It leaves row_thr = (k-th smallest of a 512-column sample) + 2E, a valid bound like every value
  // written there, so the 2 * splits lists of the main pass start inside a window instead of at +inf and take a
  // fraction of the list inserts (437 -> 99 per row at 10 k x 10 k, k = 5).  Off by default: the out-of-line hit loop is
  // warp-uniform over 32 rows, so 4.4x fewer inserts only removed a third of its iterations and the extra launch cost
  // more than that (0.362 -> 0.400 ms measured).  Kept as the first half of a per-lane candidate queue.
  int seed_tiles = (kSeedTiles > 0 && tiles_n >= 4 * kSeedTiles) ? kSeedTiles : 0;
  if (const char* e = sepb_dev_env("SEPB_K1_SEED")) seed_tiles = std::max(0, std::min(atoi(e), tiles_n));   // developer override
  if (seed_tiles > 0) {
    GemmParams sp = gp;
    sp.tiles_n = seed_tiles;
    sp.n = std::min(n, seed_tiles * BN);
    sp.splits = 1;
    sp.n_items = row_blocks;
    sp.lists = 2;
    sp.dbg_v = nullptr;
    sp.stats = nullptr;
    netvlad_gemm_kernel<<<std::min(ctx->sm_count, row_blocks), kThreads, smem, ctx->stream>>>(
        tmQ, *reinterpret_cast<const CUtensorMap*>(db->tmap), sp);
    ctx->launches++;
    SEPB_CUDA(ctx, cudaGetLastError());
  }
  netvlad_gemm_kernel<<<grid, kThreads, smem, ctx->stream>>>(tmQ, *reinterpret_cast<const CUtensorMap*>(db->tmap), gp);
  ctx->launches++;
  SEPB_CUDA(ctx, cudaGetLastError());
  RerankParams rp{};
  rp.q = d_q; rp.db = db->rows; rp.nq = nq; rp.n = n; rp.D = D; rp.k = k; rp.lists = lists; rp.base = db->base; rp.C = C;
  rp.q_norm2 = q_norm2; rp.d_maxnorm2 = db->maxnorm2; rp.row_mask = d_row_mask;
  rp.list_v = gp.list_v; rp.list_i = gp.list_i; rp.list_drop = gp.list_drop;
  rp.out_idx = d_out_idx; rp.out_dist = d_out_dist; rp.fb_rows = fb_rows; rp.fb_count = fb_count; rp.stats = d_stats;
  netvlad_rerank_kernel<<<(nq + kRerankWarps - 1) / kRerankWarps, kRerankWarps * 32, 0, ctx->stream>>>(rp);
  ctx->launches++;
  SEPB_CUDA(ctx, cudaGetLastError());
  // ---- rows the filter could not prove: full f64 scan (blocks beyond the device-side count exit at once) ----
  return sepb_launch_netvlad_exact(ctx, d_q, nq, db->rows, n, D, k, db->base, d_row_mask,
                                   db->has_mask ? db->col_mask : nullptr, d_ign_start, d_ign_col, fb_splits,
                                   reinterpret_cast<int32_t*>(s + o_pi), reinterpret_cast<double*>(s + o_pd), d_out_idx,
                                   d_out_dist, fb_rows, fb_count);
}
