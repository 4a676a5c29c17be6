// K2T: brute-force 256-bit Hamming 2-NN on the 5th-generation tensor cores (tcgen05, kind::i8).
//
// Same contract and bit-identical results as hamming_knn.cu (the POPC-pipe kernel), which replaces
// cv::BFMatcher(NORM_HAMMING).knnMatch(k=2) + NNDR + uniqueness at R/src/myRegistrationVis.cpp:839-894.
// The POPC pipe (16 lanes/clk/SM) caps that kernel at ~0.56 M pairs/s per GPU; here the 10^6 distances
// of a pair become an exact integer GEMM whose accumulators ARE the sort keys:
//
//   every descriptor bit b becomes the int8 value +11 (b = 0) or -11 (b = 1), so for a To row q and a
//   From row t   sum_k q_k t_k = 121 * (256 - 2 * hamming(q, t));
//   one extra K-block of 32 int8 carries the From row's column inside its 128-column accumulator tile:
//   A_x = (-64, -1, -128, -128, 0...), B_x(col) = (col >> 6, col & 63, pad, pad, 0...) contributing -col
//   (and -32512 for the padding rows of a partial tile, pad = 127).  The s32 accumulator is
//       acc = 30976 - (242 * hamming + col)           in [-31199, 30976]: it fits 16 bits,
//   its maximum over a row is BFMatcher's best match (smallest distance, ties to the lowest train index)
//   and the runner-up its second neighbour.  The epilogue reads TMEM with .pack::16b (two columns per
//   register) and keeps the running top-2 with packed VIMNMX(3).S16x2: 1.25 instructions per distance,
//   no key building, no masking, and everything stays exact.
//
// One CTA per keyframe pair, warp-specialised:
//   warps 0-7   epilogue: TMEM -> registers (tcgen05.ld, a whole 128-column half tile per round trip; the stage goes
//               back to the MMA warp before the scan), running top-2 per To row; two groups of four warps (TMEM
//               lane quarter = warp & 3): each drains one half of every 256-column tile (128-column geometry: the
//               tiles alternate between the groups)
//   warp  8     MMA issuer (one elected lane): per (To tile, From tile) 8 SWIZZLE_128B K-steps + 1 index K-step
//               of M=128 x N=256 into one of TWO 256-column TMEM stages (N=128: four stages; all 512 columns)
//   warps 9-12  producers: expand descriptor bits into the +-11 operand tiles in shared memory (PRMT lookups)
// A resident chunk of 512 From rows (two 256-row stages, each with its own full/empty barrier so the next
// chunk is expanded while the last To tiles still run; N=128: three 128-row stages) is expanded once per chunk; the
// 128-row To tiles stream through a 2-deep ring against it.  struct Geo holds both geometries; the launch takes the
// wide one whenever its shared memory fits (To frames up to ~1400 features).
#include <limits.h>

#include "common.cuh"
#include "match_tail.cuh"
#include "umma.cuh"

namespace {

constexpr int kEpiWarps = 8;
constexpr int kMmaWarp = 8;
constexpr int kProdWarp0 = 9;
constexpr int kProdWarps = 4;
constexpr int kProdThreads = kProdWarps * 32;
constexpr int kThreadsT = (kEpiWarps + 1 + kProdWarps) * 32;  // 416
constexpr int TM = 128;             // To rows per A tile (= TMEM lanes)
constexpr int kHalf = 128;          // columns one epilogue group drains; the column code of the index K-block is
                                    // relative to this (242 > 127)
// Geometry for an accumulator tile of TN From rows (= one TMEM stage).  TN = 256 halves the To-tile operand reads per
// multiply-add (the tensor pipe of one SM is operand-fetch limited at M128 x N128: a bare issue loop measures
// 2.79 POP/s per GPU for N = 128 against 3.43 for N = 256, tools/mma_peak.py) and both epilogue groups then drain one
// half of every tile; TN = 128 (tiles alternate between the groups) needs 32 KB less shared memory and takes the
// frames the wide geometry cannot hold.
template <int TN_>
struct Geo {
  static_assert(TN_ == 128 || TN_ == 256, "accumulator tile width");
  static constexpr int TN = TN_;
  static constexpr int kGroupsPerTile = TN / kHalf;
  static constexpr int kBStages = TN == 128 ? 3 : 2;  // From-row stages resident in shared memory
  static constexpr int CH = kBStages * TN;            // From rows per resident chunk
  static constexpr int kTStages = 512 / TN;           // TMEM accumulator stages of TN columns (all 512 columns)
  static constexpr int kTShift = TN == 128 ? 2 : 1;
  static constexpr int kBBytes = CH * 256;            // main K-blocks of the resident chunk
  // index K-block, interleaved (no-swizzle) layout.  Only the first 16 of its 32 K bytes carry anything and the To
  // side (A_x) is zero in the other 16, so the From side keeps ONE 8-row x 16-byte core matrix per 8 rows
  // (LBO = SBO = 128: the "second" core matrix of a group is the first of the next one, whatever it holds is
  // multiplied by zero; the last group's runs into A_x, which follows it in shared memory).
  static constexpr int kBxBytes = CH * 16;
  static constexpr size_t smem(int max_feat_to) {
    return (size_t)kBBytes + 2 * (TM * 256) + kBxBytes + TM * 32 + (size_t)max_feat_to * 16 + 1024;
  }
};
constexpr int kABytes = TM * 256;
constexpr int kAxBytes = TM * 32;
constexpr uint32_t kNoneKey = 0xFFFFFFFFu;

constexpr int kScale = 121;           // product of the two operand magnitudes (11 * 11)
constexpr int kStep = 2 * kScale;     // accumulator decrease per unit of Hamming distance (242 > TN - 1)
constexpr int32_t kNoneV = -32000;    // accumulators at or below this are padding / "no neighbour"
constexpr uint32_t kNone2 = 0x80008000u;

// expand one 256-bit descriptor into 256 int8 (+11 / -11) at row r of a SWIZZLE_128B tile of `rows` rows
// (shared-memory address `tile`).  Any fixed permutation of the 256 bit positions along K is fine as long
// as both operands use the same one; the one used here costs 13 instructions per 32 bits.
__device__ __forceinline__ void expand_row(uint32_t tile, int rows, int r, const uint4 a, const uint4 b) {
  const uint32_t w[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
  // byte lookups through PRMT: a selector nibble v (3 bits of the descriptor) picks byte v of an 8-byte table
  // holding +11 where the wanted bit of v is 0 and -11 (0xF5) where it is 1
  constexpr uint32_t T0 = 0xF50BF50Bu;                      // bit 0 of v
  constexpr uint32_t T1 = 0xF5F50B0Bu;                      // bit 1 of v
  constexpr uint32_t T2a = 0x0B0B0B0Bu, T2b = 0xF5F5F5F5u;  // bit 2 of v
#pragma unroll
  for (int wi = 0; wi < 8; ++wi) {
    const uint32_t wm = w[wi] & 0x77777777u, wn = (w[wi] >> 1) & 0x77777777u;
    const uint32_t wmh = wm >> 16, wnh = wn >> 16;
    const uint32_t base = tile + (uint32_t)((wi >> 2) * rows * 128 + r * 128);
    const int c0 = (wi & 3) * 2;
    umma::sts128(base + (uint32_t)(((c0) ^ (r & 7)) << 4), __byte_perm(T0, T0, wm), __byte_perm(T1, T1, wm),
                 __byte_perm(T2a, T2b, wm), __byte_perm(T2a, T2b, wn));
    umma::sts128(base + (uint32_t)(((c0 + 1) ^ (r & 7)) << 4), __byte_perm(T0, T0, wmh), __byte_perm(T1, T1, wmh),
                 __byte_perm(T2a, T2b, wmh), __byte_perm(T2a, T2b, wnh));
  }
}
__device__ __forceinline__ void zero_row(uint32_t tile, int rows, int r) {
#pragma unroll
  for (int kb = 0; kb < 2; ++kb)
#pragma unroll
    for (int c = 0; c < 8; ++c) umma::sts128(tile + (uint32_t)(kb * rows * 128 + r * 128 + c * 16), 0, 0, 0, 0);
}

__device__ __forceinline__ uint32_t max2(uint32_t a, uint32_t b) {
  uint32_t r;
  asm("max.s16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
  return r;
}
__device__ __forceinline__ uint32_t min2(uint32_t a, uint32_t b) {
  uint32_t r;
  asm("min.s16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
  return r;
}
// fold two packed registers (four accumulators) into the packed running top-2: 5 instructions
__device__ __forceinline__ void push2(uint32_t& m1, uint32_t& m2, uint32_t x, uint32_t y) {
  const uint32_t hi = max2(x, y), lo = min2(x, y);
  m2 = max2(max2(min2(m1, hi), m2), lo);
  m1 = max2(m1, hi);
}
__device__ __forceinline__ void merge2(uint32_t& a1, uint32_t& a2, uint32_t b1, uint32_t b2) {
  const uint32_t lo = min2(a1, b1);
  a1 = max2(a1, b1);
  a2 = max2(lo, max2(a2, b2));
}
__device__ __forceinline__ void merge_min(uint32_t& a1, uint32_t& a2, uint32_t b1, uint32_t b2) {
  const uint32_t hi = max(a1, b1);
  a1 = min(a1, b1);
  a2 = min(hi, min(a2, b2));
}

// 16-bit accumulator -> (hamming << 16 | global From index), the key order of hamming_knn.cu
__device__ __forceinline__ uint32_t acc_to_key(int v, int tile_base) {
  if (v <= kNoneV) return kNoneKey;
  const int kp = kScale * 256 - v;  // 242 * hamming + column
  const int ham = kp / kStep;
  return ((uint32_t)ham << 16) | (uint32_t)(tile_base + kp - ham * kStep);
}

__device__ __forceinline__ void scan32(const uint32_t (&r)[16], uint32_t (&m1)[2], uint32_t (&m2)[2]) {
#pragma unroll
  for (int j = 0; j < 16; j += 4) {
    push2(m1[0], m2[0], r[j], r[j + 1]);
    push2(m1[1], m2[1], r[j + 2], r[j + 3]);
  }
}

// Developer timeline (-DSEPB_K2T_TRACE=<pair>): the CTA of that pair stamps clock64 at every pipeline hand-off.
// Slots: 0 start, 1 setup done, 2 main loop done, 3 tail done; per accumulator tile g (< 64): 16+g MMA warp starts
// waiting for the TMEM stage, 80+g MMAs issued, 144+g epilogue (lane quarter 0) starts waiting, 208+g accumulator
// ready, 272+g stage released; per To tile `it` (< 32): 336 producer waits, 368 buffer free, 400 expanded;
// 432 + 4 * chunk + j: From stage j expanded; 464 + it: MMA warp starts waiting for To tile `it`.  tools/k2t_trace.py prints it.
#ifdef SEPB_K2T_TRACE
__device__ long long g_k2t_trace[512];
#define K2T_STAMP(cond, slot)                                              \
  do {                                                                     \
    if (blockIdx.x == (SEPB_K2T_TRACE) && (cond) && (slot) < 512) g_k2t_trace[(slot)] = clock64(); \
  } while (0)
#else
#define K2T_STAMP(cond, slot) do { } while (0)
#endif

template <int TN>
__global__ void __launch_bounds__(kThreadsT, 1)
hamming_mma_kernel(FramesView from, FramesView to, const int32_t* __restrict__ pairs, int n_pairs, float nndr, int strict,
                   int32_t* __restrict__ knn_idx, int32_t* __restrict__ knn_dist, int32_t* __restrict__ n_matches,
                   uint16_t* __restrict__ match_from, uint16_t* __restrict__ match_to,
                   int32_t* __restrict__ n_unique_to, const LazyPull lp) {
  using G = Geo<TN>;
  constexpr int kGroupsPerTile = G::kGroupsPerTile, kBStages = G::kBStages, CH = G::CH, kTStages = G::kTStages,
                kTShift = G::kTShift, kBBytes = G::kBBytes, kBxBytes = G::kBxBytes;
  extern __shared__ uint8_t smem_raw[];
  // 1024-byte alignment by pointer arithmetic on the shared array (keeps the shared address space: LDS/STS, not LD.E)
  uint8_t* smem = smem_raw + ((1024u - (umma::smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t* sB = smem;                                  // [2 K-blocks][CH][128]
  uint8_t* sA = sB + kBBytes;                          // [2 buffers][2 K-blocks][TM][128]
  uint8_t* sBx = sA + 2 * kABytes;                     // [CH/8][2][8][16]
  uint8_t* sAx = sBx + kBxBytes;                       // [TM/8][2][8][16]
  uint32_t* s_top = reinterpret_cast<uint32_t*>(sAx + kAxBytes);  // [2][to.max_feat][2] keys
  // tail arrays alias the operand tiles (all MMAs have completed by then)
  int32_t* s_assigned = reinterpret_cast<int32_t*>(sB);
  uint32_t* s_count = reinterpret_cast<uint32_t*>(s_assigned + to.max_feat);
  uint16_t* s_owner = reinterpret_cast<uint16_t*>(s_count + from.max_feat);
  __shared__ uint64_t bar_a_full[2], bar_a_empty[2], bar_b_full[kBStages], bar_b_empty[kBStages], bar_t_full[kTStages],
      bar_t_empty[kTStages];
  __shared__ uint32_t tmem_slot;
  __shared__ int s_scan[kThreadsT / 32];
  __shared__ int s_dropped;
  __shared__ int s_claim[2];

  const int pair = blockIdx.x;
  if (pair >= n_pairs) return;
  K2T_STAMP(threadIdx.x == 0, 0);
  const int fi = pairs[2 * pair], ti = pairs[2 * pair + 1];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const bool valid_pair = fi >= 0 && fi < from.size && ti >= 0 && ti < to.size;
  const int nf = valid_pair ? min(from.count[fi], from.max_feat) : 0;
  const int nt = valid_pair ? min(to.count[ti], to.max_feat) : 0;
  const uint4* g_train = reinterpret_cast<const uint4*>(from.desc + (size_t)(valid_pair ? fi : 0) * from.max_feat * 32);
  const uint4* g_query = reinterpret_cast<const uint4*>(to.desc + (size_t)(valid_pair ? ti : 0) * to.max_feat * 32);

  // producers: the rows of the first To tile and of the first From chunk are requested before the setup (and
  // before the feature counts arrive: any row below max_feat is readable), so their L2 latency hides behind it
  constexpr int kRowsB = CH / kHalf;  // From rows of a chunk per producer thread
  uint4 pre_a[2], pre_b[kRowsB][2];
  if (warp >= kProdWarp0) {
    const int p = tid - kProdWarp0 * 32;
    if (p < to.max_feat) {
      pre_a[0] = __ldg(g_query + 2 * (size_t)p);
      pre_a[1] = __ldg(g_query + 2 * (size_t)p + 1);
    }
#pragma unroll
    for (int i = 0; i < kRowsB; ++i) {
      const int r = i * kHalf + p;
      if (r < from.max_feat) {
        pre_b[i][0] = __ldg(g_train + 2 * (size_t)r);
        pre_b[i][1] = __ldg(g_train + 2 * (size_t)r + 1);
      }
    }
  }

  // ---- setup -----------------------------------------------------------------------------------
  for (int i = tid; i < 4 * to.max_feat; i += kThreadsT) s_top[i] = kNoneKey;
  for (int i = tid; i < (kBxBytes + kAxBytes) / 16; i += kThreadsT)
    reinterpret_cast<uint4*>(sBx)[i] = make_uint4(0, 0, 0, 0);
  __syncthreads();
  for (int r = tid; r < CH; r += kThreadsT) {
    const int col = r & (kHalf - 1);
    uint8_t* p = sBx + (r >> 3) * 128 + (r & 7) * 16;
    p[0] = (uint8_t)(col >> 6);
    p[1] = (uint8_t)(col & 63);
  }
  for (int r = tid; r < TM; r += kThreadsT) {
    *reinterpret_cast<uint32_t*>(sAx + (r >> 3) * 256 + (r & 7) * 16) = 0x8080FFC0u;  // (-64, -1, -128, -128)
  }
  if (tid == 0) {
    for (int s = 0; s < 2; ++s) {
      umma::mbar_init(&bar_a_full[s], kProdThreads);
      umma::mbar_init(&bar_a_empty[s], 1);
    }
    for (int s = 0; s < kBStages; ++s) {
      umma::mbar_init(&bar_b_full[s], kProdThreads);
      umma::mbar_init(&bar_b_empty[s], 1);
    }
    for (int s = 0; s < kTStages; ++s) {
      umma::mbar_init(&bar_t_full[s], 1);
      umma::mbar_init(&bar_t_empty[s], 4 * kGroupsPerTile);  // one arrive per epilogue warp that drains the stage
    }
    umma::fence_mbar_init();
    s_dropped = 0;
  }
  if (warp == kMmaWarp) umma::tmem_alloc(&tmem_slot, 512);
  umma::fence_proxy_async();  // index K-blocks were written through the generic proxy
  umma::tc_fence_before();
  __syncthreads();
  umma::tc_fence_after();
  const uint32_t tmem = tmem_slot;
  K2T_STAMP(tid == 0, 1);
  if (nt == 0 || nf == 0) {  // block-uniform: nothing to multiply (checked after the setup, which does not need the
                             // feature counts and so overlaps the pairs -> counts round trips)
    for (int qi = tid; qi < nt; qi += kThreadsT) {
      const size_t o = ((size_t)pair * to.max_feat + qi) * 2;
      if (knn_idx) knn_idx[o] = knn_idx[o + 1] = -1;
      if (knn_dist) knn_dist[o] = knn_dist[o + 1] = -1;
    }
    if (tid == 0) {
      n_matches[pair] = 0;
      if (n_unique_to) n_unique_to[pair] = nt;
    }
    if (warp == kMmaWarp) umma::tmem_dealloc(tmem, 512);  // nothing was issued against it
    return;
  }
  const int n_chunks = (nf + CH - 1) / CH, n_mt = (nt + TM - 1) / TM;

  if (warp >= kProdWarp0) {
    // ---- producers ---------------------------------------------------------------------------
    const int p = tid - kProdWarp0 * 32;
    const uint32_t sB_a = umma::smem_u32(sB), sA_a = umma::smem_u32(sA);
    int it = 0;
    for (int c = 0; c < n_chunks; ++c) {
      const int rows = min(CH, nf - c * CH);
      const int rows_pad = (rows + 31) & ~31;  // the last tile is padded to 32 columns with "no neighbour" rows
      for (int m = 0; m < n_mt; ++m, ++it) {
        const int s = it & 1;
        K2T_STAMP(p == 0 && it < 32, 336 + it);
        // every row is requested before the wait for its buffer (the first ones before the setup)
        const int row = m * TM + p;
        uint4 qa, qb;
        if (it == 0) {
          qa = pre_a[0];
          qb = pre_a[1];
        } else if (row < nt) {
          qa = __ldg(g_query + 2 * (size_t)row);
          qb = __ldg(g_query + 2 * (size_t)row + 1);
        }
        umma::mbar_wait(&bar_a_empty[s], (uint32_t)(((it >> 1) & 1) ^ 1));
        K2T_STAMP(p == 0 && it < 32, 368 + it);
        if (row < nt) expand_row(sA_a + (uint32_t)(s * kABytes), TM, p, qa, qb);
        umma::fence_proxy_async();
        umma::mbar_arrive(&bar_a_full[s]);
        K2T_STAMP(p == 0 && it < 32, 400 + it);
        if (m == 0) {
          // resident From stages of this chunk, each as soon as the previous chunk's last tile on it retired
#pragma unroll
          for (int j = 0; j < kBStages; ++j) {
            if (j * TN >= rows) break;
            uint4 ta[kGroupsPerTile], tb[kGroupsPerTile];
#pragma unroll
            for (int h = 0; h < kGroupsPerTile; ++h) {
              const int r = j * TN + h * kHalf + p;
              if (c == 0) {
                ta[h] = pre_b[j * kGroupsPerTile + h][0];
                tb[h] = pre_b[j * kGroupsPerTile + h][1];
              } else if (r < rows) {
                ta[h] = __ldg(g_train + 2 * (size_t)(c * CH + r));
                tb[h] = __ldg(g_train + 2 * (size_t)(c * CH + r) + 1);
              }
            }
            umma::mbar_wait(&bar_b_empty[j], (uint32_t)((c & 1) ^ 1));
#pragma unroll
            for (int h = 0; h < kGroupsPerTile; ++h) {
              const int r = j * TN + h * kHalf + p;
              if (r < rows) {
                expand_row(sB_a, CH, r, ta[h], tb[h]);
              } else if (r < rows_pad) {
                zero_row(sB_a, CH, r);
                uint8_t* px = sBx + (r >> 3) * 128 + (r & 7) * 16;
                px[2] = 127;
                px[3] = 127;
              }
            }
            umma::fence_proxy_async();
            umma::mbar_arrive(&bar_b_full[j]);
            K2T_STAMP(p == 0 && c < 4, 432 + 4 * c + j);
          }
        }
      }
    }
  } else if (warp == kMmaWarp) {
    // ---- MMA issuer: the whole warp walks the pipeline, one elected lane issues -----------------
    // descriptors differ only in the 14-bit start-address field: precompute the low words once
    const uint32_t a_lo = ((umma::smem_u32(sA) & 0x3FFFFu) >> 4) | (1u << 16);
    const uint32_t b_lo = ((umma::smem_u32(sB) & 0x3FFFFu) >> 4) | (1u << 16);
    const uint32_t sw_hi = (uint32_t)(1024 >> 4) | (1u << 14) | (2u << 29);         // SBO, version, SWIZZLE_128B
    const uint32_t il_hi = (uint32_t)(256 >> 4) | (1u << 14);                        // SBO, version, no swizzle
    const uint32_t ilb_hi = (uint32_t)(128 >> 4) | (1u << 14);                       // From side: one core matrix per group
    const uint32_t ax_lo = ((umma::smem_u32(sAx) & 0x3FFFFu) >> 4) | ((128u >> 4) << 16);
    const uint32_t bx_lo = ((umma::smem_u32(sBx) & 0x3FFFFu) >> 4) | ((128u >> 4) << 16);
    auto desc = [](uint32_t lo, uint32_t hi) { return ((uint64_t)hi << 32) | lo; };
    int it = 0;
    uint32_t g = 0;  // accumulator tiles issued so far: TMEM stage g & 3, use count g >> 2
    for (int c = 0; c < n_chunks; ++c) {
      const int rows = min(CH, nf - c * CH);
      for (int m = 0; m < n_mt; ++m, ++it) {
        const int s = it & 1;
        K2T_STAMP(lane == 0 && it < 32, 464 + it);
        umma::mbar_wait(&bar_a_full[s], (uint32_t)((it >> 1) & 1));
        const uint32_t a_s = a_lo + (uint32_t)((s * kABytes) >> 4);
#pragma unroll
        for (int j = 0; j < kBStages; ++j) {
          const int ncols = min(TN, rows - j * TN);
          if (ncols <= 0) break;
          if (m == 0) umma::mbar_wait(&bar_b_full[j], (uint32_t)(c & 1));
          const uint32_t st = g & (kTStages - 1);
          K2T_STAMP(lane == 0 && g < 64, 16 + g);
          umma::mbar_wait(&bar_t_empty[st], ((g >> kTShift) & 1) ^ 1);
          umma::tc_fence_after();
          if (umma::elect_one()) {
            const uint32_t idesc = umma::instr_desc(2, 1, 1, TM, (uint32_t)((ncols + 31) & ~31));
            const uint32_t d = tmem + st * TN;
#pragma unroll
            for (int kb = 0; kb < 2; ++kb)
#pragma unroll
              for (int k = 0; k < 4; ++k)
                umma::mma_i8(d, desc(a_s + (uint32_t)((kb * TM * 128 + k * 32) >> 4), sw_hi),
                             desc(b_lo + (uint32_t)((kb * CH * 128 + j * TN * 128 + k * 32) >> 4), sw_hi), idesc,
                             (kb | k) != 0);
            umma::mma_i8(d, desc(ax_lo, il_hi), desc(bx_lo + (uint32_t)((j * TN * 16) >> 4), ilb_hi), idesc, true);
            umma::mma_commit(&bar_t_full[st]);
            if (m == n_mt - 1) umma::mma_commit(&bar_b_empty[j]);  // last To tile against this From stage
          }
          __syncwarp();
          K2T_STAMP(lane == 0 && g < 64, 80 + g);
          ++g;
        }
        if (umma::elect_one()) umma::mma_commit(&bar_a_empty[s]);
        __syncwarp();
      }
    }
  } else {
    // ---- epilogue ------------------------------------------------------------------------------
    const int grp = warp >> 2, lq = warp & 3;
    uint32_t g = 0;
    for (int c = 0; c < n_chunks; ++c) {
      const int rows = min(CH, nf - c * CH);
      for (int m = 0; m < n_mt; ++m) {
        for (int j = 0; j * TN < rows; ++j, ++g) {
          // warp-uniform: 128-column tiles alternate between the groups, 256-column tiles are split between them
          if (kGroupsPerTile == 1 && (int)(g & 1) != grp) continue;
          const int off = kGroupsPerTile == 1 ? 0 : grp * kHalf;
          const int ncols = max(0, min(kHalf, rows - j * TN - off));
          const int nch = (ncols + 31) >> 5;  // the producers padded the tile to a multiple of 32 columns
          const uint32_t st = g & (kTStages - 1);
          const uint32_t taddr = tmem + ((uint32_t)(lq * 32) << 16) + st * TN + off;
          K2T_STAMP(lq == 0 && lane == 0 && g < 64, 144 + g);
          umma::mbar_wait(&bar_t_full[st], (g >> kTShift) & 1);
          K2T_STAMP(lq == 0 && lane == 0 && g < 64, 208 + g);
          umma::tc_fence_after();
          uint32_t m1[2] = {kNone2, kNone2}, m2[2] = {kNone2, kNone2};
          // the group's whole half tile goes to registers at once (one TMEM round trip instead of four) and the
          // stage returns to the MMA warp before the scan, not after it
          uint32_t r0[16], r1[16], r2[16], r3[16];
          if (nch > 0) umma::tmem_ld32_pack16(taddr, r0);
          if (nch > 1) umma::tmem_ld32_pack16(taddr + 32u, r1);
          if (nch > 2) umma::tmem_ld32_pack16(taddr + 64u, r2);
          if (nch > 3) umma::tmem_ld32_pack16(taddr + 96u, r3);
          umma::tmem_ld_wait();
          umma::tc_fence_before();
          __syncwarp();
          if (lane == 0) umma::mbar_arrive(&bar_t_empty[st]);
          K2T_STAMP(lq == 0 && lane == 0 && g < 64, 272 + g);
          if (nch > 0) scan32(r0, m1, m2);
          if (nch > 1) scan32(r1, m1, m2);
          if (nch > 2) scan32(r2, m1, m2);
          if (nch > 3) scan32(r3, m1, m2);
          merge2(m1[0], m2[0], m1[1], m2[1]);
          // unpack: even columns in the low halves, odd columns in the high halves
          int v1 = (int)(int16_t)(m1[0] & 0xFFFFu), v2 = (int)(int16_t)(m2[0] & 0xFFFFu);
          {
            const int w1 = (int)(int16_t)(m1[0] >> 16), w2 = (int)(int16_t)(m2[0] >> 16);
            const int lo = min(v1, w1);
            v1 = max(v1, w1);
            v2 = max(lo, max(v2, w2));
          }
          const int row = m * TM + lq * 32 + lane;
          if (row < nt) {
            uint32_t* slot = s_top + ((size_t)grp * to.max_feat + row) * 2;
            uint32_t k1 = slot[0], k2 = slot[1];
            merge_min(k1, k2, acc_to_key(v1, c * CH + j * TN + off), acc_to_key(v2, c * CH + j * TN + off));
            slot[0] = k1;
            slot[1] = k2;
          }
        }
      }
    }
  }

  // ---- tail: NNDR, uniqueness, correspondence list (all warps) --------------------------------------
  umma::tc_fence_before();
  __syncthreads();
  K2T_STAMP(tid == 0, 2);
  if (warp == kMmaWarp) {
    umma::tc_fence_after();
    umma::tmem_dealloc(tmem, 512);
  }
  for (int i = tid; i < nf; i += kThreadsT) s_count[i] = 0;
  int any_assigned = 0;
  for (int qi = tid; qi < nt; qi += kThreadsT) {
    uint32_t k1 = s_top[(size_t)qi * 2], k2 = s_top[(size_t)qi * 2 + 1];
    merge_min(k1, k2, s_top[((size_t)to.max_feat + qi) * 2], s_top[((size_t)to.max_feat + qi) * 2 + 1]);
    const int i1 = k1 == kNoneKey ? -1 : (int)(k1 & 0xFFFFu), d1 = k1 == kNoneKey ? -1 : (int)(k1 >> 16);
    const int i2 = k2 == kNoneKey ? -1 : (int)(k2 & 0xFFFFu), d2 = k2 == kNoneKey ? -1 : (int)(k2 >> 16);
    const size_t o = ((size_t)pair * to.max_feat + qi) * 2;
    if (knn_idx) {
      knn_idx[o] = i1;
      knn_idx[o + 1] = i2;
    }
    if (knn_dist) {
      knn_dist[o] = d1;
      knn_dist[o + 1] = d2;
    }
    const int as = mtail::nndr_assign(i1, d1, i2, d2, nndr, strict);
    s_assigned[qi] = as;
    any_assigned |= as >= 0;
  }
  if (!__syncthreads_or(any_assigned)) {  // no To row passed the ratio test (most non-matching pairs): empty list
    if (tid == 0) {
      n_matches[pair] = 0;
      if (n_unique_to) n_unique_to[pair] = nt;
    }
    K2T_STAMP(tid == 0, 3);
    return;
  }
  const float* p3 = from.kp3d_early(fi);
  const int total = mtail::unique_and_compact<kThreadsT>(tid, nf, nt, p3, s_assigned, s_count, s_owner, s_scan, &s_dropped,
                                                         (size_t)pair * from.max_feat, pair, n_matches, match_from, match_to,
                                                         n_unique_to);
  // host-bound stores: this pair will run PnP -> pull its From 3-D points and To pixels now, while the other CTAs are
  // still matching (the first CTA that wants a row copies it; K3 is a later kernel, so every copy has landed by then)
  if (lp.min_matches > 0 && total >= lp.min_matches) {
    const unsigned rf = (unsigned)(fi - lp.from_first), rt = (unsigned)(ti - lp.to_first);
    if (tid == 0) {
      s_claim[0] = lp.from_kp3d_host && rf < (unsigned)lp.from_n && sepb_claim_row(lp.from_have, fi, 2u);
      s_claim[1] = lp.to_kp2d_host && rt < (unsigned)lp.to_n && sepb_claim_row(lp.to_have, ti, 1u);
    }
    __syncthreads();
    const int bf = s_claim[0] ? nf * 12 : 0, bt = s_claim[1] ? nt * 8 : 0;
    if (bf | bt) {
      sepb_pull_rows<5>(lp.from_kp3d + (size_t)fi * from.max_feat * 3, bf ? lp.from_kp3d_host + (size_t)rf * from.max_feat * 3 : nullptr,
                     bf, lp.to_kp2d + (size_t)ti * to.max_feat * 2, bt ? lp.to_kp2d_host + (size_t)rt * to.max_feat * 2 : nullptr, bt,
                     tid, kThreadsT);
      if (tid == 0 && bf) atomicAdd(lp.from_fetched, (unsigned long long)bf);
      if (tid == 0 && bt) atomicAdd(lp.to_fetched, (unsigned long long)bt);
    }
  }
  K2T_STAMP(tid == 0, 3);
}


// ---- persistent form ---------------------------------------------------------------------------------------------
// One CTA per SM walks the pairs blockIdx.x, blockIdx.x + gridDim.x, ...; the three roles run through the SAME pair
// sequence on their own, coupled only by the operand / accumulator rings (whose phases simply keep counting across
// pairs).  What the one-CTA-per-pair kernel pays per pair in the open -- setup (TMEM allocation, barrier and index-block
// initialisation), the expansion of the first To tile and From stage before the first MMA, and the tail after the last
// one (15 % of a 1000 x 1000 pair) -- overlaps here: while the eight epilogue warps run pair p's tail (ratio test,
// uniqueness, compaction: on 16-bit scratch of its own, the operand tiles belong to the next pair by then), the
// producers have expanded pair p + 1's first tiles and the MMA warp fills both TMEM stages with them.
__device__ __forceinline__ void bar_sync(int id, int nthreads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}
constexpr int kEpiThreads = kEpiWarps * 32;   // the tail's 256 threads (named barrier 1)
// Host-bound stores (LazyPull): three more warps of the persistent CTA copy the key-point rows of the pairs that will run
// PnP out of mapped host memory.  A load over PCIe takes microseconds on a quiet bus and ~100 us while the next step's bulk
// upload saturates it -- several pairs' worth of matching -- so the epilogue only posts (row, length) requests to a small
// ring in shared memory and goes on; the pullers drain it (12 loads of 16 bytes per lane in flight: 18 KB per SM) and finish what is
// left when the pair loop ends.  K3 is a later kernel: every copy has landed by then.  A full ring falls back to the
// inline copy by the epilogue's 256 threads.
constexpr int kPullWarps = 3;
constexpr int kPullWarp0 = kProdWarp0 + kProdWarps;
constexpr int kThreadsP = kThreadsT + kPullWarps * 32;   // 512
constexpr int kPullQ = 24;                               // ring entries (8 bytes each; static shared memory must stay <= 512 B)
static_assert(kPullQ % kPullWarps == 0, "request j belongs to puller j % kPullWarps, also after the ring wraps");
__device__ __forceinline__ int ld_vol(const int* p) { return *reinterpret_cast<const volatile int*>(p); }
__device__ __forceinline__ void st_vol(int* p, int v) { *reinterpret_cast<volatile int*>(p) = v; }

struct PairGeom {
  int fi, ti, nf, nt;
};
__device__ __forceinline__ PairGeom pair_geom(const FramesView& from, const FramesView& to, const int32_t* pairs, int pair) {
  PairGeom g;
  g.fi = pairs[2 * pair];
  g.ti = pairs[2 * pair + 1];
  const bool valid = g.fi >= 0 && g.fi < from.size && g.ti >= 0 && g.ti < to.size;
  if (!valid) g.fi = g.ti = 0;
  g.nf = valid ? min(from.count[g.fi], from.max_feat) : 0;
  g.nt = valid ? min(to.count[g.ti], to.max_feat) : 0;
  return g;
}
template <int TN>
__global__ void __launch_bounds__(kThreadsP, 1)
hamming_mma_persist_kernel(FramesView from, FramesView to, const int32_t* __restrict__ pairs, int n_pairs, float nndr, int strict,
                           int32_t* __restrict__ knn_idx, int32_t* __restrict__ knn_dist, int32_t* __restrict__ n_matches,
                           uint16_t* __restrict__ match_from, uint16_t* __restrict__ match_to,
                           int32_t* __restrict__ n_unique_to, const LazyPull lp) {
  using G = Geo<TN>;
  constexpr int kGroupsPerTile = G::kGroupsPerTile, kBStages = G::kBStages, CH = G::CH, kTStages = G::kTStages,
                kTShift = G::kTShift, kBBytes = G::kBBytes, kBxBytes = G::kBxBytes;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (umma::smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t* sB = smem;                                  // [2 K-blocks][CH][128]
  uint8_t* sA = sB + kBBytes;                          // [2 buffers][2 K-blocks][TM][128]
  uint8_t* sBx = sA + 2 * kABytes;                     // [CH/8][2][8][16]
  uint8_t* sAx = sBx + kBxBytes;                       // [TM/8][2][8][16]
  uint32_t* s_top = reinterpret_cast<uint32_t*>(sAx + kAxBytes);  // [2][to.max_feat][2] keys
  // the tail's own scratch (16-bit: To rows and From rows are < 65536)
  int16_t* s_asg = reinterpret_cast<int16_t*>(s_top + 4 * (size_t)to.max_feat);   // [to.max_feat] From row or -1
  uint16_t* s_cnt = reinterpret_cast<uint16_t*>(s_asg + ((to.max_feat + 1) & ~1));  // [from.max_feat] uses of a From row
  uint16_t* s_own = s_cnt + ((from.max_feat + 1) & ~1);                            // [from.max_feat] its To row
  __shared__ uint64_t bar_a_full[2], bar_a_empty[2], bar_b_full[kBStages], bar_b_empty[kBStages], bar_t_full[kTStages],
      bar_t_empty[kTStages];
  __shared__ uint32_t tmem_slot;
  __shared__ int s_scan[kEpiWarps];
  __shared__ int s_dropped, s_any;
  __shared__ int s_claim[2];
  __shared__ int2 s_pq[kPullQ];              // pull requests: {row kind << 31 | slot, features}
  __shared__ int s_phead, s_pdone, s_ptail[kPullWarps];

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  // ---- setup, once per CTA ---------------------------------------------------------------------------
  for (int i = tid; i < 4 * to.max_feat; i += kThreadsP) s_top[i] = kNoneKey;
  for (int i = tid; i < (kBxBytes + kAxBytes) / 16; i += kThreadsP)
    reinterpret_cast<uint4*>(sBx)[i] = make_uint4(0, 0, 0, 0);
  __syncthreads();
  for (int r = tid; r < CH; r += kThreadsP) {
    const int col = r & (kHalf - 1);
    uint8_t* p = sBx + (r >> 3) * 128 + (r & 7) * 16;
    p[0] = (uint8_t)(col >> 6);
    p[1] = (uint8_t)(col & 63);
  }
  for (int r = tid; r < TM; r += kThreadsP)
    *reinterpret_cast<uint32_t*>(sAx + (r >> 3) * 256 + (r & 7) * 16) = 0x8080FFC0u;  // (-64, -1, -128, -128)
  if (tid == 0) {
    for (int s = 0; s < 2; ++s) {
      umma::mbar_init(&bar_a_full[s], kProdThreads);
      umma::mbar_init(&bar_a_empty[s], 1);
    }
    for (int s = 0; s < kBStages; ++s) {
      umma::mbar_init(&bar_b_full[s], kProdThreads);
      umma::mbar_init(&bar_b_empty[s], 1);
    }
    for (int s = 0; s < kTStages; ++s) {
      umma::mbar_init(&bar_t_full[s], 1);
      umma::mbar_init(&bar_t_empty[s], 4 * kGroupsPerTile);
    }
    umma::fence_mbar_init();
    s_dropped = 0;
    s_any = 0;
    s_phead = 0;
    s_pdone = 0;
    for (int w = 0; w < kPullWarps; ++w) s_ptail[w] = w;   // next request puller w consumes
  }
  if (warp == kMmaWarp) umma::tmem_alloc(&tmem_slot, 512);
  umma::fence_proxy_async();
  umma::tc_fence_before();
  __syncthreads();
  umma::tc_fence_after();
  const uint32_t tmem = tmem_slot;

  if (warp >= kPullWarp0) {
    // ---- pullers: copy the requested rows out of mapped host memory --------------------------------------------
    const int w = warp - kPullWarp0;
    for (int next = w;; next += kPullWarps) {
      bool have = false;
      while (true) {
        if (ld_vol(&s_phead) > next) { have = true; break; }
        if (ld_vol(&s_pdone)) { have = ld_vol(&s_phead) > next; break; }
        __nanosleep(256);
      }
      if (!have) break;
      __threadfence_block();
      const int2 rq = s_pq[next % kPullQ];
      const int slot = rq.x & 0x7fffffff, n = rq.y;
      if (rq.x < 0) {   // To pixels
        const unsigned rel = (unsigned)(slot - lp.to_first);
        sepb_pull_rows<12>(lp.to_kp2d + (size_t)slot * to.max_feat * 2, lp.to_kp2d_host + (size_t)rel * to.max_feat * 2, n * 8,
                           nullptr, nullptr, 0, lane, 32);
        if (lane == 0) atomicAdd(lp.to_fetched, (unsigned long long)n * 8);
      } else {          // From 3-D points
        const unsigned rel = (unsigned)(slot - lp.from_first);
        sepb_pull_rows<12>(lp.from_kp3d + (size_t)slot * from.max_feat * 3, lp.from_kp3d_host + (size_t)rel * from.max_feat * 3,
                           n * 12, nullptr, nullptr, 0, lane, 32);
        if (lane == 0) atomicAdd(lp.from_fetched, (unsigned long long)n * 12);
      }
      __syncwarp();
      if (lane == 0) st_vol(&s_ptail[w], next + kPullWarps);   // the ring entry may be reused
    }
  } else if (warp >= kProdWarp0) {
    // ---- producers: run ahead through the pair sequence, held back only by the operand rings --------------------
    const int p = tid - kProdWarp0 * 32;
    const uint32_t sB_a = umma::smem_u32(sB), sA_a = umma::smem_u32(sA);
    int it = 0;                 // To tiles expanded so far (all pairs): buffer it & 1, use count it >> 1
    int ub[kBStages];           // fills of From stage j so far
#pragma unroll
    for (int j = 0; j < kBStages; ++j) ub[j] = 0;
    for (int pair = blockIdx.x; pair < n_pairs; pair += gridDim.x) {
      const PairGeom pg = pair_geom(from, to, pairs, pair);
      const int nf = pg.nf, nt = pg.nt;
      if (nt == 0 || nf == 0) continue;
      const uint4* g_train = reinterpret_cast<const uint4*>(from.desc + (size_t)pg.fi * from.max_feat * 32);
      const uint4* g_query = reinterpret_cast<const uint4*>(to.desc + (size_t)pg.ti * to.max_feat * 32);
      const int n_chunks = (nf + CH - 1) / CH, n_mt = (nt + TM - 1) / TM;
      for (int c = 0; c < n_chunks; ++c) {
        const int rows = min(CH, nf - c * CH);
        const int rows_pad = (rows + 31) & ~31;
        for (int m = 0; m < n_mt; ++m, ++it) {
          const int s = it & 1;
          const int row = m * TM + p;
          uint4 qa, qb;
          if (row < nt) {
            qa = __ldg(g_query + 2 * (size_t)row);
            qb = __ldg(g_query + 2 * (size_t)row + 1);
          }
          umma::mbar_wait(&bar_a_empty[s], (uint32_t)(((it >> 1) & 1) ^ 1));
          if (row < nt) expand_row(sA_a + (uint32_t)(s * kABytes), TM, p, qa, qb);
          umma::fence_proxy_async();
          umma::mbar_arrive(&bar_a_full[s]);
          if (m == 0) {
#pragma unroll
            for (int j = 0; j < kBStages; ++j) {
              if (j * TN >= rows) break;
              uint4 ta[kGroupsPerTile], tb[kGroupsPerTile];
#pragma unroll
              for (int h = 0; h < kGroupsPerTile; ++h) {
                const int r = j * TN + h * kHalf + p;
                if (r < rows) {
                  ta[h] = __ldg(g_train + 2 * (size_t)(c * CH + r));
                  tb[h] = __ldg(g_train + 2 * (size_t)(c * CH + r) + 1);
                }
              }
              umma::mbar_wait(&bar_b_empty[j], (uint32_t)((ub[j] & 1) ^ 1));
#pragma unroll
              for (int h = 0; h < kGroupsPerTile; ++h) {
                const int r = j * TN + h * kHalf + p;
                uint8_t* px = sBx + (r >> 3) * 128 + (r & 7) * 16;
                if (r < rows) {
                  expand_row(sB_a, CH, r, ta[h], tb[h]);
                  px[2] = 0;      // a row of an earlier pair may have left the padding code here
                  px[3] = 0;
                } else if (r < rows_pad) {
                  zero_row(sB_a, CH, r);
                  px[2] = 127;
                  px[3] = 127;
                }
              }
              umma::fence_proxy_async();
              umma::mbar_arrive(&bar_b_full[j]);
              ++ub[j];
            }
          }
        }
      }
    }
  } else if (warp == kMmaWarp) {
    // ---- MMA issuer --------------------------------------------------------------------------------------------
    const uint32_t a_lo = ((umma::smem_u32(sA) & 0x3FFFFu) >> 4) | (1u << 16);
    const uint32_t b_lo = ((umma::smem_u32(sB) & 0x3FFFFu) >> 4) | (1u << 16);
    const uint32_t sw_hi = (uint32_t)(1024 >> 4) | (1u << 14) | (2u << 29);
    const uint32_t il_hi = (uint32_t)(256 >> 4) | (1u << 14);
    const uint32_t ilb_hi = (uint32_t)(128 >> 4) | (1u << 14);
    const uint32_t ax_lo = ((umma::smem_u32(sAx) & 0x3FFFFu) >> 4) | ((128u >> 4) << 16);
    const uint32_t bx_lo = ((umma::smem_u32(sBx) & 0x3FFFFu) >> 4) | ((128u >> 4) << 16);
    auto desc = [](uint32_t lo, uint32_t hi) { return ((uint64_t)hi << 32) | lo; };
    int it = 0;
    uint32_t g = 0;
    int ub[kBStages];
#pragma unroll
    for (int j = 0; j < kBStages; ++j) ub[j] = 0;
    for (int pair = blockIdx.x; pair < n_pairs; pair += gridDim.x) {
      const PairGeom pg = pair_geom(from, to, pairs, pair);
      const int nf = pg.nf, nt = pg.nt;
      if (nt == 0 || nf == 0) continue;
      const int n_chunks = (nf + CH - 1) / CH, n_mt = (nt + TM - 1) / TM;
      for (int c = 0; c < n_chunks; ++c) {
        const int rows = min(CH, nf - c * CH);
        for (int m = 0; m < n_mt; ++m, ++it) {
          const int s = it & 1;
          umma::mbar_wait(&bar_a_full[s], (uint32_t)((it >> 1) & 1));
          const uint32_t a_s = a_lo + (uint32_t)((s * kABytes) >> 4);
#pragma unroll
          for (int j = 0; j < kBStages; ++j) {
            const int ncols = min(TN, rows - j * TN);
            if (ncols <= 0) break;
            if (m == 0) umma::mbar_wait(&bar_b_full[j], (uint32_t)(ub[j] & 1));
            const uint32_t st = g & (kTStages - 1);
            umma::mbar_wait(&bar_t_empty[st], ((g >> kTShift) & 1) ^ 1);
            umma::tc_fence_after();
            if (umma::elect_one()) {
              const uint32_t idesc = umma::instr_desc(2, 1, 1, TM, (uint32_t)((ncols + 31) & ~31));
              const uint32_t d = tmem + st * TN;
#pragma unroll
              for (int kb = 0; kb < 2; ++kb)
#pragma unroll
                for (int k = 0; k < 4; ++k)
                  umma::mma_i8(d, desc(a_s + (uint32_t)((kb * TM * 128 + k * 32) >> 4), sw_hi),
                               desc(b_lo + (uint32_t)((kb * CH * 128 + j * TN * 128 + k * 32) >> 4), sw_hi), idesc,
                               (kb | k) != 0);
              umma::mma_i8(d, desc(ax_lo, il_hi), desc(bx_lo + (uint32_t)((j * TN * 16) >> 4), ilb_hi), idesc, true);
              umma::mma_commit(&bar_t_full[st]);
              if (m == n_mt - 1) umma::mma_commit(&bar_b_empty[j]);
            }
            __syncwarp();
            ++g;
          }
          if (umma::elect_one()) umma::mma_commit(&bar_a_empty[s]);
          __syncwarp();
        }
#pragma unroll
        for (int j = 0; j < kBStages; ++j)
          if (j * TN < rows) ++ub[j];
      }
    }
  } else {
    // ---- epilogue warps: drain the accumulator tiles of a pair, then run its tail -------------------------------
    const int grp = warp >> 2, lq = warp & 3;
    uint32_t g = 0;
    for (int pair = blockIdx.x; pair < n_pairs; pair += gridDim.x) {
      const PairGeom pg = pair_geom(from, to, pairs, pair);
      const int nf = pg.nf, nt = pg.nt, fi = pg.fi, ti = pg.ti;
      if (nt == 0 || nf == 0) {   // nothing to multiply (the other roles skip the pair as well)
        for (int qi = tid; qi < nt; qi += kEpiThreads) {
          const size_t o = ((size_t)pair * to.max_feat + qi) * 2;
          if (knn_idx) knn_idx[o] = knn_idx[o + 1] = -1;
          if (knn_dist) knn_dist[o] = knn_dist[o + 1] = -1;
        }
        if (tid == 0) {
          n_matches[pair] = 0;
          if (n_unique_to) n_unique_to[pair] = nt;
        }
        continue;
      }
      const int n_chunks = (nf + CH - 1) / CH, n_mt = (nt + TM - 1) / TM;
      for (int c = 0; c < n_chunks; ++c) {
        const int rows = min(CH, nf - c * CH);
        for (int m = 0; m < n_mt; ++m) {
          for (int j = 0; j * TN < rows; ++j, ++g) {
            if (kGroupsPerTile == 1 && (int)(g & 1) != grp) continue;
            const int off = kGroupsPerTile == 1 ? 0 : grp * kHalf;
            const int ncols = max(0, min(kHalf, rows - j * TN - off));
            const int nch = (ncols + 31) >> 5;
            const uint32_t st = g & (kTStages - 1);
            const uint32_t taddr = tmem + ((uint32_t)(lq * 32) << 16) + st * TN + off;
            umma::mbar_wait(&bar_t_full[st], (g >> kTShift) & 1);
            umma::tc_fence_after();
            uint32_t m1[2] = {kNone2, kNone2}, m2[2] = {kNone2, kNone2};
            uint32_t r0[16], r1[16], r2[16], r3[16];
            if (nch > 0) umma::tmem_ld32_pack16(taddr, r0);
            if (nch > 1) umma::tmem_ld32_pack16(taddr + 32u, r1);
            if (nch > 2) umma::tmem_ld32_pack16(taddr + 64u, r2);
            if (nch > 3) umma::tmem_ld32_pack16(taddr + 96u, r3);
            umma::tmem_ld_wait();
            umma::tc_fence_before();
            __syncwarp();
            if (lane == 0) umma::mbar_arrive(&bar_t_empty[st]);
            if (nch > 0) scan32(r0, m1, m2);
            if (nch > 1) scan32(r1, m1, m2);
            if (nch > 2) scan32(r2, m1, m2);
            if (nch > 3) scan32(r3, m1, m2);
            merge2(m1[0], m2[0], m1[1], m2[1]);
            int v1 = (int)(int16_t)(m1[0] & 0xFFFFu), v2 = (int)(int16_t)(m2[0] & 0xFFFFu);
            {
              const int w1 = (int)(int16_t)(m1[0] >> 16), w2 = (int)(int16_t)(m2[0] >> 16);
              const int lo = min(v1, w1);
              v1 = max(v1, w1);
              v2 = max(lo, max(v2, w2));
            }
            const int row = m * TM + lq * 32 + lane;
            if (row < nt) {
              uint32_t* slot = s_top + ((size_t)grp * to.max_feat + row) * 2;
              uint32_t k1 = slot[0], k2 = slot[1];
              merge_min(k1, k2, acc_to_key(v1, c * CH + j * TN + off), acc_to_key(v2, c * CH + j * TN + off));
              slot[0] = k1;
              slot[1] = k2;
            }
          }
        }
      }
      // ---- tail of this pair, on the 256 epilogue threads (the other warps are already in the next pair) -----------
      bar_sync(1, kEpiThreads);
      for (int i = tid; i < (nf + 1) / 2; i += kEpiThreads) reinterpret_cast<uint32_t*>(s_cnt)[i] = 0;
      int any_assigned = 0;
      for (int qi = tid; qi < nt; qi += kEpiThreads) {
        uint32_t* t0 = s_top + (size_t)qi * 2;
        uint32_t* t1 = s_top + ((size_t)to.max_feat + qi) * 2;
        uint32_t k1 = t0[0], k2 = t0[1];
        merge_min(k1, k2, t1[0], t1[1]);
        t0[0] = t0[1] = t1[0] = t1[1] = kNoneKey;   // ready for the next pair
        const int i1 = k1 == kNoneKey ? -1 : (int)(k1 & 0xFFFFu), d1 = k1 == kNoneKey ? -1 : (int)(k1 >> 16);
        const int i2 = k2 == kNoneKey ? -1 : (int)(k2 & 0xFFFFu), d2 = k2 == kNoneKey ? -1 : (int)(k2 >> 16);
        const size_t o = ((size_t)pair * to.max_feat + qi) * 2;
        if (knn_idx) {
          knn_idx[o] = i1;
          knn_idx[o + 1] = i2;
        }
        if (knn_dist) {
          knn_dist[o] = d1;
          knn_dist[o + 1] = d2;
        }
        const int as = mtail::nndr_assign(i1, d1, i2, d2, nndr, strict);
        s_asg[qi] = (int16_t)as;
        any_assigned |= as >= 0;
      }
      if (__any_sync(0xFFFFFFFFu, any_assigned) && lane == 0) atomicOr(&s_any, 1);
      bar_sync(1, kEpiThreads);
      if (!s_any) {   // no To row passed the ratio test (most non-matching pairs): empty list
        if (tid == 0) {
          n_matches[pair] = 0;
          if (n_unique_to) n_unique_to[pair] = nt;
        }
        continue;
      }
      // To rows claiming the same From row are dropped (:882-894); 16-bit counters, two per word
      for (int qi = tid; qi < nt; qi += kEpiThreads) {
        const int a = s_asg[qi];
        if (a >= 0) {
          atomicAdd(reinterpret_cast<unsigned int*>(s_cnt) + (a >> 1), 1u << (16 * (a & 1)));
          s_own[a] = (uint16_t)qi;   // only read back where the count is 1 (single writer)
        }
      }
      bar_sync(1, kEpiThreads);
      const float* p3 = from.kp3d_early(fi);
      const int chunk = (nf + kEpiThreads - 1) / kEpiThreads;
      const int lo = min(nf, tid * chunk), hi = min(nf, lo + chunk);
      int local = 0, dropped = 0;
      uint32_t keep_mask = 0;  // chunk <= 32
      for (int i = lo; i < hi; ++i) {
        const uint32_t cnt = s_cnt[i];
        if (cnt >= 2) dropped += (int)cnt;
        if (cnt == 1) {
          const float x = p3[3 * i], y = p3[3 * i + 1], z = p3[3 * i + 2];
          if (isfinite(x) && isfinite(y) && isfinite(z)) {
            keep_mask |= 1u << (i - lo);
            ++local;
          }
        }
      }
      if (dropped) atomicAdd(&s_dropped, dropped);
      int incl = local;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int v = __shfl_up_sync(0xFFFFFFFFu, incl, o);
        if (lane >= o) incl += v;
      }
      if (lane == 31) s_scan[warp] = incl;
      bar_sync(1, kEpiThreads);
      int base = 0, total = 0;
#pragma unroll
      for (int w = 0; w < kEpiWarps; ++w) {
        const int v = s_scan[w];
        if (w < warp) base += v;
        total += v;
      }
      const size_t obase = (size_t)pair * from.max_feat;
      int pos = base + incl - local;
      for (int i = lo; i < hi; ++i) {
        if (keep_mask & (1u << (i - lo))) {
          if (match_from) match_from[obase + pos] = (uint16_t)i;
          if (match_to) match_to[obase + pos] = s_own[i];
          ++pos;
        }
      }
      const bool pull = lp.min_matches > 0 && total >= lp.min_matches;
      const unsigned rf = (unsigned)(fi - lp.from_first), rt = (unsigned)(ti - lp.to_first);
      if (tid == 0) {
        n_matches[pair] = total;
        if (n_unique_to) n_unique_to[pair] = nt - s_dropped;
        s_dropped = 0;
        s_any = 0;
        if (pull) {
          // rows this CTA is first to want: hand them to the pullers; s_claim keeps what the ring could not take
          bool cf = lp.from_kp3d_host && rf < (unsigned)lp.from_n && sepb_claim_row(lp.from_have, fi, 2u);
          bool ct = lp.to_kp2d_host && rt < (unsigned)lp.to_n && sepb_claim_row(lp.to_have, ti, 1u);
          int head = s_phead;
          auto post = [&](int x, int n) {
            // entry head % kPullQ last held request head - kPullQ, which puller head % kPullWarps consumes
            if (ld_vol(&s_ptail[head % kPullWarps]) <= head - kPullQ) return false;   // ring full: copy inline
            s_pq[head % kPullQ] = make_int2(x, n);
            __threadfence_block();
            st_vol(&s_phead, ++head);
            return true;
          };
          if (cf && post(fi, nf)) cf = false;
          if (ct && post(ti | (int)0x80000000, nt)) ct = false;
          s_claim[0] = cf;
          s_claim[1] = ct;
        }
      }
      bar_sync(1, kEpiThreads);   // also: s_scan / s_any / s_dropped are free again
      if (pull) {
        // both rows in one batch of loads: one PCIe round trip on the epilogue's critical path, not five
        const int bf = s_claim[0] ? nf * 12 : 0, bt = s_claim[1] ? nt * 8 : 0;
        if (bf | bt) {
          sepb_pull_rows<5>(lp.from_kp3d + (size_t)fi * from.max_feat * 3,
                         bf ? lp.from_kp3d_host + (size_t)rf * from.max_feat * 3 : nullptr, bf,
                         lp.to_kp2d + (size_t)ti * to.max_feat * 2,
                         bt ? lp.to_kp2d_host + (size_t)rt * to.max_feat * 2 : nullptr, bt, tid, kEpiThreads);
          if (tid == 0 && bf) atomicAdd(lp.from_fetched, (unsigned long long)bf);
          if (tid == 0 && bt) atomicAdd(lp.to_fetched, (unsigned long long)bt);
        }
        bar_sync(1, kEpiThreads);   // s_claim is read before the next pair's tail may rewrite it
      }
    }
    if (tid == 0) {
      __threadfence_block();
      st_vol(&s_pdone, 1);   // the pullers finish the requests still in the ring, then join the barrier below
    }
  }
  umma::tc_fence_before();
  __syncthreads();
  if (warp == kMmaWarp) {
    umma::tc_fence_after();
    umma::tmem_dealloc(tmem, 512);
  }
}

}  // namespace

#ifdef SEPB_K2T_TRACE
// developer build only (not in include/sepb200.h): copy the timeline out
extern "C" __attribute__((visibility("default"))) int sepb_debug_k2t_trace(long long* out, int n) {
  if (n > 512) n = 512;
  return (int)cudaMemcpyFromSymbol(out, g_k2t_trace, (size_t)n * sizeof(long long));
}
#endif

// shared-memory budget of one CTA decides the geometry: wide tiles when the To frames are short enough
template <int TN>
static bool geo_fits(const FramesView& from, const FramesView& to) {
  const size_t tail = (size_t)to.max_feat * 4 + (size_t)from.max_feat * 6;
  return Geo<TN>::smem(to.max_feat) + 256 <= 232448 && tail <= (size_t)Geo<TN>::kBBytes && from.max_feat <= 65535;
}

bool sepb_hamming_mma_supported(const FramesView& from, const FramesView& to) { return geo_fits<128>(from, to); }

// the persistent form needs the tail's own scratch on top of the operand tiles
template <int TN>
static size_t persist_smem(const FramesView& from, const FramesView& to) {
  return Geo<TN>::smem(to.max_feat) + 2 * (size_t)((to.max_feat + 1) & ~1) + 4 * (size_t)((from.max_feat + 1) & ~1) + 16;
}
template <int TN>
static bool persist_fits(const FramesView& from, const FramesView& to) {
  return geo_fits<TN>(from, to) && persist_smem<TN>(from, to) + 512 <= 232448;
}

int sepb_launch_hamming_mma(sepb_ctx* ctx, FramesView from, FramesView to, const int32_t* d_pairs, int n_pairs,
                            float nndr, int strict, int32_t* d_knn_idx, int32_t* d_knn_dist, int32_t* d_n_matches,
                            uint16_t* d_match_from, uint16_t* d_match_to, int32_t* d_n_unique_to, const LazyPull* lazy) {
  if (n_pairs == 0) return SEPB_OK;
  const LazyPull lp = lazy ? *lazy : LazyPull{};
  if (!sepb_hamming_mma_supported(from, to))
    return sepb_fail(ctx, SEPB_E_SHAPE, "hamming_mma: frame too large for the tensor kernel", __FILE__, __LINE__);
  bool wide = geo_fits<256>(from, to);
  if (ctx->opt_match_tile == 128) wide = false;   // SEPB_OPT_MATCH_TILE: the narrow geometry on request (tests, A/B)
  // SEPB_OPT_MATCH_SCHEDULER: persistent CTAs (default) whenever the tail's scratch fits beside the operand tiles
  bool persist = ctx->opt_match_sched == SEPB_MATCH_SCHED_PERSISTENT &&
                 (wide ? persist_fits<256>(from, to) : persist_fits<128>(from, to));
#ifdef SEPB_K2T_TRACE
  persist = false;   // the timeline build instruments the one-CTA-per-pair kernel
#endif
  if (persist) {
    const size_t smem = wide ? persist_smem<256>(from, to) : persist_smem<128>(from, to);
    size_t& cfg = wide ? ctx->cfg_hamming_mma_psmem_wide : ctx->cfg_hamming_mma_psmem;
    if (smem > cfg) {
      SEPB_CUDA(ctx, wide ? cudaFuncSetAttribute(hamming_mma_persist_kernel<256>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)
                          : cudaFuncSetAttribute(hamming_mma_persist_kernel<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      cfg = smem;
    }
    const int ctas = n_pairs < ctx->sm_count ? n_pairs : ctx->sm_count;
    if (wide)
      hamming_mma_persist_kernel<256><<<ctas, kThreadsP, smem, ctx->stream>>>(from, to, d_pairs, n_pairs, nndr, strict, d_knn_idx,
                                                                               d_knn_dist, d_n_matches, d_match_from, d_match_to,
                                                                               d_n_unique_to, lp);
    else
      hamming_mma_persist_kernel<128><<<ctas, kThreadsP, smem, ctx->stream>>>(from, to, d_pairs, n_pairs, nndr, strict, d_knn_idx,
                                                                               d_knn_dist, d_n_matches, d_match_from, d_match_to,
                                                                               d_n_unique_to, lp);
    ctx->launches++;
    SEPB_CUDA(ctx, cudaGetLastError());
    return SEPB_OK;
  }
  const size_t smem = wide ? Geo<256>::smem(to.max_feat) : Geo<128>::smem(to.max_feat);
  size_t& cfg = wide ? ctx->cfg_hamming_mma_smem_wide : ctx->cfg_hamming_mma_smem;
  if (smem > cfg) {
    SEPB_CUDA(ctx, wide ? cudaFuncSetAttribute(hamming_mma_kernel<256>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)
                        : cudaFuncSetAttribute(hamming_mma_kernel<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    cfg = smem;
  }
  if (wide)
    hamming_mma_kernel<256><<<n_pairs, kThreadsT, smem, ctx->stream>>>(from, to, d_pairs, n_pairs, nndr, strict, d_knn_idx,
                                                                      d_knn_dist, d_n_matches, d_match_from, d_match_to,
                                                                      d_n_unique_to, lp);
  else
    hamming_mma_kernel<128><<<n_pairs, kThreadsT, smem, ctx->stream>>>(from, to, d_pairs, n_pairs, nndr, strict, d_knn_idx,
                                                                      d_knn_dist, d_n_matches, d_match_from, d_match_to,
                                                                      d_n_unique_to, lp);
  ctx->launches++;
  SEPB_CUDA(ctx, cudaGetLastError());
  return SEPB_OK;
}
