// Shared host-side plumbing for the C ABI (context, frame store, error capture).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include <mutex>
#include <string>

#include "../../include/sepb200.h"

struct sepb_ctx {
  int device = 0;
  cudaStream_t stream = nullptr;
  cudaStream_t copy_stream = nullptr;   // H2D of sepb_frames_upload_async, concurrent with the kernels on `stream`
  std::string err;
  int64_t launches = 0;
  std::mutex mu;
  // grow-only device scratch (pairs, outputs of host-level calls)
  void* scratch = nullptr;
  size_t scratch_bytes = 0;
  // second grow-only buffer: staging of the host-buffer entry points whose launches use `scratch` underneath
  void* io = nullptr;
  size_t io_bytes = 0;
  int sm_count = 148;
  // cv::RNG(-1) minimal-sample index tables [count][table_h][mp] (K3): [0] 5-point sets, [1] 4-point sets (P3P flags)
  uint16_t* table = nullptr;
  int table_m = -1, table_h = 0;
  uint16_t* table4 = nullptr;
  int table4_m = -1, table4_h = 0;
  // K3 work queue (pair list, worker states, chunk slots of the helper protocol; pnp_ransac.cu), grow-only, in int32 units
  int32_t* k3_order = nullptr;
  size_t k3_order_cap = 0;
  // per-device kernel attribute bookkeeping (cudaFuncSetAttribute is per device)
  size_t cfg_hamming_smem = 0, cfg_hamming_mma_smem = 0, cfg_hamming_mma_smem_wide = 0, cfg_guess_smem = 0;
  size_t cfg_hamming_mma_psmem = 0, cfg_hamming_mma_psmem_wide = 0, cfg_ba_smem = 0;
  sepb_ba_params ba{};          // ba.type == 0: no bundle adjustment after PnP (sepb_ctx_set_bundle_adjustment)
  int opt_match_kernel = SEPB_MATCH_AUTO;
  int opt_match_tile = 0;       // 0: widest K2T accumulator tile that fits, 128: force the narrow geometry
  int opt_pnp_sched = SEPB_PNP_SCHED_PERSISTENT;
  int opt_match_sched = SEPB_MATCH_SCHED_PERSISTENT;
  int opt_nn_mode = -1;  // reserved
  bool cfg_pnp = false, cfg_exact = false, cfg_gemm = false;
};

struct sepb_db {
  sepb_ctx* ctx = nullptr;
  int dim = 0, capacity = 0, size = 0;
  int64_t base = 0;            // global index of row 0 (row-sharded databases)
  double* rows = nullptr;      // [capacity][dim] exact copy of the float64[] wire data
  uint8_t* col_mask = nullptr; // [capacity] other_kf_already_used
  bool has_mask = false;
  // tensor-core filter operands (K1): split-bf16 rows [capacity][3*dim_pad] and squared norms, filled lazily
  uint16_t* split = nullptr;   // [capacity][2*dim_pad] bf16: hi | lo
  float* norm2 = nullptr;      // [capacity] squared norms (f32, from the f64 sum)
  float* maxnorm2 = nullptr;   // device scalar: max of norm2 over converted rows (bit pattern ordered as u32)
  int dim_pad = 0;             // dim rounded up to 64
  int split_rows = 0;          // rows [0, split_rows) are converted
  alignas(64) unsigned char tmap[128];  // CUtensorMap over `split` with tmap_rows rows
  int tmap_rows = -1;
};

struct sepb_frames {
  sepb_ctx* ctx = nullptr;
  int capacity = 0;
  int max_feat = 0;
  int size = 0;  // highest uploaded slot + 1
  uint8_t* desc = nullptr;   // [capacity][max_feat][32]
  float* kp2d = nullptr;     // [capacity][max_feat][2]
  float* kp3d = nullptr;     // [capacity][max_feat][3]
  int32_t* octave = nullptr; // [capacity][max_feat]
  int32_t* count = nullptr;  // [capacity]
  // asynchronous uploads: kernels wait for ev_uploaded, the next upload waits for ev_last_use
  cudaEvent_t ev_uploaded = nullptr, ev_last_use = nullptr;
  mutable bool has_upload = false, has_use = false;
  // lazy keypoints (sepb_frames_bind_host_async, frames_fetch.cu): device-visible addresses of the caller's pinned
  // arrays for slots [lazy_first, lazy_first + lazy_n); verification pulls the rows it needs
  bool lazy = false;
  int lazy_first = 0, lazy_n = 0;
  const float* h_kp2d = nullptr;
  const float* h_kp3d = nullptr;
  const int32_t* h_octave = nullptr;
  uint8_t* need = nullptr;                // [capacity rounded up to 4] bit 0 kp2d, 1 kp3d, 2 octave: wanted by the current call
  uint8_t* have = nullptr;                // [capacity rounded up to 4] same bits: resident for the current binding
  unsigned long long* fetched = nullptr;  // device counter: bytes pulled since the binding
};

// device view of a frame store passed by value to kernels
struct FramesView {
  const uint8_t* desc;
  const float* kp2d;
  const float* kp3d;
  const int32_t* octave;
  const int32_t* count;
  int max_feat;
  int size;
  // host-bound slots [lazy_first, lazy_first + lazy_n): their 3-D points are readable in place (pinned host memory
  // mapped into the device address space) before the fetch kernel has made them resident -- K2's list compaction only
  // asks whether the 3-D point of a matched From feature is finite
  const float* kp3d_host;
  int lazy_first, lazy_n;
#ifdef __CUDACC__
  // 3-D points of frame `slot` for a reader that may run BEFORE the pull (the matchers)
  __device__ __forceinline__ const float* kp3d_early(int slot) const {
    const unsigned rel = (unsigned)(slot - lazy_first);
    return rel < (unsigned)lazy_n ? kp3d_host + (size_t)rel * max_feat * 3 : kp3d + (size_t)slot * max_feat * 3;
  }
#endif
};

// host-bound stores: what a matcher needs to pull the key points of a pair that will run PnP while the rest of the
// batch is still being matched (the PCIe reads spread over K2 instead of a fetch phase between K2 and K3)
struct LazyPull {
  int min_matches;                 // 0: off
  float* from_kp3d;                // device rows / host rows / resident bits / byte counter of the From store
  const float* from_kp3d_host;
  uint8_t* from_have;
  unsigned long long* from_fetched;
  int from_first, from_n;
  float* to_kp2d;
  const float* to_kp2d_host;
  uint8_t* to_have;
  unsigned long long* to_fetched;
  int to_first, to_n;
};

#ifdef __CUDACC__
// Copy of up to two rows out of mapped host memory by `nthreads` threads (row B may be empty).  A load from host memory
// is a PCIe round trip (microseconds; tens of them while a bulk cudaMemcpyAsync saturates the bus), so every 16-byte
// load of a batch -- U per thread, both rows in one index space -- is issued before the first store: a 12 KB + 8 KB
// pair of rows costs 256 threads ONE round trip instead of five dependent ones.
template <int U = 8>
__device__ __forceinline__ void sepb_pull_rows(void* dA, const void* sA, int bytesA, void* dB, const void* sB, int bytesB,
                                               int tid, int nthreads) {
  if (((((uintptr_t)dA | (uintptr_t)sA) & 15) == 0) && (bytesB == 0 || ((((uintptr_t)dB | (uintptr_t)sB) & 15) == 0))) {
    const int nA = bytesA >> 4, nB = bytesB >> 4, n = nA + nB;
    const int4* a16 = reinterpret_cast<const int4*>(sA);
    const int4* b16 = reinterpret_cast<const int4*>(sB);
    int4* da16 = reinterpret_cast<int4*>(dA);
    int4* db16 = reinterpret_cast<int4*>(dB);
    // the (< 16 byte) ends of the rows: one word per thread, loaded up front as well
    const int wA = (bytesA & 15) >> 2, wB = (bytesB & 15) >> 2;
    int32_t tail = 0;
    const bool tail_a = tid < wA, tail_b = !tail_a && tid >= 4 && tid - 4 < wB;
    if (tail_a) tail = reinterpret_cast<const int32_t*>(sA)[4 * nA + tid];
    if (tail_b) tail = reinterpret_cast<const int32_t*>(sB)[4 * nB + tid - 4];
    for (int base = 0; base < n; base += U * nthreads) {
      int4 v[U];
#pragma unroll
      for (int j = 0; j < U; ++j) {
        const int i = base + j * nthreads + tid;
        if (i < n) v[j] = __ldcs(i < nA ? a16 + i : b16 + (i - nA));
      }
#pragma unroll
      for (int j = 0; j < U; ++j) {
        const int i = base + j * nthreads + tid;
        if (i < n) {
          if (i < nA) da16[i] = v[j];
          else db16[i - nA] = v[j];
        }
      }
    }
    if (tail_a) reinterpret_cast<int32_t*>(dA)[4 * nA + tid] = tail;
    if (tail_b) reinterpret_cast<int32_t*>(dB)[4 * nB + tid - 4] = tail;
  } else {
    for (int i = 4 * tid; i < bytesA; i += 4 * nthreads)
      *reinterpret_cast<int32_t*>((char*)dA + i) = *reinterpret_cast<const int32_t*>((const char*)sA + i);
    for (int i = 4 * tid; i < bytesB; i += 4 * nthreads)
      *reinterpret_cast<int32_t*>((char*)dB + i) = *reinterpret_cast<const int32_t*>((const char*)sB + i);
  }
}
// CTA-wide copy of one row
__device__ __forceinline__ void sepb_pull_row(void* dst, const void* src, int bytes) {
  sepb_pull_rows(dst, src, bytes, nullptr, nullptr, 0, threadIdx.x, blockDim.x);
}
// first CTA to ask for `bit` of a slot's row claims the copy (one lane calls; the resident bits are bytes)
__device__ __forceinline__ bool sepb_claim_row(uint8_t* have, int slot, unsigned bit) {
  const unsigned sh = 8u * ((unsigned)slot & 3u);
  const unsigned old = atomicOr(reinterpret_cast<unsigned int*>(have + (slot & ~3)), bit << sh);
  return !((old >> sh) & bit);
}
#endif

static inline FramesView view_of(const sepb_frames* f) {
  return FramesView{f->desc, f->kp2d, f->kp3d, f->octave, f->count, f->max_feat, f->size,
                    f->lazy ? f->h_kp3d : nullptr, f->lazy ? f->lazy_first : 0, f->lazy ? f->lazy_n : 0};
}

// order the compute stream after pending asynchronous uploads of the frame stores a launch is about to read ...
static inline void sepb_frames_acquire(sepb_ctx* ctx, const sepb_frames* a, const sepb_frames* b) {
  if (a->has_upload) cudaStreamWaitEvent(ctx->stream, a->ev_uploaded, 0);
  if (b != a && b->has_upload) cudaStreamWaitEvent(ctx->stream, b->ev_uploaded, 0);
}
// ... and remember the last kernel that read them, so that a later upload does not overwrite data still in use
static inline void sepb_frames_release(sepb_ctx* ctx, const sepb_frames* a, const sepb_frames* b) {
  cudaEventRecord(a->ev_last_use, ctx->stream);
  a->has_use = true;
  if (b != a) {
    cudaEventRecord(b->ev_last_use, ctx->stream);
    b->has_use = true;
  }
}

// scope guard: one acquire when a call starts enqueuing readers of the stores, one release after the LAST of them on
// every return path (a release recorded too early lets a following sepb_frames_upload_async overwrite data in use)
struct FramesUse {
  sepb_ctx* ctx;
  const sepb_frames *a, *b;
  FramesUse(sepb_ctx* c, const sepb_frames* a_, const sepb_frames* b_) : ctx(c), a(a_), b(b_) { sepb_frames_acquire(ctx, a, b); }
  ~FramesUse() { sepb_frames_release(ctx, a, b); }
  FramesUse(const FramesUse&) = delete;
  FramesUse& operator=(const FramesUse&) = delete;
};

// developer A/B switches read from the environment exist only in -DSEPB_DEVELOPER builds (SEPB_DEVELOPER=1 python -m
// ...build); a production library never calls getenv on a launch path
#ifdef SEPB_DEVELOPER
#include <stdlib.h>
static inline const char* sepb_dev_env(const char* name) { return getenv(name); }
#else
static inline const char* sepb_dev_env(const char*) { return nullptr; }
#endif

static inline int sepb_fail(sepb_ctx* ctx, int code, const char* what, const char* file, int line) {
  char buf[512];
  if (code > 0)
    snprintf(buf, sizeof buf, "%s: %s (%s:%d)", what, cudaGetErrorString((cudaError_t)code), file, line);
  else
    snprintf(buf, sizeof buf, "%s (%s:%d)", what, file, line);
  if (ctx) ctx->err = buf;
  return code;
}

#define SEPB_CUDA(ctx, expr)                                                     \
  do {                                                                           \
    cudaError_t _e = (expr);                                                     \
    if (_e != cudaSuccess) return sepb_fail((ctx), (int)_e, #expr, __FILE__, __LINE__); \
  } while (0)

#define SEPB_REQUIRE(ctx, cond, code)                                            \
  do {                                                                           \
    if (!(cond)) return sepb_fail((ctx), (code), "requirement failed: " #cond, __FILE__, __LINE__); \
  } while (0)

// grow-only scratch; returns nullptr on failure (error recorded)
static inline void* sepb_scratch(sepb_ctx* ctx, size_t bytes) {
  if (bytes <= ctx->scratch_bytes) return ctx->scratch;
  if (ctx->scratch) {
    cudaStreamSynchronize(ctx->stream);
    cudaFree(ctx->scratch);
    ctx->scratch = nullptr;
    ctx->scratch_bytes = 0;
  }
  size_t want = bytes + bytes / 4 + 4096;
  cudaError_t e = cudaMalloc(&ctx->scratch, want);
  if (e != cudaSuccess) {
    sepb_fail(ctx, (int)e, "cudaMalloc(scratch)", __FILE__, __LINE__);
    return nullptr;
  }
  ctx->scratch_bytes = want;
  return ctx->scratch;
}

static inline void* sepb_scratch_io(sepb_ctx* ctx, size_t bytes) {
  if (bytes <= ctx->io_bytes) return ctx->io;
  if (ctx->io) {
    cudaStreamSynchronize(ctx->stream);
    cudaFree(ctx->io);
    ctx->io = nullptr;
    ctx->io_bytes = 0;
  }
  const size_t want = bytes + bytes / 4 + 4096;
  cudaError_t e = cudaMalloc(&ctx->io, want);
  if (e != cudaSuccess) {
    sepb_fail(ctx, (int)e, "cudaMalloc(io)", __FILE__, __LINE__);
    return nullptr;
  }
  ctx->io_bytes = want;
  return ctx->io;
}

static inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

// kernel launchers implemented in the .cu files
int sepb_launch_hamming_match(sepb_ctx* ctx, FramesView from, FramesView to, const int32_t* d_pairs, int n_pairs,
                              float nndr, int strict, int32_t* d_knn_idx, int32_t* d_knn_dist,
                              int32_t* d_n_matches, uint16_t* d_match_from, uint16_t* d_match_to,
                              int32_t* d_n_unique_to);

// tensor-core variant (hamming_mma.cu); same contract, bit-identical outputs
bool sepb_hamming_mma_supported(const FramesView& from, const FramesView& to);
int sepb_launch_hamming_mma(sepb_ctx* ctx, FramesView from, FramesView to, const int32_t* d_pairs, int n_pairs,
                            float nndr, int strict, int32_t* d_knn_idx, int32_t* d_knn_dist, int32_t* d_n_matches,
                            uint16_t* d_match_from, uint16_t* d_match_to, int32_t* d_n_unique_to,
                            const LazyPull* lazy = nullptr);
// dispatch on ctx->opt_match_kernel.  lazy (optional, host-bound stores): the tensor matcher pulls the key points of
// the pairs with at least lazy->min_matches correspondences in its tail (the POPC kernel leaves that to the fetch kernels)
int sepb_launch_match(sepb_ctx* ctx, FramesView from, FramesView to, const int32_t* d_pairs, int n_pairs, float nndr,
                      int strict, int32_t* d_knn_idx, int32_t* d_knn_dist, int32_t* d_n_matches,
                      uint16_t* d_match_from, uint16_t* d_match_to, int32_t* d_n_unique_to, const LazyPull* lazy = nullptr);

int sepb_launch_guess_match(sepb_ctx* ctx, FramesView from, FramesView to, const int32_t* d_pairs, int n_pairs,
                            const int32_t* d_status1, const float* d_transform1, const sepb_pnp_params* p, int width,
                            int height, float win, float nndr, int32_t* d_n_matches, uint16_t* d_match_from,
                            uint16_t* d_match_to, int32_t* d_n_unique_to, double* d_guess);
int sepb_launch_pose_cov(sepb_ctx* ctx, FramesView from, FramesView to, const int32_t* d_pairs, int n_pairs,
                         const int32_t* d_status, const int32_t* d_n_matches, const uint16_t* d_match_from,
                         const uint16_t* d_match_to, const uint32_t* d_inlier_mask, int mask_words,
                         const float* d_transform, double* d_cov);
LazyPull sepb_lazy_pull_of(sepb_frames* from, sepb_frames* to, int min_matches);
int sepb_launch_bundle_adjust(sepb_ctx* ctx, FramesView from, FramesView to, const int32_t* d_pairs, int n_pairs,
                              const int32_t* d_n_matches, const uint16_t* d_match_from, const uint16_t* d_match_to,
                              const double* d_cov, const sepb_pnp_params* pprm, const sepb_ba_params* bprm,
                              int32_t* d_status, int32_t* d_n_inliers, double* d_pose, float* d_transform,
                              uint32_t* d_inlier_mask, int mask_words, double* d_stats);
int sepb_launch_frames_fetch(sepb_ctx* ctx, sepb_frames* from, sepb_frames* to, const int32_t* d_pairs, int n_pairs,
                             const int32_t* d_n_matches, int min_matches, const int32_t* d_status, int bits_from, int bits_to);
int sepb_launch_sample_table(sepb_ctx* ctx, uint16_t* d_table, int max_count, int H, int model_points);
int sepb_launch_pnp_ransac(sepb_ctx* ctx, FramesView from, FramesView to, const int32_t* d_pairs, int n_pairs,
                           const int32_t* d_n_matches, const uint16_t* d_match_from, const uint16_t* d_match_to,
                           const int32_t* d_n_unique_to, const uint16_t* d_table, int table_h,
                           const int32_t* d_custom_samples, int custom_h, const double* d_guess,
                           const int32_t* d_only_if_ok, const sepb_pnp_params* p, int32_t* d_status, int32_t* d_n_inliers, double* d_pose,
                           float* d_transform, uint32_t* d_inlier_mask, int mask_words, int32_t* d_best_hyp,
                           int32_t* d_iters_run);

int sepb_launch_netvlad_exact(sepb_ctx* ctx, const double* d_q, int nq, const double* d_db, int n, int D, int k,
                              int64_t base, const uint8_t* d_row_mask, const uint8_t* d_col_mask,
                              const int32_t* d_ign_start, const int32_t* d_ign_col, int splits, int32_t* d_part_idx,
                              double* d_part_dist, int32_t* d_out_idx, double* d_out_dist,
                              const int32_t* d_row_list = nullptr, const int32_t* d_n_rows = nullptr);
int sepb_launch_topk_merge(sepb_ctx* ctx, const int32_t* d_in_idx, const double* d_in_dist, int lists, int nq, int k,
                           int32_t* d_out_idx, double* d_out_dist);
