// Roofline denominators that MEASURED_PEAKS.json does not carry: POPC (K2) and FP64 FMA (K3)
// issue-rate microbenchmarks.  Measurement support only; not on the data path.
#include "common.cuh"
#include "umma.cuh"

namespace {

__global__ void __launch_bounds__(256) popc_peak_kernel(uint32_t* out, int iters, uint32_t seed) {
  uint32_t x0 = seed + threadIdx.x, x1 = x0 * 3u, x2 = x0 * 5u, x3 = x0 * 7u;
  uint32_t x4 = x0 * 11u, x5 = x0 * 13u, x6 = x0 * 17u, x7 = x0 * 19u;
  for (int i = 0; i < iters; ++i) {
    // 8 independent dependent-chains of POPC: x = popc(x) ^ c keeps the pipe busy without folding
    x0 = __popc(x0) ^ 0x9E3779B9u; x1 = __popc(x1) ^ 0x7F4A7C15u; x2 = __popc(x2) ^ 0xF39CC060u; x3 = __popc(x3) ^ 0x5CEDC834u;
    x4 = __popc(x4) ^ 0x1082276Bu; x5 = __popc(x5) ^ 0xF3A5C9E1u; x6 = __popc(x6) ^ 0x3C6EF372u; x7 = __popc(x7) ^ 0xA54FF53Au;
  }
  out[blockIdx.x * blockDim.x + threadIdx.x] = x0 ^ x1 ^ x2 ^ x3 ^ x4 ^ x5 ^ x6 ^ x7;
}

__global__ void __launch_bounds__(256) dfma_peak_kernel(double* out, int iters, double seed) {
  double a0 = seed + threadIdx.x, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3, a4 = a0 + 4, a5 = a0 + 5, a6 = a0 + 6, a7 = a0 + 7;
  const double m = 0.999999, c = 1e-9;
  for (int i = 0; i < iters; ++i) {
    a0 = fma(a0, m, c); a1 = fma(a1, m, c); a2 = fma(a2, m, c); a3 = fma(a3, m, c);
    a4 = fma(a4, m, c); a5 = fma(a5, m, c); a6 = fma(a6, m, c); a7 = fma(a7, m, c);
  }
  out[blockIdx.x * blockDim.x + threadIdx.x] = a0 + a1 + a2 + a3 + a4 + a5 + a6 + a7;
}

// TMEM -> register read rate (tcgen05.ld 32x32b.x32, and the .pack::16b form K2T uses): 8 warps of one CTA per SM, two
// per lane quarter, sweep the 512 columns `iters` times.  Uninitialised TMEM content is fine: only the rate matters.
__global__ void __launch_bounds__(256, 1) tmem_read_peak_kernel(uint32_t* out, int iters, int packed) {
  __shared__ uint32_t slot;
  const int warp = threadIdx.x >> 5;
  if (warp == 0) umma::tmem_alloc(&slot, 512);
  umma::tc_fence_before();
  __syncthreads();
  umma::tc_fence_after();
  const uint32_t base = slot + ((uint32_t)((warp & 3) * 32) << 16);
  uint32_t acc = 0;
  for (int i = 0; i < iters; ++i) {
    for (int c = 0; c < 512; c += 64) {
      if (packed) {
        uint32_t a[16], b[16];
        umma::tmem_ld32_pack16(base + c, a);
        umma::tmem_ld32_pack16(base + c + 32, b);
        umma::tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 16; ++j) acc ^= a[j] + b[j];
      } else {
        uint32_t a[32], b[32];
        umma::tmem_ld32(base + c, a);
        umma::tmem_ld32(base + c + 32, b);
        umma::tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 32; ++j) acc ^= a[j] + b[j];
      }
    }
  }
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
  umma::tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    umma::tc_fence_after();
    umma::tmem_dealloc(slot, 512);
  }
}

// Tensor-pipe issue rate of the MMA shapes the matchers use: one CTA per SM, one elected thread issues `tiles`
// accumulator tiles of 8 K-steps each (SWIZZLE_128B operands, 2 K-blocks of 128 B like K2T / K1) round-robin over
// the TMEM stages, nothing drains them.  mode 0: kind::i8 M128 N128 K32; 1: kind::i8 M128 N256 K32;
// 2: kind::f16 (bf16) M128 N256 K16; 3: mode 0 while four more warps store into shared memory the way K2T's
// producers do (operand-fetch contention).  Operand content is whatever shared memory holds: only the rate matters.
__global__ void __launch_bounds__(160, 1) mma_peak_kernel(uint32_t* out, int tiles, int mode) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (umma::smem_u32(smem_raw) & 1023u)) & 1023u);
  __shared__ uint64_t bar;
  __shared__ uint32_t slot;
  __shared__ volatile int done;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int N = (mode == 1 || mode == 2) ? 256 : 128;
  const uint32_t sA = umma::smem_u32(smem), sB = sA + 2 * 128 * 128, sScratch = sB + 2 * 256 * 128;
  if (threadIdx.x == 0) {
    umma::mbar_init(&bar, 1);
    umma::fence_mbar_init();
    done = 0;
  }
  if (warp == 0) umma::tmem_alloc(&slot, 512);
  umma::tc_fence_before();
  __syncthreads();
  umma::tc_fence_after();
  const uint32_t tmem = slot;
  if (warp == 0) {
    if (umma::elect_one()) {
      const uint32_t idesc = mode == 2 ? umma::instr_desc(1, 1, 1, 128, 256) : umma::instr_desc(2, 1, 1, 128, (uint32_t)N);
      for (int t = 0; t < tiles; ++t) {
        const uint32_t d = tmem + (uint32_t)((t * N) & 511);
#pragma unroll
        for (int kb = 0; kb < 2; ++kb)
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            const uint64_t a = umma::smem_desc_sw128(sA + (uint32_t)(kb * 128 * 128 + k * 32));
            const uint64_t b = umma::smem_desc_sw128(sB + (uint32_t)(kb * N * 128 + k * 32));
            if (mode == 2) umma::mma_f16(d, a, b, idesc, (kb | k) != 0);
            else umma::mma_i8(d, a, b, idesc, (kb | k) != 0);
          }
      }
      umma::mma_commit(&bar);
    }
    __syncwarp();
    umma::mbar_wait(&bar, 0);
    done = 1;
  } else if (mode == 3) {
    uint32_t i = 0;
    while (!done) {  // 128 threads x 16 B, rows of 128 B: the store pattern of expand_row
      umma::sts128(sScratch + (uint32_t)(((threadIdx.x - 32) * 128 + (i & 7) * 16)), i, i, i, i);
      ++i;
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = i;
  }
  (void)lane;
  umma::tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    umma::tc_fence_after();
    umma::tmem_dealloc(tmem, 512);
  }
}


// The same bare issue loop for a CTA PAIR: tcgen05.mma.cta_group::2.kind::i8, M256 (128 To rows per CTA) x N256 (each
// CTA holds 128 of the 256 B rows) x K32.  Only the leader CTA (cluster rank 0) issues; per SM and multiply-add the
// operand fetch is 8 KB instead of the 12 KB of the one-CTA N256 shape.  mode 1: kind::f16 (bf16) K16.
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(128, 1) mma_peak2_kernel(uint32_t* out, int tiles, int mode) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (umma::smem_u32(smem_raw) & 1023u)) & 1023u);
  __shared__ uint64_t bar;
  __shared__ uint32_t slot;
  const int warp = threadIdx.x >> 5;
  uint32_t rank;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(rank));
  const uint32_t sA = umma::smem_u32(smem), sB = sA + 2 * 128 * 128;
  if (threadIdx.x == 0) {
    umma::mbar_init(&bar, 1);
    umma::fence_mbar_init();
  }
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(umma::smem_u32(&slot)), "r"(512u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
  }
  umma::tc_fence_before();
  __syncthreads();
  umma::tc_fence_after();
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
  const uint32_t tmem = slot;
  if (warp == 0) {
    if (rank == 0 && umma::elect_one()) {
      const uint32_t idesc = mode == 1 ? umma::instr_desc(1, 1, 1, 256, 256) : umma::instr_desc(2, 1, 1, 256, 256);
      for (int t = 0; t < tiles; ++t) {
        const uint32_t d = tmem + (uint32_t)((t * 256) & 511);
#pragma unroll
        for (int kb = 0; kb < 2; ++kb)
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            const uint64_t a = umma::smem_desc_sw128(sA + (uint32_t)(kb * 128 * 128 + k * 32));
            const uint64_t b = umma::smem_desc_sw128(sB + (uint32_t)(kb * 128 * 128 + k * 32));
            const uint32_t acc = (kb | k) != 0;
            if (mode == 1)
              asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d),
                           "l"(a), "l"(b), "r"(idesc), "r"(acc) : "memory");
            else
              asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::2.kind::i8 [%0], %1, %2, %3, p;\n\t}" ::"r"(d),
                           "l"(a), "l"(b), "r"(idesc), "r"(acc) : "memory");
          }
      }
      // completion arrives on the barrier of BOTH CTAs (same shared-memory offset)
      asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(umma::smem_u32(&bar)),
                   "h"((uint16_t)3) : "memory");
    }
    __syncwarp();
    umma::mbar_wait(&bar, 0);
  }
  if (threadIdx.x == 0) out[blockIdx.x] = rank;
  umma::tc_fence_before();
  __syncthreads();
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
  if (warp == 0) {
    umma::tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512u) : "memory");
  }
}

}  // namespace

extern "C" {

// kind 0: POPC ops/s; kind 1: FP64 FMA/s (one FMA = 2 FLOP); kind 2 / 3: TMEM bytes/s read by tcgen05.ld (32-bit cells,
// plain / .pack::16b); kind 4..7: tensor-pipe ops/s (2 per multiply-add) of mma_peak_kernel modes 0..3.
// Best of `reps` launches.
int sepb_measure_peak(sepb_ctx* ctx, int kind, int reps, double* ops_per_s) {
  if (!ctx || !ops_per_s) return SEPB_E_INVALID;
  SEPB_CUDA(ctx, cudaSetDevice(ctx->device));
  if (kind == 8 || kind == 9) {   // CTA-pair issue loop (cta_group::2): int8 / bf16, M256 x N256
    const int tiles = 4096, blocks = ctx->sm_count & ~1;
    const size_t smem = 4 * 128 * 128 + 1024;
    SEPB_CUDA(ctx, cudaFuncSetAttribute(mma_peak2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    void* buf = sepb_scratch(ctx, (size_t)blocks * 4);
    if (!buf) return (int)cudaErrorMemoryAllocation;
    cudaEvent_t e0, e1;
    SEPB_CUDA(ctx, cudaEventCreate(&e0));
    SEPB_CUDA(ctx, cudaEventCreate(&e1));
    double best = 0;
    for (int r = 0; r < reps + 1; ++r) {
      SEPB_CUDA(ctx, cudaEventRecord(e0, ctx->stream));
      mma_peak2_kernel<<<blocks, 128, smem, ctx->stream>>>((uint32_t*)buf, tiles, kind - 8);
      ctx->launches++;
      SEPB_CUDA(ctx, cudaEventRecord(e1, ctx->stream));
      SEPB_CUDA(ctx, cudaEventSynchronize(e1));
      SEPB_CUDA(ctx, cudaGetLastError());
      float ms = 0;
      SEPB_CUDA(ctx, cudaEventElapsedTime(&ms, e0, e1));
      const double ops = (double)(blocks / 2) * tiles * 8 * (2.0 * 256 * 256 * (kind == 9 ? 16.0 : 32.0)) / (ms * 1e-3);
      if (r > 0 && ops > best) best = ops;
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    *ops_per_s = best;
    return SEPB_OK;
  }
  const bool tm = kind == 2 || kind == 3, mm = kind >= 4 && kind <= 7;
  const int mma_tiles = 4096;
  const size_t mma_smem = 2 * 128 * 128 + 2 * 256 * 128 + 128 * 128 + 1024;
  if (mm) SEPB_CUDA(ctx, cudaFuncSetAttribute(mma_peak_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)mma_smem));
  const int blocks = (tm || mm) ? ctx->sm_count : ctx->sm_count * 8, threads = 256, iters = kind == 0 ? 4096 : tm ? 512 : 2048;
  void* buf = sepb_scratch(ctx, (size_t)blocks * threads * 8);
  if (!buf) return (int)cudaErrorMemoryAllocation;
  cudaEvent_t e0, e1;
  SEPB_CUDA(ctx, cudaEventCreate(&e0));
  SEPB_CUDA(ctx, cudaEventCreate(&e1));
  double best = 0;
  for (int r = 0; r < reps + 1; ++r) {
    SEPB_CUDA(ctx, cudaEventRecord(e0, ctx->stream));
    if (kind == 0) popc_peak_kernel<<<blocks, threads, 0, ctx->stream>>>((uint32_t*)buf, iters, 12345u + r);
    else if (kind == 1) dfma_peak_kernel<<<blocks, threads, 0, ctx->stream>>>((double*)buf, iters, 1.0 + r);
    else if (mm) mma_peak_kernel<<<blocks, 160, mma_smem, ctx->stream>>>((uint32_t*)buf, mma_tiles, kind - 4);
    else tmem_read_peak_kernel<<<blocks, threads, 0, ctx->stream>>>((uint32_t*)buf, iters, kind == 3);
    ctx->launches++;
    SEPB_CUDA(ctx, cudaEventRecord(e1, ctx->stream));
    SEPB_CUDA(ctx, cudaEventSynchronize(e1));
    float ms = 0;
    SEPB_CUDA(ctx, cudaEventElapsedTime(&ms, e0, e1));
    // TMEM: every warp reads 32 lanes x 512 columns x 4 B per sweep (the packed form returns half the registers)
    const double mma_k = kind == 6 ? 16.0 : 32.0, mma_n = (kind == 5 || kind == 6) ? 256.0 : 128.0;
    const double ops = mm ? (double)blocks * mma_tiles * 8 * (2.0 * 128 * mma_n * mma_k) / (ms * 1e-3) : tm ? (double)blocks * 8 * iters * (32.0 * 512 * 4) / (ms * 1e-3)
                          : (double)blocks * threads * iters * 8.0 / (ms * 1e-3);
    if (r > 0 && ops > best) best = ops;
  }
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  *ops_per_s = best;
  return SEPB_OK;
}

}  // extern "C"
