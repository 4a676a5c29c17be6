// Bring-up check of csrc/umma.cuh on a B200: one CTA computes D = A * B^T (M=128, N=256) through
// tcgen05.mma with (1) thread-written SWIZZLE_128B int8 operands, (2) thread-written bf16 operands,
// (3) TMA-loaded bf16 operands, and compares with a host reference.  Developer tool, not shipped:
//   nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o gpurun_out/umma_selftest tools/umma_selftest.cu
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <vector>

#include "../multi_robot_slam_separators_b200/csrc/umma.cuh"

#define CK(x)                                                                          \
  do {                                                                                 \
    cudaError_t e_ = (x);                                                              \
    if (e_ != cudaSuccess) {                                                           \
      printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__);  \
      exit(2);                                                                         \
    }                                                                                  \
  } while (0)

constexpr int M = 128, N = 256;

// byte offset of (row r, byte kb of the row) inside a SW128 tile of `rows` rows
__device__ __host__ inline uint32_t sw128_off(int rows, int r, int kb) {
  const int blk = kb >> 7, in = kb & 127;
  return (uint32_t)blk * rows * 128 + r * 128 + ((((in >> 4) ^ (r & 7)) << 4) | (in & 15));
}

// mode 0: int8 (K bytes per row = KB), mode 1: bf16 thread-written, mode 2: bf16 via TMA
template <int MODE>
__global__ void __launch_bounds__(128) gemm_kernel(const uint8_t* __restrict__ A, const uint8_t* __restrict__ B, int KB,
                                                   uint32_t* __restrict__ D, const __grid_constant__ CUtensorMap tmA,
                                                   const __grid_constant__ CUtensorMap tmB) {
  extern __shared__ uint8_t raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)raw + 1023) & ~(uintptr_t)1023);
  uint8_t* sA = smem;
  uint8_t* sB = smem + (size_t)M * KB;
  __shared__ uint64_t bar_mma, bar_tma;
  __shared__ uint32_t tmem_slot;
  const int tid = threadIdx.x, warp = tid >> 5;
  if (tid == 0) {
    umma::mbar_init(&bar_mma, 1);
    umma::mbar_init(&bar_tma, 1);
    umma::fence_mbar_init();
  }
  if (warp == 0) umma::tmem_alloc(&tmem_slot, 256);
  umma::tc_fence_before();
  __syncthreads();
  umma::tc_fence_after();
  const uint32_t tmem = tmem_slot;

  if (MODE == 2) {
    if (tid == 0) {
      umma::mbar_arrive_expect_tx(&bar_tma, (uint32_t)((M + N) * KB));
      for (int blk = 0; blk < KB / 128; ++blk) {
        umma::tma_load_2d(sA + (size_t)blk * M * 128, &tmA, &bar_tma, blk * 64, 0);
        umma::tma_load_2d(sB + (size_t)blk * N * 128, &tmB, &bar_tma, blk * 64, 0);
      }
    }
    umma::mbar_wait(&bar_tma, 0);
  } else {
    for (int i = tid; i < M * KB / 16; i += 128) {
      const int r = i / (KB / 16), c = i % (KB / 16);
      *(uint4*)(sA + sw128_off(M, r, c * 16)) = *(const uint4*)(A + (size_t)r * KB + c * 16);
    }
    for (int i = tid; i < N * KB / 16; i += 128) {
      const int r = i / (KB / 16), c = i % (KB / 16);
      *(uint4*)(sB + sw128_off(N, r, c * 16)) = *(const uint4*)(B + (size_t)r * KB + c * 16);
    }
    umma::fence_proxy_async();
    __syncthreads();
  }
  if (tid == 0) {
    umma::tc_fence_after();
    const uint32_t idesc = MODE == 0 ? umma::instr_desc(2, 1, 1, M, N) : umma::instr_desc(1, 1, 1, M, N);
    const uint32_t a0 = umma::smem_u32(sA), b0 = umma::smem_u32(sB);
    for (int blk = 0; blk < KB / 128; ++blk)
      for (int k = 0; k < 4; ++k) {
        const uint64_t ad = umma::smem_desc_sw128(a0 + blk * M * 128 + k * 32);
        const uint64_t bd = umma::smem_desc_sw128(b0 + blk * N * 128 + k * 32);
        if (MODE == 0)
          umma::mma_i8(tmem, ad, bd, idesc, (blk | k) != 0);
        else
          umma::mma_f16(tmem, ad, bd, idesc, (blk | k) != 0);
      }
    umma::mma_commit(&bar_mma);
  }
  umma::mbar_wait(&bar_mma, 0);
  umma::tc_fence_after();
  for (int c = 0; c < N; c += 32) {
    uint32_t r[32];
    umma::tmem_ld32(tmem + ((uint32_t)(warp * 32) << 16) + c, r);
    umma::tmem_ld_wait();
    for (int j = 0; j < 32; ++j) D[(size_t)tid * N + c + j] = r[j];
  }
  umma::tc_fence_before();
  __syncthreads();
  if (warp == 0) umma::tmem_dealloc(tmem, 256);
}

// mode 3: int8, K = 256 (SW128) + 32 "extra" bytes per row in the interleaved no-swizzle layout, N = 224
__global__ void __launch_bounds__(128) gemm_extras_kernel(const uint8_t* __restrict__ A, const uint8_t* __restrict__ B,
                                                          uint32_t* __restrict__ D, int NN) {
  extern __shared__ uint8_t raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)raw + 1023) & ~(uintptr_t)1023);
  const int KB = 288;
  uint8_t* sA = smem;                 // [2][128][128]
  uint8_t* sB = sA + M * 256;         // [2][N][128]
  uint8_t* sAx = sB + N * 256;        // [16 groups][2][8][16]
  uint8_t* sBx = sAx + M * 32;        // [32 groups][2][8][16]
  __shared__ uint64_t bar_mma;
  __shared__ uint32_t tmem_slot;
  const int tid = threadIdx.x, warp = tid >> 5;
  if (tid == 0) {
    umma::mbar_init(&bar_mma, 1);
    umma::fence_mbar_init();
  }
  if (warp == 0) umma::tmem_alloc(&tmem_slot, 256);
  umma::tc_fence_before();
  __syncthreads();
  umma::tc_fence_after();
  const uint32_t tmem = tmem_slot;
  for (int i = tid; i < M * 16; i += 128) {
    const int r = i / 16, c = i % 16;
    *(uint4*)(sA + sw128_off(M, r, c * 16)) = *(const uint4*)(A + (size_t)r * KB + c * 16);
  }
  for (int i = tid; i < N * 16; i += 128) {
    const int r = i / 16, c = i % 16;
    *(uint4*)(sB + sw128_off(N, r, c * 16)) = *(const uint4*)(B + (size_t)r * KB + c * 16);
  }
  for (int i = tid; i < M * 2; i += 128) {
    const int r = i / 2, c = i % 2;
    *(uint4*)(sAx + (r >> 3) * 256 + c * 128 + (r & 7) * 16) = *(const uint4*)(A + (size_t)r * KB + 256 + c * 16);
  }
  for (int i = tid; i < N * 2; i += 128) {
    const int r = i / 2, c = i % 2;
    *(uint4*)(sBx + (r >> 3) * 256 + c * 128 + (r & 7) * 16) = *(const uint4*)(B + (size_t)r * KB + 256 + c * 16);
  }
  umma::fence_proxy_async();
  __syncthreads();
  if (tid == 0) {
    umma::tc_fence_after();
    const uint32_t idesc = umma::instr_desc(2, 1, 1, M, NN);
    const uint32_t a0 = umma::smem_u32(sA), b0 = umma::smem_u32(sB);
    for (int blk = 0; blk < 2; ++blk)
      for (int k = 0; k < 4; ++k)
        umma::mma_i8(tmem, umma::smem_desc_sw128(a0 + blk * M * 128 + k * 32), umma::smem_desc_sw128(b0 + blk * N * 128 + k * 32),
                     idesc, (blk | k) != 0);
    umma::mma_i8(tmem, umma::smem_desc_interleave(umma::smem_u32(sAx), 128, 256),
                 umma::smem_desc_interleave(umma::smem_u32(sBx), 128, 256), idesc, true);
    umma::mma_commit(&bar_mma);
  }
  umma::mbar_wait(&bar_mma, 0);
  umma::tc_fence_after();
  for (int c = 0; c < N; c += 32) {
    uint32_t r[32];
    umma::tmem_ld32(tmem + ((uint32_t)(warp * 32) << 16) + c, r);
    umma::tmem_ld_wait();
    for (int j = 0; j < 32; ++j) D[(size_t)tid * N + c + j] = r[j];
  }
  umma::tc_fence_before();
  __syncthreads();
  if (warp == 0) umma::tmem_dealloc(tmem, 256);
}

// probe of tcgen05.ld ... .pack::16b : D (s32, small values) read back packed; host checks the column order
__global__ void __launch_bounds__(128) pack_probe_kernel(const uint8_t* __restrict__ A, const uint8_t* __restrict__ B,
                                                         uint32_t* __restrict__ D) {
  extern __shared__ uint8_t raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)raw + 1023) & ~(uintptr_t)1023);
  const int KB = 256;
  uint8_t* sA = smem;
  uint8_t* sB = smem + (size_t)M * KB;
  __shared__ uint64_t bar_mma;
  __shared__ uint32_t tmem_slot;
  const int tid = threadIdx.x, warp = tid >> 5;
  if (tid == 0) {
    umma::mbar_init(&bar_mma, 1);
    umma::fence_mbar_init();
  }
  if (warp == 0) umma::tmem_alloc(&tmem_slot, 256);
  umma::tc_fence_before();
  __syncthreads();
  umma::tc_fence_after();
  const uint32_t tmem = tmem_slot;
  for (int i = tid; i < M * KB / 16; i += 128) {
    const int r = i / (KB / 16), c = i % (KB / 16);
    *(uint4*)(sA + sw128_off(M, r, c * 16)) = *(const uint4*)(A + (size_t)r * KB + c * 16);
  }
  for (int i = tid; i < N * KB / 16; i += 128) {
    const int r = i / (KB / 16), c = i % (KB / 16);
    *(uint4*)(sB + sw128_off(N, r, c * 16)) = *(const uint4*)(B + (size_t)r * KB + c * 16);
  }
  umma::fence_proxy_async();
  __syncthreads();
  if (tid == 0) {
    umma::tc_fence_after();
    const uint32_t idesc = umma::instr_desc(2, 1, 1, M, N);
    const uint32_t a0 = umma::smem_u32(sA), b0 = umma::smem_u32(sB);
    for (int blk = 0; blk < 2; ++blk)
      for (int k = 0; k < 4; ++k)
        umma::mma_i8(tmem, umma::smem_desc_sw128(a0 + blk * M * 128 + k * 32), umma::smem_desc_sw128(b0 + blk * N * 128 + k * 32),
                     idesc, (blk | k) != 0);
    umma::mma_commit(&bar_mma);
  }
  umma::mbar_wait(&bar_mma, 0);
  umma::tc_fence_after();
  for (int c = 0; c < N; c += 32) {
    uint32_t r[16];
    umma::tmem_ld32_pack16(tmem + ((uint32_t)(warp * 32) << 16) + c, r);
    umma::tmem_ld_wait();
    for (int j = 0; j < 16; ++j) {
      D[(size_t)tid * N + c + 2 * j] = (uint32_t)(int32_t)(int16_t)(r[j] & 0xFFFFu);
      D[(size_t)tid * N + c + 2 * j + 1] = (uint32_t)(int32_t)(int16_t)(r[j] >> 16);
    }
  }
  umma::tc_fence_before();
  __syncthreads();
  if (warp == 0) umma::tmem_dealloc(tmem, 256);
}

typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                             const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                             CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static CUtensorMap make_map(EncodeFn enc, void* base, int rows, int cols_elems, int box_rows) {
  CUtensorMap m;
  cuuint64_t dims[2] = {(cuuint64_t)cols_elems, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)cols_elems * 2};
  cuuint32_t box[2] = {64, (cuuint32_t)box_rows};
  cuuint32_t es[2] = {1, 1};
  CUresult r = enc(&m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, base, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    printf("cuTensorMapEncodeTiled failed: %d\n", (int)r);
    exit(3);
  }
  return m;
}

int main() {
  CK(cudaSetDevice(0));
  EncodeFn enc = nullptr;
  cudaDriverEntryPointQueryResult qres;
  CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", (void**)&enc, cudaEnableDefault, &qres));
  if (!enc) {
    printf("no cuTensorMapEncodeTiled\n");
    return 3;
  }
  int fails = 0;
  for (int mode = 0; mode < 5; ++mode) {
    const int KB = (mode == 0 || mode == 4) ? 256 : mode == 3 ? 288 : 384;
    const int NN = mode == 3 ? 224 : N;  // columns actually computed  // bytes per row: int8 K=256; bf16 K=192
    std::vector<uint8_t> hA((size_t)M * KB), hB((size_t)N * KB);
    std::vector<double> ref((size_t)M * N);
    srand(1234 + mode);
    if (mode == 0 || mode == 3 || mode == 4) {
      for (auto& v : hA) v = (uint8_t)(int8_t)((rand() % 5) - 2);
      for (auto& v : hB) v = (uint8_t)(int8_t)(mode == 3 ? (rand() % 127) - 63 : (rand() & 1) ? 1 : -1);
      for (int i = 0; i < M; ++i)
        for (int j = 0; j < N; ++j) {
          long s = 0;
          for (int k = 0; k < KB; ++k) s += (long)(int8_t)hA[(size_t)i * KB + k] * (int8_t)hB[(size_t)j * KB + k];
          ref[(size_t)i * N + j] = (double)s;
        }
    } else {
      __nv_bfloat16* a = (__nv_bfloat16*)hA.data();
      __nv_bfloat16* b = (__nv_bfloat16*)hB.data();
      for (int i = 0; i < M * KB / 2; ++i) a[i] = __float2bfloat16((float)((rand() % 17) - 8) / 8.f);
      for (int i = 0; i < N * KB / 2; ++i) b[i] = __float2bfloat16((float)((rand() % 17) - 8) / 4.f);
      for (int i = 0; i < M; ++i)
        for (int j = 0; j < N; ++j) {
          double s = 0;
          for (int k = 0; k < KB / 2; ++k)
            s += (double)__bfloat162float(a[(size_t)i * (KB / 2) + k]) * (double)__bfloat162float(b[(size_t)j * (KB / 2) + k]);
          ref[(size_t)i * N + j] = s;
        }
    }
    uint8_t *dA, *dB;
    uint32_t* dD;
    CK(cudaMalloc(&dA, hA.size()));
    CK(cudaMalloc(&dB, hB.size()));
    CK(cudaMalloc(&dD, (size_t)M * N * 4));
    CK(cudaMemcpy(dA, hA.data(), hA.size(), cudaMemcpyHostToDevice));
    CK(cudaMemcpy(dB, hB.data(), hB.size(), cudaMemcpyHostToDevice));
    CK(cudaMemset(dD, 0xCD, (size_t)M * N * 4));
    CUtensorMap tmA, tmB;
    memset(&tmA, 0, sizeof tmA);
    memset(&tmB, 0, sizeof tmB);
    if (mode == 2) {
      tmA = make_map(enc, dA, M, KB / 2, M);
      tmB = make_map(enc, dB, N, KB / 2, N);
    }
    const size_t smem = (size_t)(M + N) * KB + 1024;
    if (mode == 0) {
      CK(cudaFuncSetAttribute(gemm_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      gemm_kernel<0><<<1, 128, smem>>>(dA, dB, KB, dD, tmA, tmB);
    } else if (mode == 1) {
      CK(cudaFuncSetAttribute(gemm_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      gemm_kernel<1><<<1, 128, smem>>>(dA, dB, KB, dD, tmA, tmB);
    } else if (mode == 4) {
      CK(cudaFuncSetAttribute(pack_probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      pack_probe_kernel<<<1, 128, smem>>>(dA, dB, dD);
    } else if (mode == 3) {
      CK(cudaFuncSetAttribute(gemm_extras_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      gemm_extras_kernel<<<1, 128, smem>>>(dA, dB, dD, NN);
    } else {
      CK(cudaFuncSetAttribute(gemm_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      gemm_kernel<2><<<1, 128, smem>>>(dA, dB, KB, dD, tmA, tmB);
    }
    CK(cudaGetLastError());
    CK(cudaDeviceSynchronize());
    std::vector<uint32_t> hD((size_t)M * N);
    CK(cudaMemcpy(hD.data(), dD, hD.size() * 4, cudaMemcpyDeviceToHost));
    int bad = 0;
    double maxerr = 0;
    for (size_t i = 0; i < hD.size(); ++i) {
      double got;
      if ((int)(i % N) >= NN) continue;
      if (mode == 0 || mode == 3 || mode == 4) {
        got = (double)(int32_t)hD[i];
      } else {
        float f;
        memcpy(&f, &hD[i], 4);
        got = f;
      }
      const double e = fabs(got - ref[i]);
      if (e > maxerr) maxerr = e;
      if (e > ((mode == 0 || mode == 3 || mode == 4) ? 0.0 : 1e-3)) {
        if (bad < 5) printf("  mode %d mismatch at (%zu,%zu): got %g want %g\n", mode, i / N, i % N, got, ref[i]);
        ++bad;
      }
    }
    printf("mode %d (%s): %d mismatches of %d, max err %g\n", mode,
           mode == 0 ? "i8 thread-written" : mode == 4 ? "i8, TMEM read with pack::16b (low half = even column)" : mode == 3 ? "i8 K=256+32 interleaved extras, N=224" : mode == 1 ? "bf16 thread-written" : "bf16 TMA", bad, M * N, maxerr);
    fails += bad != 0;
    cudaFree(dA);
    cudaFree(dB);
    cudaFree(dD);
  }
  printf(fails ? "UMMA SELFTEST FAILED\n" : "UMMA SELFTEST OK\n");
  return fails ? 1 : 0;
}
