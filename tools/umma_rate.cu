// Microbenchmark: cycles per tcgen05.mma (cta_group::1, SS operands) for kind::i8 / kind::f16 / kind::f8f6f4 at
// several N, one CTA per SM on all SMs.  Developer tool:  nvcc -gencode arch=compute_100a,code=sm_100a -O2 ...
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>
#include "../multi_robot_slam_separators_b200/csrc/umma.cuh"

__device__ __forceinline__ void mma_f8(uint32_t d, uint64_t a, uint64_t b, uint32_t idesc, bool acc) {
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
               "tcgen05.mma.cta_group::1.kind::f8f6f4 [%0], %1, %2, %3, p;\n\t}" ::"r"(d), "l"(a), "l"(b), "r"(idesc),
               "r"((uint32_t)acc) : "memory");
}

// kind 0: i8 (s32 acc), 1: bf16 (f32 acc), 2: e4m3 (f32 acc)
__global__ void __launch_bounds__(128) rate_kernel(int kind, int N, int iters, long long* out) {
  extern __shared__ uint8_t raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)raw + 1023) & ~(uintptr_t)1023);
  __shared__ uint64_t bar;
  __shared__ uint32_t slot;
  const int tid = threadIdx.x, warp = tid >> 5;
  for (int i = tid; i < (128 + 256) * 256 / 4; i += 128) ((uint32_t*)smem)[i] = 0x01010101u * (i & 1);
  if (tid == 0) { umma::mbar_init(&bar, 1); umma::fence_mbar_init(); }
  if (warp == 0) umma::tmem_alloc(&slot, 512);
  umma::fence_proxy_async();
  umma::tc_fence_before();
  __syncthreads();
  umma::tc_fence_after();
  const uint32_t tmem = slot;
  long long t0 = 0, t1 = 0;
  if (warp == 0) {
    const uint32_t a0 = umma::smem_u32(smem), b0 = a0 + 128 * 256;
    const uint32_t idesc = kind == 0 ? umma::instr_desc(2, 1, 1, 128, N) : kind == 1 ? umma::instr_desc(1, 1, 1, 128, N)
                                                                                     : umma::instr_desc(1, 0, 0, 128, N);
    t0 = clock64();
    if (umma::elect_one()) {
      for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          const uint64_t ad = umma::smem_desc_sw128(a0 + (k >> 2) * 128 * 128 + (k & 3) * 32);
          const uint64_t bd = umma::smem_desc_sw128(b0 + (k >> 2) * 256 * 128 + (k & 3) * 32);
          const uint32_t d = tmem + (uint32_t)((i & 1) * 256);
          if (kind == 0) umma::mma_i8(d, ad, bd, idesc, k != 0);
          else if (kind == 1) umma::mma_f16(d, ad, bd, idesc, k != 0);
          else mma_f8(d, ad, bd, idesc, k != 0);
        }
      }
      umma::mma_commit(&bar);
    }
    __syncwarp();
    umma::mbar_wait(&bar, 0);
    t1 = clock64();
  }
  umma::tc_fence_before();
  __syncthreads();
  if (warp == 0) umma::tmem_dealloc(tmem, 512);
  if (tid == 0) out[blockIdx.x] = t1 - t0;
}

int main() {
  cudaSetDevice(0);
  long long* d;
  cudaMalloc(&d, 148 * 8);
  const size_t smem = (128 + 256) * 256 + 1024;
  cudaFuncSetAttribute(rate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  const char* names[3] = {"i8", "bf16", "e4m3"};
  for (int grid : {1, 148})
    for (int kind = 0; kind < 3; ++kind)
      for (int N : {64, 128, 224, 256}) {
        const int iters = 2000;
        rate_kernel<<<grid, 128, smem>>>(kind, N, iters, d);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
        long long h[148];
        cudaMemcpy(h, d, grid * 8, cudaMemcpyDeviceToHost);
        long long mx = 0;
        for (int i = 0; i < grid; ++i) mx = h[i] > mx ? h[i] : mx;
        const double per = (double)mx / (iters * 8.0);
        const int kel = kind == 1 ? 16 : 32;
        printf("grid %3d %-5s M=128 N=%3d: %.1f clk/MMA -> %.0f MAC/clk/SM\n", grid, names[kind], N, per, 128.0 * N * kel / per);
      }
  return 0;
}
