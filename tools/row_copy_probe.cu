// Probe for a prefetch design: 6000 key-point rows of ~10 KB from pinned host memory as individual
// cudaMemcpyAsync calls -- device time, host time, alone and with a 640 MB bulk copy in flight on another stream -- against
// one gathered copy of the same bytes.
//   nvcc -O2 -o tools/_row_copy_probe tools/row_copy_probe.cu && tools/_row_copy_probe
#include <cuda_runtime.h>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e_)); return 1; } } while (0)

int main() {
  const size_t rows = 20000, row_bytes = 12000, n_pick = 6000;
  char *h, *d, *hb, *db, *stage;
  CK(cudaMallocHost(&h, rows * row_bytes));
  CK(cudaMalloc(&d, rows * row_bytes));
  const size_t bulk = 640u << 20;
  CK(cudaMallocHost(&hb, bulk));
  CK(cudaMalloc(&db, bulk));
  CK(cudaMallocHost(&stage, n_pick * row_bytes));
  for (size_t i = 0; i < rows * row_bytes; i += 4096) h[i] = (char)i;
  cudaStream_t s1, s2;
  CK(cudaStreamCreateWithFlags(&s1, cudaStreamNonBlocking));
  CK(cudaStreamCreateWithFlags(&s2, cudaStreamNonBlocking));
  std::vector<size_t> pick(n_pick);
  srand(1);
  for (size_t i = 0; i < n_pick; ++i) pick[i] = (size_t)rand() % rows;
  cudaEvent_t e0, e1, b0, b1;
  CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1)); CK(cudaEventCreate(&b0)); CK(cudaEventCreate(&b1));
  for (int busy = 0; busy < 2; ++busy)
    for (int rep = 0; rep < 3; ++rep) {
      CK(cudaDeviceSynchronize());
      if (busy) {
        CK(cudaEventRecord(b0, s2));
        CK(cudaMemcpyAsync(db, hb, bulk, cudaMemcpyHostToDevice, s2));
        CK(cudaEventRecord(b1, s2));
      }
      CK(cudaEventRecord(e0, s1));
      const auto t0 = std::chrono::steady_clock::now();
      for (size_t i = 0; i < n_pick; ++i)
        CK(cudaMemcpyAsync(d + pick[i] * row_bytes, h + pick[i] * row_bytes, row_bytes, cudaMemcpyHostToDevice, s1));
      const auto t1 = std::chrono::steady_clock::now();
      CK(cudaEventRecord(e1, s1));
      CK(cudaDeviceSynchronize());
      float ms = 0, msb = 0;
      CK(cudaEventElapsedTime(&ms, e0, e1));
      if (busy) CK(cudaEventElapsedTime(&msb, b0, b1));
      printf("%s: %zu row copies of %zu B = %.1f MB: device %.3f ms, host %.3f ms", busy ? "busy " : "quiet", n_pick, row_bytes,
             n_pick * row_bytes / 1e6, ms, std::chrono::duration<double, std::milli>(t1 - t0).count());
      if (busy) printf(", bulk 640 MB alongside %.3f ms", msb);
      printf("\n");
    }
  for (int rep = 0; rep < 3; ++rep) {
    CK(cudaDeviceSynchronize());
    const auto t0 = std::chrono::steady_clock::now();
    for (size_t i = 0; i < n_pick; ++i) memcpy(stage + i * row_bytes, h + pick[i] * row_bytes, row_bytes);
    const auto t1 = std::chrono::steady_clock::now();
    CK(cudaEventRecord(e0, s1));
    CK(cudaMemcpyAsync(d, stage, n_pick * row_bytes, cudaMemcpyHostToDevice, s1));
    CK(cudaEventRecord(e1, s1));
    CK(cudaDeviceSynchronize());
    float ms = 0;
    CK(cudaEventElapsedTime(&ms, e0, e1));
    printf("gathered: host memcpy of the rows %.3f ms, one copy %.3f ms\n", std::chrono::duration<double, std::milli>(t1 - t0).count(), ms);
  }
  return 0;
}
