// C ABI: context, keyframe feature store, host-buffer wrappers around the kernel launchers.
#include "common.cuh"

#include <new>

int sepb_launch_match(sepb_ctx* ctx, FramesView from, FramesView to, const int32_t* d_pairs, int n_pairs, float nndr,
                      int strict, int32_t* d_knn_idx, int32_t* d_knn_dist, int32_t* d_n_matches,
                      uint16_t* d_match_from, uint16_t* d_match_to, int32_t* d_n_unique_to, const LazyPull* lazy) {
  const int opt = ctx->opt_match_kernel;
  const bool tensor = opt == SEPB_MATCH_TENSOR || (opt == SEPB_MATCH_AUTO && sepb_hamming_mma_supported(from, to));
  if (tensor)
    return sepb_launch_hamming_mma(ctx, from, to, d_pairs, n_pairs, nndr, strict, d_knn_idx, d_knn_dist, d_n_matches,
                                   d_match_from, d_match_to, d_n_unique_to, lazy);
  return sepb_launch_hamming_match(ctx, from, to, d_pairs, n_pairs, nndr, strict, d_knn_idx, d_knn_dist, d_n_matches,
                                   d_match_from, d_match_to, d_n_unique_to);
}

extern "C" {

static int ctx_create_impl(int device, bool high_priority, sepb_ctx** out);

int sepb_ctx_create(int device, sepb_ctx** out) { return ctx_create_impl(device, false, out); }
int sepb_ctx_create_priority(int device, sepb_ctx** out) { return ctx_create_impl(device, true, out); }

}  // extern "C"

static int ctx_create_impl(int device, bool high_priority, sepb_ctx** out) {
  if (!out) return SEPB_E_INVALID;
  *out = nullptr;
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess) return (int)e;  // no CPU fallback: fail loudly
  if (device < 0 || device >= n) return SEPB_E_INVALID;
  e = cudaSetDevice(device);
  if (e != cudaSuccess) return (int)e;
  cudaDeviceProp prop;
  e = cudaGetDeviceProperties(&prop, device);
  if (e != cudaSuccess) return (int)e;
  if (prop.major != 10) return (int)cudaErrorNoKernelImageForDevice;  // sm_100a only
  sepb_ctx* c = new (std::nothrow) sepb_ctx();
  if (!c) return (int)cudaErrorMemoryAllocation;
  c->device = device;
  c->sm_count = prop.multiProcessorCount;
  int prio_lo = 0, prio_hi = 0;   // numerically lower = more urgent
  if (high_priority) cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi);
  e = cudaStreamCreateWithPriority(&c->stream, cudaStreamNonBlocking, high_priority ? prio_hi : 0);
  if (e == cudaSuccess) e = cudaStreamCreateWithPriority(&c->copy_stream, cudaStreamNonBlocking, high_priority ? prio_hi : 0);
  if (e != cudaSuccess) {
    if (c->stream) cudaStreamDestroy(c->stream);
    delete c;
    return (int)e;
  }
  *out = c;
  return SEPB_OK;
}

extern "C" {

void sepb_ctx_destroy(sepb_ctx* ctx) {
  if (!ctx) return;
  cudaSetDevice(ctx->device);
  cudaStreamSynchronize(ctx->stream);
  cudaStreamSynchronize(ctx->copy_stream);
  cudaStreamDestroy(ctx->copy_stream);
  if (ctx->scratch) cudaFree(ctx->scratch);
  if (ctx->io) cudaFree(ctx->io);
  if (ctx->table) cudaFree(ctx->table);
  if (ctx->table4) cudaFree(ctx->table4);
  if (ctx->k3_order) cudaFree(ctx->k3_order);
  cudaStreamDestroy(ctx->stream);
  delete ctx;
}

const char* sepb_last_error(const sepb_ctx* ctx) { return ctx ? ctx->err.c_str() : "null context"; }
void* sepb_ctx_stream(const sepb_ctx* ctx) { return ctx ? (void*)ctx->stream : nullptr; }
int64_t sepb_ctx_launch_count(const sepb_ctx* ctx) { return ctx ? ctx->launches : 0; }

int sepb_ctx_sync(sepb_ctx* ctx) {
  if (!ctx) return SEPB_E_INVALID;
  SEPB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return SEPB_OK;
}

int sepb_ctx_set_option(sepb_ctx* ctx, int option, int value) {
  if (!ctx) return SEPB_E_INVALID;
  if (option == SEPB_OPT_MATCH_KERNEL) {
    SEPB_REQUIRE(ctx, value == SEPB_MATCH_AUTO || value == SEPB_MATCH_POPC || value == SEPB_MATCH_TENSOR, SEPB_E_INVALID);
    ctx->opt_match_kernel = value;
    return SEPB_OK;
  }
  if (option == SEPB_OPT_PNP_SCHEDULER) {
    SEPB_REQUIRE(ctx, value == SEPB_PNP_SCHED_PERSISTENT || value == SEPB_PNP_SCHED_PER_PAIR, SEPB_E_INVALID);
    ctx->opt_pnp_sched = value;
    return SEPB_OK;
  }
  if (option == SEPB_OPT_MATCH_TILE) {
    SEPB_REQUIRE(ctx, value == 0 || value == 128 || value == 256, SEPB_E_INVALID);
    ctx->opt_match_tile = value == 256 ? 0 : value;
    return SEPB_OK;
  }
  if (option == SEPB_OPT_MATCH_SCHEDULER) {
    SEPB_REQUIRE(ctx, value == SEPB_MATCH_SCHED_PERSISTENT || value == SEPB_MATCH_SCHED_PER_PAIR, SEPB_E_INVALID);
    ctx->opt_match_sched = value;
    return SEPB_OK;
  }
  return sepb_fail(ctx, SEPB_E_INVALID, "unknown option", __FILE__, __LINE__);
}

int sepb_ctx_get_option(const sepb_ctx* ctx, int option, int* value) {
  if (!ctx || !value) return SEPB_E_INVALID;
  if (option == SEPB_OPT_MATCH_KERNEL) {
    *value = ctx->opt_match_kernel;
    return SEPB_OK;
  }
  if (option == SEPB_OPT_PNP_SCHEDULER) {
    *value = ctx->opt_pnp_sched;
    return SEPB_OK;
  }
  if (option == SEPB_OPT_MATCH_TILE) {
    *value = ctx->opt_match_tile;
    return SEPB_OK;
  }
  if (option == SEPB_OPT_MATCH_SCHEDULER) {
    *value = ctx->opt_match_sched;
    return SEPB_OK;
  }
  return SEPB_E_INVALID;
}

// ---- frame store ---------------------------------------------------------------------------
int sepb_frames_create(sepb_ctx* ctx, int capacity, int max_feat, sepb_frames** out) {
  if (!ctx || !out) return SEPB_E_INVALID;
  *out = nullptr;
  SEPB_REQUIRE(ctx, capacity > 0 && max_feat > 0 && max_feat <= SEPB_MAX_FEATURES, SEPB_E_SHAPE);
  SEPB_CUDA(ctx, cudaSetDevice(ctx->device));
  sepb_frames* f = new (std::nothrow) sepb_frames();
  if (!f) return (int)cudaErrorMemoryAllocation;
  f->ctx = ctx;
  f->capacity = capacity;
  f->max_feat = max_feat;
  const size_t nf = (size_t)capacity * max_feat;
  cudaError_t e;
  if ((e = cudaMalloc(&f->desc, nf * 32)) != cudaSuccess || (e = cudaMalloc(&f->kp2d, nf * 8)) != cudaSuccess ||
      (e = cudaMalloc(&f->kp3d, nf * 12)) != cudaSuccess || (e = cudaMalloc(&f->octave, nf * 4)) != cudaSuccess ||
      (e = cudaMalloc(&f->count, (size_t)capacity * 4)) != cudaSuccess ||
      (e = cudaMalloc(&f->need, ((size_t)capacity + 3) / 4 * 4)) != cudaSuccess ||
      (e = cudaMalloc(&f->have, ((size_t)capacity + 3) / 4 * 4)) != cudaSuccess ||
      (e = cudaMalloc(&f->fetched, 8)) != cudaSuccess) {
    sepb_frames_destroy(f);
    return sepb_fail(ctx, (int)e, "cudaMalloc(frames)", __FILE__, __LINE__);
  }
  if ((e = cudaEventCreateWithFlags(&f->ev_uploaded, cudaEventDisableTiming)) != cudaSuccess ||
      (e = cudaEventCreateWithFlags(&f->ev_last_use, cudaEventDisableTiming)) != cudaSuccess) {
    sepb_frames_destroy(f);
    return sepb_fail(ctx, (int)e, "cudaEventCreate(frames)", __FILE__, __LINE__);
  }
  cudaMemsetAsync(f->count, 0, (size_t)capacity * 4, ctx->stream);
  cudaMemsetAsync(f->octave, 0, nf * 4, ctx->stream);
  cudaMemsetAsync(f->need, 0, ((size_t)capacity + 3) / 4 * 4, ctx->stream);
  cudaMemsetAsync(f->have, 0, ((size_t)capacity + 3) / 4 * 4, ctx->stream);
  cudaMemsetAsync(f->fetched, 0, 8, ctx->stream);
  cudaStreamSynchronize(ctx->stream);   // uploads run on the copy stream: the store must be initialised before the first one
  *out = f;
  return SEPB_OK;
}

void sepb_frames_destroy(sepb_frames* f) {
  if (!f) return;
  if (f->ctx) {
    cudaSetDevice(f->ctx->device);
    cudaStreamSynchronize(f->ctx->stream);
    cudaStreamSynchronize(f->ctx->copy_stream);
  }
  if (f->ev_uploaded) cudaEventDestroy(f->ev_uploaded);
  if (f->ev_last_use) cudaEventDestroy(f->ev_last_use);
  cudaFree(f->desc);
  cudaFree(f->kp2d);
  cudaFree(f->kp3d);
  cudaFree(f->octave);
  cudaFree(f->count);
  cudaFree(f->need);
  cudaFree(f->have);
  cudaFree(f->fetched);
  delete f;
}

int sepb_frames_size(const sepb_frames* f) { return f ? f->size : 0; }

// lazy == true (sepb_frames_bind_host_async): only descriptors and counts are copied; kp2d / kp3d / octave stay in the
// caller's pinned arrays, whose device-visible addresses are remembered for frames_fetch.cu
static int frames_upload_impl(sepb_ctx* ctx, sepb_frames* f, int first, int n, const uint8_t* desc, const float* kp2d,
                              const float* kp3d, const int32_t* octave, const int32_t* count, bool async, bool lazy = false) {
  if (!ctx || !f) return SEPB_E_INVALID;
  SEPB_REQUIRE(ctx, n >= 0 && first >= 0 && first + n <= f->capacity, SEPB_E_SHAPE);
  if (n == 0) return SEPB_OK;
  SEPB_REQUIRE(ctx, desc && kp2d && kp3d && count, SEPB_E_INVALID);
  for (int i = 0; i < n; ++i) SEPB_REQUIRE(ctx, count[i] >= 0 && count[i] <= f->max_feat, SEPB_E_SHAPE);
  std::lock_guard<std::mutex> lk(ctx->mu);
  SEPB_CUDA(ctx, cudaSetDevice(ctx->device));
  const float *d_kp2d = nullptr, *d_kp3d = nullptr;
  const int32_t* d_oct = nullptr;
  if (lazy) {
    // the arrays must be pinned (cudaHostAlloc / cudaHostRegister): the device reads them in place
    if (cudaHostGetDevicePointer((void**)&d_kp2d, (void*)kp2d, 0) != cudaSuccess ||
        cudaHostGetDevicePointer((void**)&d_kp3d, (void*)kp3d, 0) != cudaSuccess ||
        (octave && cudaHostGetDevicePointer((void**)&d_oct, (void*)octave, 0) != cudaSuccess)) {
      cudaGetLastError();
      return sepb_fail(ctx, SEPB_E_INVALID, "sepb_frames_bind_host_async: kp2d / kp3d / octave must be pinned host memory", __FILE__, __LINE__);
    }
  }
  // the copy stream must not overwrite slots a kernel enqueued earlier is still reading
  cudaStream_t cs = ctx->copy_stream;
  if (f->has_use) SEPB_CUDA(ctx, cudaStreamWaitEvent(cs, f->ev_last_use, 0));
  const size_t off = (size_t)first * f->max_feat, nf = (size_t)n * f->max_feat;
  SEPB_CUDA(ctx, cudaMemcpyAsync(f->desc + off * 32, desc, nf * 32, cudaMemcpyHostToDevice, cs));
  if (!lazy) {
    SEPB_CUDA(ctx, cudaMemcpyAsync(f->kp2d + off * 2, kp2d, nf * 8, cudaMemcpyHostToDevice, cs));
    SEPB_CUDA(ctx, cudaMemcpyAsync(f->kp3d + off * 3, kp3d, nf * 12, cudaMemcpyHostToDevice, cs));
    if (octave) SEPB_CUDA(ctx, cudaMemcpyAsync(f->octave + off, octave, nf * 4, cudaMemcpyHostToDevice, cs));
    SEPB_CUDA(ctx, cudaMemsetAsync(f->have + first, 7, (size_t)n, cs));   // everything resident: nothing to pull
  } else {
    f->lazy = true;
    f->lazy_first = first;
    f->lazy_n = n;
    f->h_kp2d = d_kp2d;
    f->h_kp3d = d_kp3d;
    f->h_octave = d_oct;
    SEPB_CUDA(ctx, cudaMemsetAsync(f->have, octave ? 0 : 4, ((size_t)f->capacity + 3) / 4 * 4, cs));
    SEPB_CUDA(ctx, cudaMemsetAsync(f->need, 0, ((size_t)f->capacity + 3) / 4 * 4, cs));
    SEPB_CUDA(ctx, cudaMemsetAsync(f->fetched, 0, 8, cs));
  }
  SEPB_CUDA(ctx, cudaMemcpyAsync(f->count + first, count, (size_t)n * 4, cudaMemcpyHostToDevice, cs));
  SEPB_CUDA(ctx, cudaEventRecord(f->ev_uploaded, cs));
  f->has_upload = true;
  if (!async) SEPB_CUDA(ctx, cudaStreamSynchronize(cs));
  if (first + n > f->size) f->size = first + n;
  return SEPB_OK;
}

int sepb_frames_upload(sepb_ctx* ctx, sepb_frames* f, int first, int n, const uint8_t* desc, const float* kp2d,
                       const float* kp3d, const int32_t* octave, const int32_t* count) {
  return frames_upload_impl(ctx, f, first, n, desc, kp2d, kp3d, octave, count, false);
}

int sepb_frames_upload_async(sepb_ctx* ctx, sepb_frames* f, int first, int n, const uint8_t* desc, const float* kp2d,
                             const float* kp3d, const int32_t* octave, const int32_t* count) {
  return frames_upload_impl(ctx, f, first, n, desc, kp2d, kp3d, octave, count, true);
}

int sepb_frames_bind_host_async(sepb_ctx* ctx, sepb_frames* f, int first, int n, const uint8_t* desc, const float* kp2d,
                                const float* kp3d, const int32_t* octave, const int32_t* count) {
  return frames_upload_impl(ctx, f, first, n, desc, kp2d, kp3d, octave, count, true, true);
}

int64_t sepb_frames_fetched_bytes(sepb_ctx* ctx, const sepb_frames* f) {
  if (!ctx || !f || !f->fetched) return -1;
  unsigned long long v = 0;
  cudaSetDevice(ctx->device);
  if (cudaMemcpyAsync(&v, f->fetched, 8, cudaMemcpyDeviceToHost, ctx->stream) != cudaSuccess) return -1;
  if (cudaStreamSynchronize(ctx->stream) != cudaSuccess) return -1;
  return (int64_t)v;
}

int sepb_frames_wait(sepb_ctx* ctx, sepb_frames* f) {
  if (!ctx || !f) return SEPB_E_INVALID;
  SEPB_CUDA(ctx, cudaSetDevice(ctx->device));
  if (f->has_upload) SEPB_CUDA(ctx, cudaEventSynchronize(f->ev_uploaded));
  return SEPB_OK;
}

// ---- K2 ------------------------------------------------------------------------------------
int sepb_match_pairs_dev(sepb_ctx* ctx, const sepb_frames* from, const sepb_frames* to, const int32_t* d_pairs,
                         int n_pairs, const sepb_match_params* prm, int32_t* d_knn_idx, int32_t* d_knn_dist,
                         int32_t* d_n_matches, uint16_t* d_match_from, uint16_t* d_match_to,
                         int32_t* d_n_unique_to) {
  if (!ctx || !from || !to || !prm) return SEPB_E_INVALID;
  SEPB_REQUIRE(ctx, n_pairs >= 0, SEPB_E_SHAPE);
  if (n_pairs == 0) return SEPB_OK;
  SEPB_REQUIRE(ctx, d_pairs && d_n_matches, SEPB_E_INVALID);
  SEPB_CUDA(ctx, cudaSetDevice(ctx->device));
  sepb_frames_acquire(ctx, from, to);
  const int rc = sepb_launch_match(ctx, view_of(from), view_of(to), d_pairs, n_pairs, prm->nndr, prm->nndr_strict,
                                   d_knn_idx, d_knn_dist, d_n_matches, d_match_from, d_match_to, d_n_unique_to);
  sepb_frames_release(ctx, from, to);
  return rc;
}

int sepb_match_pairs(sepb_ctx* ctx, const sepb_frames* from, const sepb_frames* to, const int32_t* pairs, int n_pairs,
                     const sepb_match_params* prm, int32_t* knn_idx, int32_t* knn_dist, int32_t* n_matches,
                     uint16_t* match_from, uint16_t* match_to, int32_t* n_unique_to) {
  if (!ctx || !from || !to || !prm) return SEPB_E_INVALID;
  SEPB_REQUIRE(ctx, n_pairs >= 0, SEPB_E_SHAPE);
  if (n_pairs == 0) return SEPB_OK;
  SEPB_REQUIRE(ctx, pairs && n_matches, SEPB_E_INVALID);
  for (int i = 0; i < n_pairs; ++i)
    SEPB_REQUIRE(ctx, pairs[2 * i] >= 0 && pairs[2 * i] < from->size && pairs[2 * i + 1] >= 0 &&
                          pairs[2 * i + 1] < to->size, SEPB_E_SHAPE);
  std::lock_guard<std::mutex> lk(ctx->mu);
  SEPB_CUDA(ctx, cudaSetDevice(ctx->device));
  const size_t P = (size_t)n_pairs;
  const size_t b_pairs = align_up(P * 8, 256), b_knn = align_up(P * to->max_feat * 8, 256),
               b_n = align_up(P * 4, 256), b_m = align_up(P * from->max_feat * 2, 256);
  const size_t total = b_pairs + (knn_idx ? b_knn : 0) + (knn_dist ? b_knn : 0) + 2 * b_n + 2 * b_m;
  uint8_t* s = (uint8_t*)sepb_scratch(ctx, total);
  if (!s) return (int)cudaErrorMemoryAllocation;
  int32_t* d_pairs = (int32_t*)s; s += b_pairs;
  int32_t* d_ki = nullptr; if (knn_idx) { d_ki = (int32_t*)s; s += b_knn; }
  int32_t* d_kd = nullptr; if (knn_dist) { d_kd = (int32_t*)s; s += b_knn; }
  int32_t* d_nm = (int32_t*)s; s += b_n;
  int32_t* d_nu = (int32_t*)s; s += b_n;
  uint16_t* d_mf = (uint16_t*)s; s += b_m;
  uint16_t* d_mt = (uint16_t*)s; s += b_m;
  SEPB_CUDA(ctx, cudaMemcpyAsync(d_pairs, pairs, P * 8, cudaMemcpyHostToDevice, ctx->stream));
  if (d_ki) SEPB_CUDA(ctx, cudaMemsetAsync(d_ki, 0xFF, P * to->max_feat * 8, ctx->stream));
  if (d_kd) SEPB_CUDA(ctx, cudaMemsetAsync(d_kd, 0xFF, P * to->max_feat * 8, ctx->stream));
  sepb_frames_acquire(ctx, from, to);
  int rc = sepb_launch_match(ctx, view_of(from), view_of(to), d_pairs, n_pairs, prm->nndr, prm->nndr_strict,
                                     d_ki, d_kd, d_nm, d_mf, d_mt, d_nu);
  sepb_frames_release(ctx, from, to);
  if (rc) return rc;
  if (knn_idx) SEPB_CUDA(ctx, cudaMemcpyAsync(knn_idx, d_ki, P * to->max_feat * 8, cudaMemcpyDeviceToHost, ctx->stream));
  if (knn_dist) SEPB_CUDA(ctx, cudaMemcpyAsync(knn_dist, d_kd, P * to->max_feat * 8, cudaMemcpyDeviceToHost, ctx->stream));
  SEPB_CUDA(ctx, cudaMemcpyAsync(n_matches, d_nm, P * 4, cudaMemcpyDeviceToHost, ctx->stream));
  if (n_unique_to) SEPB_CUDA(ctx, cudaMemcpyAsync(n_unique_to, d_nu, P * 4, cudaMemcpyDeviceToHost, ctx->stream));
  if (match_from) SEPB_CUDA(ctx, cudaMemcpyAsync(match_from, d_mf, P * from->max_feat * 2, cudaMemcpyDeviceToHost, ctx->stream));
  if (match_to) SEPB_CUDA(ctx, cudaMemcpyAsync(match_to, d_mt, P * from->max_feat * 2, cudaMemcpyDeviceToHost, ctx->stream));
  SEPB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return SEPB_OK;
}

}  // extern "C"
