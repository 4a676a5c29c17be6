// C ABI: NetVLAD database + nearest-neighbour entry points (K1 / K1b) and the selection loop.
#include "common.cuh"

#include <algorithm>
#include <new>
#include <numeric>
#include <vector>

// netvlad_gemm.cu
int sepb_netvlad_tensor_supported(int dim, int k);
int sepb_netvlad_tensor_topk(sepb_ctx* ctx, const sepb_db* db, const double* d_q, int nq, int k,
                             const uint8_t* d_row_mask, const int32_t* d_ign_start, const int32_t* d_ign_col,
                             int32_t* d_out_idx, double* d_out_dist, float* d_dbg_v, int32_t* d_stats);

namespace {
int exact_splits(const sepb_ctx* ctx, int nq, int n) {
  const int row_blocks = (nq + 127) / 128;
  int s = (2 * ctx->sm_count + row_blocks - 1) / row_blocks;
  const int tiles = (n + 31) / 32;
  s = std::max(1, std::min(std::min(s, 64), tiles));
  return s;
}
}  // namespace

extern "C" {

int sepb_db_create(sepb_ctx* ctx, int dim, int capacity, sepb_db** out) {
  if (!ctx || !out) return SEPB_E_INVALID;
  *out = nullptr;
  SEPB_REQUIRE(ctx, dim > 0 && dim <= 65536 && capacity > 0, SEPB_E_SHAPE);
  SEPB_CUDA(ctx, cudaSetDevice(ctx->device));
  sepb_db* db = new (std::nothrow) sepb_db();
  if (!db) return (int)cudaErrorMemoryAllocation;
  db->ctx = ctx;
  db->dim = dim;
  db->capacity = capacity;
  cudaError_t e;
  if ((e = cudaMalloc(&db->rows, (size_t)capacity * dim * 8)) != cudaSuccess ||
      (e = cudaMalloc(&db->col_mask, (size_t)capacity)) != cudaSuccess) {
    sepb_db_destroy(db);
    return sepb_fail(ctx, (int)e, "cudaMalloc(db)", __FILE__, __LINE__);
  }
  cudaMemsetAsync(db->col_mask, 0, capacity, ctx->stream);
  *out = db;
  return SEPB_OK;
}

void sepb_db_destroy(sepb_db* db) {
  if (!db) return;
  if (db->ctx) {
    cudaSetDevice(db->ctx->device);
    cudaStreamSynchronize(db->ctx->stream);
  }
  cudaFree(db->rows);
  cudaFree(db->col_mask);
  cudaFree(db->split);
  cudaFree(db->norm2);
  cudaFree(db->maxnorm2);
  delete db;
}

int sepb_db_size(const sepb_db* db) { return db ? db->size : 0; }

int sepb_db_clear(sepb_db* db) {
  if (!db) return SEPB_E_INVALID;
  db->size = 0;
  db->split_rows = 0;
  db->has_mask = false;
  return SEPB_OK;
}

int sepb_db_set_base(sepb_db* db, int64_t base) {
  if (!db || base < 0) return SEPB_E_INVALID;
  db->base = base;
  return SEPB_OK;
}

int sepb_db_append(sepb_ctx* ctx, sepb_db* db, const double* rows, int n) {
  if (!ctx || !db) return SEPB_E_INVALID;
  SEPB_REQUIRE(ctx, n >= 0 && db->size + n <= db->capacity, SEPB_E_SHAPE);
  if (n == 0) return SEPB_OK;
  SEPB_REQUIRE(ctx, rows != nullptr, SEPB_E_INVALID);
  std::lock_guard<std::mutex> lk(ctx->mu);
  SEPB_CUDA(ctx, cudaSetDevice(ctx->device));
  SEPB_CUDA(ctx, cudaMemcpyAsync(db->rows + (size_t)db->size * db->dim, rows, (size_t)n * db->dim * 8,
                                 cudaMemcpyHostToDevice, ctx->stream));
  SEPB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  db->size += n;
  return SEPB_OK;
}

int sepb_db_set_col_mask(sepb_ctx* ctx, sepb_db* db, const uint8_t* mask) {
  if (!ctx || !db) return SEPB_E_INVALID;
  std::lock_guard<std::mutex> lk(ctx->mu);
  SEPB_CUDA(ctx, cudaSetDevice(ctx->device));
  if (!mask) {
    SEPB_CUDA(ctx, cudaMemsetAsync(db->col_mask, 0, db->capacity, ctx->stream));
    db->has_mask = false;
  } else {
    SEPB_CUDA(ctx, cudaMemcpyAsync(db->col_mask, mask, db->size, cudaMemcpyHostToDevice, ctx->stream));
    db->has_mask = true;
  }
  SEPB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return SEPB_OK;
}

int sepb_netvlad_topk_dev(sepb_ctx* ctx, const sepb_db* db, const double* d_q, int nq, int k, int mode,
                          const uint8_t* d_row_mask, const int32_t* d_ign_start, const int32_t* d_ign_col,
                          int32_t* d_out_idx, double* d_out_dist) {
  if (!ctx || !db) return SEPB_E_INVALID;
  SEPB_REQUIRE(ctx, nq >= 0 && k >= 1 && k <= 16, SEPB_E_SHAPE);
  SEPB_REQUIRE(ctx, mode == SEPB_NN_EXACT || mode == SEPB_NN_TENSOR, SEPB_E_INVALID);
  if (nq == 0) return SEPB_OK;
  SEPB_REQUIRE(ctx, d_q && d_out_idx && d_out_dist, SEPB_E_INVALID);
  SEPB_CUDA(ctx, cudaSetDevice(ctx->device));
  // the filter's per-row lists hold k <= 8; larger k takes the exact scan (same results either way)
  if (mode == SEPB_NN_TENSOR && sepb_netvlad_tensor_supported(db->dim, k))
    return sepb_netvlad_tensor_topk(ctx, db, d_q, nq, k, d_row_mask, d_ign_start, d_ign_col, d_out_idx, d_out_dist,
                                    nullptr, nullptr);
  const int splits = exact_splits(ctx, nq, db->size);
  int32_t* part_i = nullptr;
  double* part_d = nullptr;
  if (splits > 1) {
    const size_t b_d = align_up((size_t)splits * nq * k * 8, 256), b_i = align_up((size_t)splits * nq * k * 4, 256);
    uint8_t* s = (uint8_t*)sepb_scratch(ctx, b_d + b_i);
    if (!s) return (int)cudaErrorMemoryAllocation;
    part_d = (double*)s;
    part_i = (int32_t*)(s + b_d);
  }
  return sepb_launch_netvlad_exact(ctx, d_q, nq, db->rows, db->size, db->dim, k, db->base, d_row_mask,
                                   db->has_mask ? db->col_mask : nullptr, d_ign_start, d_ign_col, splits, part_i, part_d,
                                   d_out_idx, d_out_dist);
}

int sepb_netvlad_topk(sepb_ctx* ctx, const sepb_db* db, const double* queries, int nq, int k, int mode,
                      const uint8_t* row_mask, const int32_t* ignored, int n_ignored, int32_t* out_idx,
                      double* out_dist) {
  if (!ctx || !db) return SEPB_E_INVALID;
  SEPB_REQUIRE(ctx, nq >= 0 && k >= 1 && k <= 16 && n_ignored >= 0, SEPB_E_SHAPE);
  if (nq == 0) return SEPB_OK;
  SEPB_REQUIRE(ctx, queries && out_idx && out_dist && (n_ignored == 0 || ignored), SEPB_E_INVALID);
  // CSR of ignored cells by query row
  std::vector<int32_t> start, cols;
  if (n_ignored > 0) {
    start.assign(nq + 1, 0);
    for (int i = 0; i < n_ignored; ++i) {
      SEPB_REQUIRE(ctx, ignored[2 * i] >= 0 && ignored[2 * i] < nq, SEPB_E_SHAPE);
      start[ignored[2 * i] + 1]++;
    }
    std::partial_sum(start.begin(), start.end(), start.begin());
    cols.resize(n_ignored);
    std::vector<int32_t> fill(start.begin(), start.end() - 1);
    for (int i = 0; i < n_ignored; ++i) cols[fill[ignored[2 * i]]++] = ignored[2 * i + 1];
  }
  int32_t* d_idx = nullptr;
  {
    std::lock_guard<std::mutex> lk(ctx->mu);
    SEPB_CUDA(ctx, cudaSetDevice(ctx->device));
    // staging lives in the context's second grow-only buffer (the launch underneath uses ctx scratch); no
    // cudaMalloc / cudaFree per call: those synchronise the whole device, including uploads in flight
    const size_t b_q = align_up((size_t)nq * db->dim * 8, 256), b_rm = align_up((size_t)nq, 256),
                 b_st = align_up((size_t)(nq + 1) * 4, 256), b_co = align_up((size_t)std::max(n_ignored, 1) * 4, 256),
                 b_oi = align_up((size_t)nq * k * 4, 256), b_od = align_up((size_t)nq * k * 8, 256);
    uint8_t* buf = (uint8_t*)sepb_scratch_io(ctx, b_q + b_rm + b_st + b_co + b_oi + b_od);
    if (!buf) return (int)cudaErrorMemoryAllocation;
    uint8_t* s = buf;
    double* d_q = (double*)s; s += b_q;
    uint8_t* d_rm = (uint8_t*)s; s += b_rm;
    int32_t* d_st = (int32_t*)s; s += b_st;
    int32_t* d_co = (int32_t*)s; s += b_co;
    d_idx = (int32_t*)s; s += b_oi;
    double* d_dist = (double*)s;
    int rc = SEPB_OK;
    auto fail = [&](cudaError_t e) { return sepb_fail(ctx, (int)e, "netvlad_topk copy", __FILE__, __LINE__); };
    cudaError_t e;
    if ((e = cudaMemcpyAsync(d_q, queries, (size_t)nq * db->dim * 8, cudaMemcpyHostToDevice, ctx->stream)) != cudaSuccess) return fail(e);
    if (row_mask && (e = cudaMemcpyAsync(d_rm, row_mask, nq, cudaMemcpyHostToDevice, ctx->stream)) != cudaSuccess) return fail(e);
    if (n_ignored > 0) {
      if ((e = cudaMemcpyAsync(d_st, start.data(), (size_t)(nq + 1) * 4, cudaMemcpyHostToDevice, ctx->stream)) != cudaSuccess) return fail(e);
      if ((e = cudaMemcpyAsync(d_co, cols.data(), (size_t)n_ignored * 4, cudaMemcpyHostToDevice, ctx->stream)) != cudaSuccess) return fail(e);
    }
    rc = sepb_netvlad_topk_dev(ctx, db, d_q, nq, k, mode, row_mask ? d_rm : nullptr, n_ignored ? d_st : nullptr,
                               n_ignored ? d_co : nullptr, d_idx, d_dist);
    if (rc == SEPB_OK) {
      if ((e = cudaMemcpyAsync(out_idx, d_idx, (size_t)nq * k * 4, cudaMemcpyDeviceToHost, ctx->stream)) != cudaSuccess) return fail(e);
      if ((e = cudaMemcpyAsync(out_dist, d_dist, (size_t)nq * k * 8, cudaMemcpyDeviceToHost, ctx->stream)) != cudaSuccess) return fail(e);
    }
    e = cudaStreamSynchronize(ctx->stream);
    if (rc) return rc;
    if (e != cudaSuccess) return sepb_fail(ctx, (int)e, "netvlad_topk sync", __FILE__, __LINE__);
  }
  return SEPB_OK;
}

// test / profiling hook of the tensor-core filter: same search as SEPB_NN_TENSOR plus the raw filter values
int sepb_netvlad_filter_debug(sepb_ctx* ctx, const sepb_db* db, const double* queries, int nq, int k, float* out_v,
                              int32_t* out_idx, double* out_dist, int32_t* stats) {
  if (!ctx || !db) return SEPB_E_INVALID;
  SEPB_REQUIRE(ctx, nq > 0 && k >= 1 && sepb_netvlad_tensor_supported(db->dim, k) && db->size > 0, SEPB_E_SHAPE);
  SEPB_REQUIRE(ctx, queries && out_idx && out_dist, SEPB_E_INVALID);
  std::lock_guard<std::mutex> lk(ctx->mu);
  SEPB_CUDA(ctx, cudaSetDevice(ctx->device));
  const size_t n = db->size;
  const size_t b_q = align_up((size_t)nq * db->dim * 8, 256), b_v = align_up(out_v ? (size_t)nq * n * 4 : 4, 256),
               b_oi = align_up((size_t)nq * k * 4, 256), b_od = align_up((size_t)nq * k * 8, 256);
  uint8_t* buf = nullptr;
  SEPB_CUDA(ctx, cudaMalloc(&buf, b_q + b_v + b_oi + b_od + 256));
  double* d_q = (double*)buf;
  float* d_v = (float*)(buf + b_q);
  int32_t* d_idx = (int32_t*)(buf + b_q + b_v);
  double* d_dist = (double*)(buf + b_q + b_v + b_oi);
  int32_t* d_stats = (int32_t*)(buf + b_q + b_v + b_oi + b_od);
  cudaMemcpyAsync(d_q, queries, (size_t)nq * db->dim * 8, cudaMemcpyHostToDevice, ctx->stream);
  cudaMemsetAsync(d_stats, 0, 16, ctx->stream);
  if (out_v) cudaMemsetAsync(d_v, 0xFF, (size_t)nq * n * 4, ctx->stream);
  int rc = sepb_netvlad_tensor_topk(ctx, db, d_q, nq, k, nullptr, nullptr, nullptr, d_idx, d_dist, out_v ? d_v : nullptr,
                                    d_stats);
  if (rc == SEPB_OK) {
    if (out_v) cudaMemcpyAsync(out_v, d_v, (size_t)nq * n * 4, cudaMemcpyDeviceToHost, ctx->stream);
    cudaMemcpyAsync(out_idx, d_idx, (size_t)nq * k * 4, cudaMemcpyDeviceToHost, ctx->stream);
    cudaMemcpyAsync(out_dist, d_dist, (size_t)nq * k * 8, cudaMemcpyDeviceToHost, ctx->stream);
    if (stats) cudaMemcpyAsync(stats, d_stats, 16, cudaMemcpyDeviceToHost, ctx->stream);
  }
  cudaError_t e = cudaStreamSynchronize(ctx->stream);
  cudaFree(buf);
  if (rc) return rc;
  if (e != cudaSuccess) return sepb_fail(ctx, (int)e, "netvlad_filter_debug", __FILE__, __LINE__);
  if (stats) stats[2] = nq - stats[1];
  return SEPB_OK;
}

int sepb_topk_merge_dev(sepb_ctx* ctx, const int32_t* d_in_idx, const double* d_in_dist, int lists, int nq, int k,
                        int32_t* d_out_idx, double* d_out_dist) {
  if (!ctx) return SEPB_E_INVALID;
  SEPB_REQUIRE(ctx, lists >= 1 && lists <= 64 && nq >= 0 && k >= 1 && k <= 16, SEPB_E_SHAPE);
  SEPB_REQUIRE(ctx, d_in_idx && d_in_dist && d_out_idx && d_out_dist, SEPB_E_INVALID);
  SEPB_CUDA(ctx, cudaSetDevice(ctx->device));
  return sepb_launch_topk_merge(ctx, d_in_idx, d_in_dist, lists, nq, k, d_out_idx, d_out_dist);
}

// data_handler.py:191-205
int sepb_select_matches(const int32_t* row_min_idx, const double* row_min_dist, int nq, double netvlad_distance,
                        int max_matches_nb, int32_t* out_pairs) {
  if (nq < 0 || max_matches_nb < 0 || (nq > 0 && (!row_min_idx || !row_min_dist)) || !out_pairs) return SEPB_E_INVALID;
  std::vector<int32_t> order(nq);
  std::iota(order.begin(), order.end(), 0);
  // np.argsort(smallest_values_each_frame): ties -> lowest row (stable), inf rows last
  std::stable_sort(order.begin(), order.end(), [&](int32_t a, int32_t b) { return row_min_dist[a] < row_min_dist[b]; });
  int n_out = 0;
  const int lim = std::min(nq, max_matches_nb);
  for (int i = 0; i < lim; ++i) {
    const int32_t idx_local = order[i], idx_other = row_min_idx[idx_local];
    bool used = false;
    for (int j = 0; j < n_out; ++j) used |= out_pairs[2 * j + 1] == idx_other;
    if (used) continue;                                   // still consumes one of the slots (:194)
    if (row_min_dist[idx_local] < netvlad_distance) {     // strict (:202)
      out_pairs[2 * n_out] = idx_local;
      out_pairs[2 * n_out + 1] = idx_other;
      ++n_out;
    } else {
      break;                                              // (:204-205)
    }
  }
  return n_out;
}

}  // extern "C"
