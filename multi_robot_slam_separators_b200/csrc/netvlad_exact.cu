// K1b: exact fp64 NetVLAD distances in scipy.cdist's summation order + masks + per-row top-k.
//
// Replaces `cdist` + row/column/cell masks + `argsort(axis=1)[:,0]` of DataHandler.find_matches
// (R/scripts/data_handler.py:170-188), extended to top-k per row (BASELINE.json configs[1]: k=5).
// cdist(euclidean) is s = 0; for j: d = u[j]-v[j]; s += d*d; return sqrt(s)  -- sequential, no FMA
// (verified bit-for-bit against scipy 1.18 for D = 77, 128, 4096), so each (query,row) distance is
// one serial chain; the parallelism is across pairs.  A thread owns one query and 32 database rows
// at a time (32 independent chains), database chunks are read as 128-bit shared-memory broadcasts,
// query chunks through an XOR-swizzled tile.  Two uses: (a) full scan of a database slice
// (exact path; also the fallback of the tensor-core filter), (b) re-ranking candidate lists.
// Output order per row: ascending (distance, index) -- ties to the lowest index like np.argmin.
#include "common.cuh"

namespace {

constexpr int QB = 128;   // queries per CTA == threads
constexpr int NB = 32;    // database rows per tile
constexpr int DC = 64;    // descriptor elements per staged chunk
constexpr int KMAX = 16;  // max k

struct ExactParams {
  const double* q;        // [nq][D]
  const double* db;       // [n][D]
  int nq, n, D;
  int k;
  int64_t base;           // global index of db row 0 (sharded databases)
  const uint8_t* row_mask;   // [nq] 1 = row ignored (local_kf_already_used)
  const uint8_t* col_mask;   // [n]  1 = column ignored (other_kf_already_used)
  const int32_t* ign_start;  // [nq+1] CSR over ignored cells (frames_kept_pairs_ignored), may be null
  const int32_t* ign_col;    // global column ids
  int splits;             // database split count (grid.y)
  int32_t* out_idx;       // [splits][nq][k]
  double* out_dist;       // [splits][nq][k]
  // gather mode (fallback rows of the tensor-core filter): scan only the rows row_list[0 .. *n_rows), results
  // are written at list position; both null = every row in order
  const int32_t* row_list;
  const int32_t* n_rows;
};

__device__ __forceinline__ bool better(double d, int64_t i, double d2, int64_t i2) {
  return d < d2 || (d == d2 && i < i2);
}

__global__ void __launch_bounds__(QB) netvlad_exact_kernel(ExactParams p) {
  extern __shared__ __align__(16) uint8_t smem[];
  double* q_s = reinterpret_cast<double*>(smem);                 // [QB][DC] swizzled in 16 B units
  double* db_s = q_s + QB * DC;                                  // [NB][DC]
  double* top_d = db_s + NB * DC;                                // [KMAX][QB]
  int32_t* top_i = reinterpret_cast<int32_t*>(top_d + KMAX * QB);  // [KMAX][QB]
  const int tid = threadIdx.x;
  const int q0 = blockIdx.x * QB;
  const int n_act = p.n_rows ? min(*p.n_rows, p.nq) : p.nq;   // rows to scan (device-side count in gather mode)
  if (q0 >= n_act) return;
  const int qpos = q0 + tid;                                    // position in the output / row list
  const int qi = qpos < n_act ? (p.row_list ? p.row_list[qpos] : qpos) : p.nq;   // query row (p.nq = none)
  const int k = p.k;
  // database slice of this split, in whole tiles
  const int tiles = (p.n + NB - 1) / NB;
  const int t_lo = (int)((int64_t)tiles * blockIdx.y / p.splits), t_hi = (int)((int64_t)tiles * (blockIdx.y + 1) / p.splits);
  for (int j = 0; j < k; ++j) {
    top_d[j * QB + tid] = INFINITY;
    top_i[j * QB + tid] = -1;
  }
  double worst = INFINITY;   // k-th best distance so far
  const bool row_dead = qi >= p.nq || (p.row_mask && p.row_mask[qi]);
  const int ig0 = (p.ign_start && qi < p.nq) ? p.ign_start[qi] : 0, ig1 = (p.ign_start && qi < p.nq) ? p.ign_start[qi + 1] : 0;

  for (int t = t_lo; t < t_hi; ++t) {
    const int n0 = t * NB;
    double acc[NB];
#pragma unroll
    for (int j = 0; j < NB; ++j) acc[j] = 0.0;
    for (int kc = 0; kc < p.D; kc += DC) {
      __syncthreads();
      // stage query chunk (swizzled) and database chunk; zero fill outside (0-0)^2 adds exactly 0
      for (int e = tid; e < QB * DC; e += QB) {
        const int r = e / DC, c = e % DC;
        double v = 0.0;
        if (q0 + r < n_act && kc + c < p.D) {
          const int qr = p.row_list ? p.row_list[q0 + r] : q0 + r;
          v = p.q[(size_t)qr * p.D + kc + c];
        }
        q_s[r * DC + ((((c >> 1) ^ (r & 31)) << 1) | (c & 1))] = v;
      }
      for (int e = tid; e < NB * DC; e += QB) {
        const int r = e / DC, c = e % DC;
        double v = 0.0;
        if (n0 + r < p.n && kc + c < p.D) v = p.db[(size_t)(n0 + r) * p.D + kc + c];
        db_s[e] = v;
      }
      __syncthreads();
      const double2* qrow = reinterpret_cast<const double2*>(q_s + tid * DC);
      const double2* dbt = reinterpret_cast<const double2*>(db_s);
#pragma unroll 2
      for (int c2 = 0; c2 < DC / 2; ++c2) {
        const double2 a = qrow[c2 ^ (tid & 31)];
#pragma unroll
        for (int j = 0; j < NB; ++j) {
          const double2 b = dbt[j * (DC / 2) + c2];
          const double d0 = __dsub_rn(a.x, b.x);
          acc[j] = __dadd_rn(acc[j], __dmul_rn(d0, d0));
          const double d1 = __dsub_rn(a.y, b.y);
          acc[j] = __dadd_rn(acc[j], __dmul_rn(d1, d1));
        }
      }
    }
    if (row_dead) continue;
#pragma unroll
    for (int j = 0; j < NB; ++j) {
      const int n = n0 + j;
      if (n >= p.n) continue;
      const double d = __dsqrt_rn(acc[j]);
      if (!(d <= worst)) continue;        // also rejects NaN
      if (p.col_mask && p.col_mask[n]) continue;
      const int64_t gi = p.base + n;
      bool ign = false;
      for (int g = ig0; g < ig1; ++g) ign |= p.ign_col[g] == (int32_t)gi;
      if (ign || isinf(d)) continue;
      // sorted insert by (d, idx); lists are short
      int pos = k - 1;
      if (!better(d, gi, top_d[pos * QB + tid], top_i[pos * QB + tid] < 0 ? INT64_MAX : (int64_t)top_i[pos * QB + tid])) continue;
      while (pos > 0) {
        const double pd = top_d[(pos - 1) * QB + tid];
        const int32_t pi = top_i[(pos - 1) * QB + tid];
        if (!better(d, gi, pd, pi < 0 ? INT64_MAX : (int64_t)pi)) break;
        top_d[pos * QB + tid] = pd;
        top_i[pos * QB + tid] = pi;
        --pos;
      }
      top_d[pos * QB + tid] = d;
      top_i[pos * QB + tid] = (int32_t)gi;
      worst = top_d[(k - 1) * QB + tid];
    }
  }
  if (qi < p.nq) {
    const size_t o = ((size_t)blockIdx.y * p.nq + qpos) * k;
    for (int j = 0; j < k; ++j) {
      p.out_idx[o + j] = top_i[j * QB + tid];
      p.out_dist[o + j] = top_d[j * QB + tid];
    }
  }
}

// merge `lists` sorted top-k lists per row (layout [lists][nq][k]) into one; thread per row
__global__ void topk_merge_kernel(const int32_t* __restrict__ in_idx, const double* __restrict__ in_dist, int lists,
                                  int nq, int k, int32_t* __restrict__ out_idx, double* __restrict__ out_dist,
                                  const int32_t* __restrict__ row_list, const int32_t* __restrict__ n_rows) {
  const int q = blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= (n_rows ? min(*n_rows, nq) : nq)) return;
  const int qo = row_list ? row_list[q] : q;   // gather mode: scatter to the original row
  int head[64];
  for (int l = 0; l < lists; ++l) head[l] = 0;
  for (int j = 0; j < k; ++j) {
    int bl = -1;
    double bd = INFINITY;
    int32_t bi = -1;
    for (int l = 0; l < lists; ++l) {
      if (head[l] >= k) continue;
      const size_t o = ((size_t)l * nq + q) * k + head[l];
      const int32_t i = in_idx[o];
      if (i < 0) continue;
      const double d = in_dist[o];
      if (bl < 0 || d < bd || (d == bd && i < bi)) {
        bl = l;
        bd = d;
        bi = i;
      }
    }
    out_idx[(size_t)qo * k + j] = bi;
    out_dist[(size_t)qo * k + j] = bl < 0 ? INFINITY : bd;
    if (bl >= 0) head[bl]++;
  }
}

}  // namespace

int sepb_launch_netvlad_exact(sepb_ctx* ctx, const double* d_q, int nq, const double* d_db, int n, int D, int k,
                              int64_t base, const uint8_t* d_row_mask, const uint8_t* d_col_mask,
                              const int32_t* d_ign_start, const int32_t* d_ign_col, int splits, int32_t* d_part_idx,
                              double* d_part_dist, int32_t* d_out_idx, double* d_out_dist, const int32_t* d_row_list,
                              const int32_t* d_n_rows) {
  if (nq == 0) return SEPB_OK;
  const bool gather = d_row_list != nullptr;
  if (gather && (!d_part_idx || !d_part_dist || !d_n_rows))
    return sepb_fail(ctx, SEPB_E_INVALID, "netvlad_exact: gather mode needs partial buffers", __FILE__, __LINE__);
  const bool two_pass = splits > 1 || gather;
  ExactParams p{d_q, d_db, nq, n, D, k, base, d_row_mask, d_col_mask, d_ign_start, d_ign_col, splits,
                two_pass ? d_part_idx : d_out_idx, two_pass ? d_part_dist : d_out_dist, d_row_list, d_n_rows};
  const size_t smem = (size_t)(QB * DC + NB * DC + KMAX * QB) * 8 + (size_t)KMAX * QB * 4;
  bool& configured = ctx->cfg_exact;
  if (!configured) {
    SEPB_CUDA(ctx, cudaFuncSetAttribute(netvlad_exact_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    configured = true;
  }
  dim3 grid((nq + QB - 1) / QB, splits);
  netvlad_exact_kernel<<<grid, QB, smem, ctx->stream>>>(p);
  ctx->launches++;
  SEPB_CUDA(ctx, cudaGetLastError());
  if (two_pass) {
    topk_merge_kernel<<<(nq + 127) / 128, 128, 0, ctx->stream>>>(d_part_idx, d_part_dist, splits, nq, k, d_out_idx,
                                                                 d_out_dist, d_row_list, d_n_rows);
    ctx->launches++;
    SEPB_CUDA(ctx, cudaGetLastError());
  }
  return SEPB_OK;
}

int sepb_launch_topk_merge(sepb_ctx* ctx, const int32_t* d_in_idx, const double* d_in_dist, int lists, int nq, int k,
                           int32_t* d_out_idx, double* d_out_dist) {
  if (nq == 0) return SEPB_OK;
  if (lists > 64) return sepb_fail(ctx, SEPB_E_SHAPE, "topk_merge: more than 64 lists", __FILE__, __LINE__);
  topk_merge_kernel<<<(nq + 127) / 128, 128, 0, ctx->stream>>>(d_in_idx, d_in_dist, lists, nq, k, d_out_idx, d_out_dist,
                                                               nullptr, nullptr);
  ctx->launches++;
  SEPB_CUDA(ctx, cudaGetLastError());
  return SEPB_OK;
}

// ---- allocation-free variant of the exact search ----------------------------------------------------------------------
// No counterpart in the reference; added for callers that cannot call cudaMalloc
// (CUDA graph capture, tight serving loops, harnesses that forbid it).  sepb_netvlad_topk_dev takes the per-split
// partial lists of the two-pass scan from the context's grow-only scratch (cudaMalloc on growth); here they live in a
// file-scope device array of fixed size that every GPU loading the library holds until exit.
//   * The array's name is only a host-side stand-in in host code: its device address is taken once per device, with that
//     device current, by cudaGetSymbolAddress, and that pointer is what the kernels receive.
//   * Every call shares the array, and it keeps the last call's contents.  Nothing in it has to start at zero: the merge
//     reads exactly the [splits][nq][k] slots the scan of the SAME call has written, so nothing is cleared.
//   * Calls that overlap on the GPU (two streams / contexts / host threads on one device) race on it.
constexpr size_t kStaticScratchBytes = size_t(8) << 20;   // 8 MiB per device: splits * nq * k * 12 bytes (+ alignment)
static __device__ __align__(256) unsigned char g_exact_static_scratch[kStaticScratchBytes];

namespace {
constexpr int kMaxDevices = 64;
std::mutex g_static_mu;
unsigned char* g_static_ptr[kMaxDevices] = {};   // device address of g_exact_static_scratch, per device
}  // namespace

extern "C" int sepb_netvlad_topk_exact_static_dev(sepb_ctx* ctx, const sepb_db* db, const double* d_q, int nq, int k,
                                                  const uint8_t* d_row_mask, const int32_t* d_ign_start,
                                                  const int32_t* d_ign_col, int32_t* d_out_idx, double* d_out_dist) {
  if (!ctx || !db) return SEPB_E_INVALID;
  SEPB_REQUIRE(ctx, nq >= 0 && k >= 1 && k <= KMAX, SEPB_E_SHAPE);
  SEPB_REQUIRE(ctx, ctx->device >= 0 && ctx->device < kMaxDevices, SEPB_E_STATE);
  SEPB_CUDA(ctx, cudaSetDevice(ctx->device));
  unsigned char* base_ptr = nullptr;
  {
    // set-up, once per device (a call with nq == 0 does only this, e.g. before a graph capture starts)
    std::lock_guard<std::mutex> lk(g_static_mu);
    if (!g_static_ptr[ctx->device]) {
      void* p = nullptr;
      SEPB_CUDA(ctx, cudaGetSymbolAddress(&p, g_exact_static_scratch));
      g_static_ptr[ctx->device] = static_cast<unsigned char*>(p);
    }
    base_ptr = g_static_ptr[ctx->device];
  }
  if (nq == 0) return SEPB_OK;
  SEPB_REQUIRE(ctx, d_q && d_out_idx && d_out_dist, SEPB_E_INVALID);
  // same database split as sepb_netvlad_topk_dev (api_netvlad.cu:exact_splits): identical launches, identical results
  const int row_blocks = (nq + QB - 1) / QB;
  int splits = (2 * ctx->sm_count + row_blocks - 1) / row_blocks;
  const int tiles = (db->size + NB - 1) / NB;
  splits = splits < 64 ? splits : 64;
  splits = splits < tiles ? splits : tiles;
  splits = splits > 1 ? splits : 1;
  int32_t* part_i = nullptr;
  double* part_d = nullptr;
  if (splits > 1) {
    const size_t b_d = align_up((size_t)splits * nq * k * 8, 256), b_i = align_up((size_t)splits * nq * k * 4, 256);
    if (b_d + b_i > kStaticScratchBytes) {
      char msg[256];
      snprintf(msg, sizeof msg,
               "sepb_netvlad_topk_exact_static_dev: %d splits x %d queries x k = %d need %zu bytes of partial lists, "
               "the static scratch holds %zu (use sepb_netvlad_topk_dev or fewer queries per call)",
               splits, nq, k, b_d + b_i, kStaticScratchBytes);
      return sepb_fail(ctx, SEPB_E_SHAPE, msg, __FILE__, __LINE__);
    }
    part_d = reinterpret_cast<double*>(base_ptr);
    part_i = reinterpret_cast<int32_t*>(base_ptr + b_d);
  }
  return sepb_launch_netvlad_exact(ctx, d_q, nq, db->rows, db->size, db->dim, k, db->base, d_row_mask,
                                   db->has_mask ? db->col_mask : nullptr, d_ign_start, d_ign_col, splits, part_i, part_d,
                                   d_out_idx, d_out_dist, nullptr, nullptr);
}
