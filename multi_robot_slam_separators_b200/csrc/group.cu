// C ABI: multi-GPU group -- an NCCL communicator clique over the devices of one box and the row-sharded NetVLAD
// search on top of it (the peer database is partitioned by rows, queries are replicated, every device
// searches its shard with GLOBAL row ids, ONE all-gather of the per-shard [Q, k] {idx, f64 dist} lists travels over
// NVLink / NVSwitch and sepb_topk_merge's kernel merges them in the (distance, index) order of the single-device search:
// the result is identical to one big database, reference semantics R/scripts/data_handler.py:166-191).
//
// NCCL is resolved at run time (dlopen("libnccl.so.2")): the library loads and every single-device entry point works
// on a machine without NCCL; the group entry points then fail with SEPB_E_STATE.  In a process that already carries
// an NCCL (e.g. torch's bundled one) the same copy is used.
#include <dlfcn.h>
#include <nccl.h>

#include <new>
#include <string>
#include <vector>

#include "common.cuh"

int sepb_launch_topk_merge(sepb_ctx* ctx, const int32_t* d_in_idx, const double* d_in_dist, int lists, int nq, int k,
                           int32_t* d_out_idx, double* d_out_dist);

namespace {

struct NcclApi {
  void* handle = nullptr;
  ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommInitAll)(ncclComm_t*, int, const int*) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*Broadcast)(const void*, void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*GroupStart)() = nullptr;
  ncclResult_t (*GroupEnd)() = nullptr;
  const char* (*GetErrorString)(ncclResult_t) = nullptr;
  ncclResult_t (*GetVersion)(int*) = nullptr;
  bool ok = false;
};

NcclApi& nccl() {
  static NcclApi api = [] {
    NcclApi a;
    for (const char* name : {"libnccl.so.2", "libnccl.so"}) {
      a.handle = dlopen(name, RTLD_NOW | RTLD_GLOBAL);
      if (a.handle) break;
    }
    if (!a.handle) return a;
#define SEPB_NCCL_SYM(field, sym) a.field = reinterpret_cast<decltype(a.field)>(dlsym(a.handle, sym))
    SEPB_NCCL_SYM(GetUniqueId, "ncclGetUniqueId");
    SEPB_NCCL_SYM(CommInitRank, "ncclCommInitRank");
    SEPB_NCCL_SYM(CommInitAll, "ncclCommInitAll");
    SEPB_NCCL_SYM(CommDestroy, "ncclCommDestroy");
    SEPB_NCCL_SYM(AllGather, "ncclAllGather");
    SEPB_NCCL_SYM(Broadcast, "ncclBroadcast");
    SEPB_NCCL_SYM(GroupStart, "ncclGroupStart");
    SEPB_NCCL_SYM(GroupEnd, "ncclGroupEnd");
    SEPB_NCCL_SYM(GetErrorString, "ncclGetErrorString");
    SEPB_NCCL_SYM(GetVersion, "ncclGetVersion");
#undef SEPB_NCCL_SYM
    a.ok = a.GetUniqueId && a.CommInitRank && a.CommInitAll && a.CommDestroy && a.AllGather && a.Broadcast &&
           a.GroupStart && a.GroupEnd && a.GetErrorString;
    return a;
  }();
  return api;
}

}  // namespace

struct sepb_group {
  int nlocal = 0;            // devices driven by this process
  int nranks = 0;            // devices in the communicator
  int rank0 = 0;             // communicator rank of local device 0 (local device i has rank0 + i)
  bool owns_ctx = false;
  std::vector<sepb_ctx*> ctx;
  std::vector<ncclComm_t> comm;
  // per local device, grow-only: this shard's lists, the gathered lists, staged queries / results of the host form
  struct Buf {
    int32_t* loc_idx = nullptr;
    double* loc_dist = nullptr;
    int32_t* all_idx = nullptr;
    double* all_dist = nullptr;
    size_t cap = 0;            // in (query, k) cells
    double* q = nullptr;
    size_t q_cap = 0;          // doubles
    int32_t* out_idx = nullptr;
    double* out_dist = nullptr;
    size_t out_cap = 0;        // cells
  };
  std::vector<Buf> buf;
  int64_t gather_bytes = 0;  // bytes this process received in the last all-gather (all local devices)
  std::string err;
};

namespace {

int group_fail(sepb_group* g, int code, const std::string& what) {
  if (g) g->err = what;
  return code;
}

#define SEPB_G_CUDA(g, expr)                                                                        \
  do {                                                                                              \
    cudaError_t _e = (expr);                                                                        \
    if (_e != cudaSuccess) return group_fail((g), (int)_e, std::string(#expr ": ") + cudaGetErrorString(_e)); \
  } while (0)
#define SEPB_G_NCCL(g, expr)                                                                        \
  do {                                                                                              \
    ncclResult_t _r = (expr);                                                                       \
    if (_r != ncclSuccess)                                                                          \
      return group_fail((g), SEPB_E_NCCL, std::string(#expr ": ") + nccl().GetErrorString(_r));     \
  } while (0)

int ensure(sepb_group* g, int i, size_t cells, size_t q_doubles, bool host_out) {
  sepb_group::Buf& b = g->buf[i];
  SEPB_G_CUDA(g, cudaSetDevice(g->ctx[i]->device));
  if (cells > b.cap) {
    SEPB_G_CUDA(g, cudaStreamSynchronize(g->ctx[i]->stream));
    cudaFree(b.loc_idx); cudaFree(b.loc_dist); cudaFree(b.all_idx); cudaFree(b.all_dist);
    b.loc_idx = nullptr; b.loc_dist = nullptr; b.all_idx = nullptr; b.all_dist = nullptr;
    b.cap = 0;
    const size_t cap = cells + cells / 4;
    SEPB_G_CUDA(g, cudaMalloc(&b.loc_idx, cap * 4));
    SEPB_G_CUDA(g, cudaMalloc(&b.loc_dist, cap * 8));
    SEPB_G_CUDA(g, cudaMalloc(&b.all_idx, cap * 4 * g->nranks));
    SEPB_G_CUDA(g, cudaMalloc(&b.all_dist, cap * 8 * g->nranks));
    b.cap = cap;
  }
  if (q_doubles > b.q_cap) {
    SEPB_G_CUDA(g, cudaStreamSynchronize(g->ctx[i]->stream));
    cudaFree(b.q);
    b.q = nullptr;
    b.q_cap = 0;
    SEPB_G_CUDA(g, cudaMalloc(&b.q, (q_doubles + q_doubles / 4) * 8));
    b.q_cap = q_doubles + q_doubles / 4;
  }
  if (host_out && cells > b.out_cap) {
    SEPB_G_CUDA(g, cudaStreamSynchronize(g->ctx[i]->stream));
    cudaFree(b.out_idx); cudaFree(b.out_dist);
    b.out_idx = nullptr; b.out_dist = nullptr;
    b.out_cap = 0;
    SEPB_G_CUDA(g, cudaMalloc(&b.out_idx, (cells + cells / 4) * 4));
    SEPB_G_CUDA(g, cudaMalloc(&b.out_dist, (cells + cells / 4) * 8));
    b.out_cap = cells + cells / 4;
  }
  return SEPB_OK;
}

}  // namespace

extern "C" {

int sepb_group_create(int ndev, const int* devices, sepb_group** out) {
  if (!out) return SEPB_E_INVALID;
  *out = nullptr;
  if (ndev < 1 || ndev > 64) return SEPB_E_SHAPE;
  if (!nccl().ok && ndev > 1) return SEPB_E_STATE;   // one device needs no communicator
  sepb_group* g = new (std::nothrow) sepb_group();
  if (!g) return (int)cudaErrorMemoryAllocation;
  g->nlocal = g->nranks = ndev;
  g->owns_ctx = true;
  g->ctx.assign(ndev, nullptr);
  g->comm.assign(ndev, nullptr);
  g->buf.resize(ndev);
  std::vector<int> devs(ndev);
  for (int i = 0; i < ndev; ++i) devs[i] = devices ? devices[i] : i;
  for (int i = 0; i < ndev; ++i) {
    const int rc = sepb_ctx_create(devs[i], &g->ctx[i]);
    if (rc != SEPB_OK) {
      sepb_group_destroy(g);
      return rc;
    }
  }
  if (ndev > 1) {
    const ncclResult_t r = nccl().CommInitAll(g->comm.data(), ndev, devs.data());   // one communicator clique, this process
    if (r != ncclSuccess) {
      sepb_group_destroy(g);
      return SEPB_E_NCCL;
    }
  }
  *out = g;
  return SEPB_OK;
}

int sepb_group_unique_id(uint8_t* id128) {
  if (!id128) return SEPB_E_INVALID;
  if (!nccl().ok) return SEPB_E_STATE;
  ncclUniqueId id;
  static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
  if (nccl().GetUniqueId(&id) != ncclSuccess) return SEPB_E_NCCL;
  memcpy(id128, &id, 128);
  return SEPB_OK;
}

int sepb_group_create_rank(sepb_ctx* ctx, const uint8_t* id128, int rank, int nranks, sepb_group** out) {
  if (!out) return SEPB_E_INVALID;
  *out = nullptr;
  if (!ctx || !id128 || nranks < 1 || nranks > 64 || rank < 0 || rank >= nranks) return SEPB_E_INVALID;
  if (!nccl().ok && nranks > 1) return SEPB_E_STATE;
  sepb_group* g = new (std::nothrow) sepb_group();
  if (!g) return (int)cudaErrorMemoryAllocation;
  g->nlocal = 1;
  g->nranks = nranks;
  g->rank0 = rank;
  g->ctx.assign(1, ctx);
  g->comm.assign(1, nullptr);
  g->buf.resize(1);
  if (nranks > 1) {
    ncclUniqueId id;
    memcpy(&id, id128, 128);
    cudaSetDevice(ctx->device);
    if (nccl().CommInitRank(&g->comm[0], nranks, id, rank) != ncclSuccess) {
      delete g;
      return SEPB_E_NCCL;
    }
  }
  *out = g;
  return SEPB_OK;
}

void sepb_group_destroy(sepb_group* g) {
  if (!g) return;
  for (int i = 0; i < g->nlocal; ++i) {
    if (!g->ctx[i]) continue;
    cudaSetDevice(g->ctx[i]->device);
    cudaStreamSynchronize(g->ctx[i]->stream);
    sepb_group::Buf& b = g->buf[i];
    cudaFree(b.loc_idx); cudaFree(b.loc_dist); cudaFree(b.all_idx); cudaFree(b.all_dist);
    cudaFree(b.q); cudaFree(b.out_idx); cudaFree(b.out_dist);
    if (g->comm[i]) nccl().CommDestroy(g->comm[i]);
  }
  if (g->owns_ctx)
    for (sepb_ctx* c : g->ctx) sepb_ctx_destroy(c);
  delete g;
}

int sepb_group_local_size(const sepb_group* g) { return g ? g->nlocal : 0; }
int sepb_group_nranks(const sepb_group* g) { return g ? g->nranks : 0; }
sepb_ctx* sepb_group_ctx(sepb_group* g, int i) { return g && i >= 0 && i < g->nlocal ? g->ctx[i] : nullptr; }
const char* sepb_group_last_error(const sepb_group* g) { return g ? g->err.c_str() : "null group"; }
int64_t sepb_group_last_gather_bytes(const sepb_group* g) { return g ? g->gather_bytes : 0; }

int sepb_group_nccl_version(void) {
  int v = 0;
  if (!nccl().ok || !nccl().GetVersion || nccl().GetVersion(&v) != ncclSuccess) return 0;
  return v;
}

int sepb_group_netvlad_topk_dev(sepb_group* g, sepb_db* const* dbs, const double* const* d_queries, int nq, int k, int mode,
                                int32_t* const* d_out_idx, double* const* d_out_dist) {
  if (!g || !dbs || !d_queries || !d_out_idx || !d_out_dist) return SEPB_E_INVALID;
  if (nq < 0 || k < 1 || k > 16) return group_fail(g, SEPB_E_SHAPE, "group_netvlad_topk: bad nq / k");
  if (nq == 0) return SEPB_OK;
  const size_t cells = (size_t)nq * k;
  for (int i = 0; i < g->nlocal; ++i) {
    if (!dbs[i] || dbs[i]->ctx != g->ctx[i] || !d_queries[i] || !d_out_idx[i] || !d_out_dist[i])
      return group_fail(g, SEPB_E_INVALID, "group_netvlad_topk: shard i must live on the group's context i");
    int rc = ensure(g, i, cells, 0, false);
    if (rc) return rc;
    // this shard's exact top-k with global row ids (K1 / K1b), on context i's stream
    rc = sepb_netvlad_topk_dev(g->ctx[i], dbs[i], d_queries[i], nq, k, mode, nullptr, nullptr, nullptr,
                               g->nranks > 1 ? g->buf[i].loc_idx : d_out_idx[i], g->nranks > 1 ? g->buf[i].loc_dist : d_out_dist[i]);
    if (rc) return group_fail(g, rc, sepb_last_error(g->ctx[i]));
  }
  g->gather_bytes = 0;
  if (g->nranks == 1) return SEPB_OK;
  // the path's only exchange: 12 bytes per (query, candidate) and shard, one fused collective per device
  SEPB_G_NCCL(g, nccl().GroupStart());
  for (int i = 0; i < g->nlocal; ++i) {
    sepb_group::Buf& b = g->buf[i];
    ncclResult_t r = nccl().AllGather(b.loc_idx, b.all_idx, cells, ncclInt32, g->comm[i], g->ctx[i]->stream);
    if (r == ncclSuccess) r = nccl().AllGather(b.loc_dist, b.all_dist, cells, ncclFloat64, g->comm[i], g->ctx[i]->stream);
    if (r != ncclSuccess) {
      nccl().GroupEnd();
      return group_fail(g, SEPB_E_NCCL, std::string("ncclAllGather: ") + nccl().GetErrorString(r));
    }
    g->gather_bytes += (int64_t)cells * 12 * (g->nranks - 1);
  }
  SEPB_G_NCCL(g, nccl().GroupEnd());
  for (int i = 0; i < g->nlocal; ++i) {
    SEPB_G_CUDA(g, cudaSetDevice(g->ctx[i]->device));
    const int rc = sepb_launch_topk_merge(g->ctx[i], g->buf[i].all_idx, g->buf[i].all_dist, g->nranks, nq, k, d_out_idx[i],
                                          d_out_dist[i]);
    if (rc) return group_fail(g, rc, sepb_last_error(g->ctx[i]));
  }
  return SEPB_OK;
}

int sepb_group_netvlad_topk(sepb_group* g, sepb_db* const* dbs, const double* queries, int nq, int k, int mode,
                            int32_t* out_idx, double* out_dist) {
  if (!g || !dbs || !queries || !out_idx || !out_dist) return SEPB_E_INVALID;
  if (nq < 0 || k < 1 || k > 16) return group_fail(g, SEPB_E_SHAPE, "group_netvlad_topk: bad nq / k");
  if (nq == 0) return SEPB_OK;
  for (int i = 0; i < g->nlocal; ++i)
    if (!dbs[i] || dbs[i]->ctx != g->ctx[i] || dbs[i]->dim != dbs[0]->dim)
      return group_fail(g, SEPB_E_INVALID, "group_netvlad_topk: shard i must live on the group's context i, same dim");
  const size_t cells = (size_t)nq * k, qd = (size_t)nq * dbs[0]->dim;
  std::vector<const double*> dq(g->nlocal);
  std::vector<int32_t*> oi(g->nlocal);
  std::vector<double*> od(g->nlocal);
  for (int i = 0; i < g->nlocal; ++i) {
    const int rc = ensure(g, i, cells, qd, true);
    if (rc) return rc;
    dq[i] = g->buf[i].q;
    oi[i] = g->buf[i].out_idx;
    od[i] = g->buf[i].out_dist;
  }
  // queries: one H2D copy, then NVLink to the other local devices (every process of a multi-process job passes its own copy)
  SEPB_G_CUDA(g, cudaSetDevice(g->ctx[0]->device));
  SEPB_G_CUDA(g, cudaMemcpyAsync(g->buf[0].q, queries, qd * 8, cudaMemcpyHostToDevice, g->ctx[0]->stream));
  if (g->nlocal > 1) {
    SEPB_G_NCCL(g, nccl().GroupStart());
    for (int i = 0; i < g->nlocal; ++i) {
      const ncclResult_t r = nccl().Broadcast(g->buf[i].q, g->buf[i].q, qd, ncclFloat64, 0, g->comm[i], g->ctx[i]->stream);
      if (r != ncclSuccess) {
        nccl().GroupEnd();
        return group_fail(g, SEPB_E_NCCL, std::string("ncclBroadcast: ") + nccl().GetErrorString(r));
      }
    }
    SEPB_G_NCCL(g, nccl().GroupEnd());
  }
  const int rc = sepb_group_netvlad_topk_dev(g, dbs, dq.data(), nq, k, mode, oi.data(), od.data());
  if (rc) return rc;
  SEPB_G_CUDA(g, cudaSetDevice(g->ctx[0]->device));
  SEPB_G_CUDA(g, cudaMemcpyAsync(out_idx, g->buf[0].out_idx, cells * 4, cudaMemcpyDeviceToHost, g->ctx[0]->stream));
  SEPB_G_CUDA(g, cudaMemcpyAsync(out_dist, g->buf[0].out_dist, cells * 8, cudaMemcpyDeviceToHost, g->ctx[0]->stream));
  for (int i = 0; i < g->nlocal; ++i) {
    SEPB_G_CUDA(g, cudaSetDevice(g->ctx[i]->device));
    SEPB_G_CUDA(g, cudaStreamSynchronize(g->ctx[i]->stream));
  }
  return SEPB_OK;
}

}  // extern "C"
