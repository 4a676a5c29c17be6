// C ABI: K3 (PnP-RANSAC) and the fused K2 -> K3 verification entry points.
// Compiled with -fmad=false like pnp_ransac.cu? No: only host-side parameter preparation here; the
// Rodrigues of the guess is computed with the same routine the kernel uses (cvm::), on the host.
#include "common.cuh"
#include "pnp_geom.cuh"

#include <vector>

namespace {

int ensure_table(sepb_ctx* ctx, int max_count, int H) {
  if (ctx->table && ctx->table_m >= max_count && ctx->table_h >= H) return SEPB_OK;
  const int m = max_count > ctx->table_m ? max_count : ctx->table_m;
  const int h = H > ctx->table_h ? H : ctx->table_h;
  if (ctx->table) {
    SEPB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    cudaFree(ctx->table);
    ctx->table = nullptr;
  }
  SEPB_CUDA(ctx, cudaMalloc(&ctx->table, (size_t)(m + 1) * h * 5 * sizeof(uint16_t)));
  ctx->table_m = m;
  ctx->table_h = h;
  return sepb_launch_sample_table(ctx, ctx->table, m, h);
}

int check_pnp_params(sepb_ctx* ctx, const sepb_pnp_params* p) {
  SEPB_REQUIRE(ctx, p->min_inliers >= 5, SEPB_E_INVALID);  // 4-point P3P path is not implemented
  SEPB_REQUIRE(ctx, p->iterations >= 1 && p->iterations <= 65536, SEPB_E_INVALID);
  SEPB_REQUIRE(ctx, p->confidence > 0 && p->confidence < 1 && p->reproj_error > 0, SEPB_E_INVALID);
  SEPB_REQUIRE(ctx, p->fx > 0 && p->fy > 0, SEPB_E_INVALID);
  return SEPB_OK;
}

}  // namespace

extern "C" {

void sepb_pnp_params_init(sepb_pnp_params* p, double fx, double fy, double cx, double cy) {
  if (!p) return;
  p->fx = fx; p->fy = fy; p->cx = cx; p->cy = cy;
  p->iterations = 300;
  p->min_inliers = 5;
  p->reproj_error = 2.0;
  p->confidence = 0.99;
  // optical (x right, y down, z forward) -> base (x forward, y left, z up)
  const double L[9] = {0, 0, 1, -1, 0, 0, 0, -1, 0};
  for (int i = 0; i < 9; ++i) p->local_R[i] = L[i];
  p->local_t[0] = p->local_t[1] = p->local_t[2] = 0;
  const double I[12] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0};
  sepb_pnp_set_guess(p, I);
}

void sepb_pnp_set_guess(sepb_pnp_params* p, const double* G) {
  if (!p || !G) return;
  // rtabmap::Transform is f32: guessCameraFrame = (guess * local)^-1 evaluated on f32-rounded entries
  double RP[9], tP[3];
  for (int i = 0; i < 3; ++i) {
    for (int j = 0; j < 3; ++j)
      RP[3 * i + j] = (double)(float)(G[4 * i] * p->local_R[j] + G[4 * i + 1] * p->local_R[3 + j] + G[4 * i + 2] * p->local_R[6 + j]);
    tP[i] = (double)(float)(G[4 * i] * p->local_t[0] + G[4 * i + 1] * p->local_t[1] + G[4 * i + 2] * p->local_t[2] + G[4 * i + 3]);
  }
  double Rinv[9];
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) Rinv[3 * i + j] = RP[3 * j + i];
  for (int i = 0; i < 3; ++i)
    p->guess_tvec[i] = (double)(float)(-(Rinv[3 * i] * tP[0] + Rinv[3 * i + 1] * tP[1] + Rinv[3 * i + 2] * tP[2]));
  double scratch[21 * cvm::kHeapStride];   // host call of the strided-heap routine
  cvm::rodrigues_mat_to_vec(cvm::Heap{scratch}, Rinv, p->guess_rvec);
}

int sepb_pnp_ransac_dev(sepb_ctx* ctx, const sepb_frames* from, const sepb_frames* to, const int32_t* d_pairs,
                        int n_pairs, const int32_t* d_n_matches, const uint16_t* d_match_from,
                        const uint16_t* d_match_to, const int32_t* d_n_unique_to, const sepb_pnp_params* prm,
                        const int32_t* d_custom_samples, int custom_h, const double* d_guess, int32_t* d_status,
                        int32_t* d_n_inliers, double* d_pose, float* d_transform, uint32_t* d_inlier_mask,
                        int mask_words, int32_t* d_best_hyp, int32_t* d_iters_run) {
  if (!ctx || !from || !to || !prm) return SEPB_E_INVALID;
  SEPB_REQUIRE(ctx, n_pairs >= 0, SEPB_E_SHAPE);
  if (n_pairs == 0) return SEPB_OK;
  SEPB_REQUIRE(ctx, d_pairs && d_n_matches && d_match_from && d_match_to && d_status && d_n_inliers && d_pose &&
                        d_transform, SEPB_E_INVALID);
  SEPB_REQUIRE(ctx, !d_inlier_mask || mask_words * 32 >= from->max_feat, SEPB_E_SHAPE);
  SEPB_REQUIRE(ctx, !d_custom_samples || custom_h >= prm->iterations, SEPB_E_SHAPE);
  int rc = check_pnp_params(ctx, prm);
  if (rc) return rc;
  SEPB_CUDA(ctx, cudaSetDevice(ctx->device));
  if (!d_custom_samples) {
    rc = ensure_table(ctx, from->max_feat, prm->iterations);
    if (rc) return rc;
  }
  sepb_frames_acquire(ctx, from, to);
  rc = sepb_launch_pnp_ransac(ctx, view_of(from), view_of(to), d_pairs, n_pairs, d_n_matches, d_match_from,
                              d_match_to, d_n_unique_to, ctx->table, ctx->table_h, d_custom_samples, custom_h, d_guess,
                              prm, d_status, d_n_inliers, d_pose, d_transform, d_inlier_mask, mask_words, d_best_hyp,
                              d_iters_run);
  sepb_frames_release(ctx, from, to);
  return rc;
}

int sepb_verify_pairs_dev(sepb_ctx* ctx, const sepb_frames* from, const sepb_frames* to, const int32_t* d_pairs,
                          int n_pairs, const sepb_match_params* mprm, const sepb_pnp_params* pprm,
                          const int32_t* d_custom_samples, int custom_h, const sepb_verify_out* o) {
  if (!ctx || !from || !to || !mprm || !pprm || !o) return SEPB_E_INVALID;
  SEPB_REQUIRE(ctx, n_pairs >= 0, SEPB_E_SHAPE);
  if (n_pairs == 0) return SEPB_OK;
  SEPB_REQUIRE(ctx, o->status && o->n_matches && o->n_inliers && o->pose && o->transform && o->match_from &&
                        o->match_to, SEPB_E_INVALID);
  SEPB_CUDA(ctx, cudaSetDevice(ctx->device));
  // n_unique_to lives in a small grow-only scratch slice owned by the context
  int32_t* d_nu = (int32_t*)sepb_scratch(ctx, align_up((size_t)n_pairs * 4, 256));
  if (!d_nu) return (int)cudaErrorMemoryAllocation;
  sepb_frames_acquire(ctx, from, to);
  int rc = sepb_launch_match(ctx, view_of(from), view_of(to), d_pairs, n_pairs, mprm->nndr, mprm->nndr_strict,
                                     nullptr, nullptr, o->n_matches, o->match_from, o->match_to, d_nu);
  if (rc) return rc;
  return sepb_pnp_ransac_dev(ctx, from, to, d_pairs, n_pairs, o->n_matches, o->match_from, o->match_to, d_nu, pprm,
                             d_custom_samples, custom_h, nullptr, o->status, o->n_inliers, o->pose, o->transform,
                             o->inlier_mask, (from->max_feat + 31) / 32, o->best_hyp, o->iters_run);
}

static int verify_pairs_impl(sepb_ctx* ctx, const sepb_frames* from, const sepb_frames* to, const int32_t* pairs,
                             int n_pairs, const sepb_match_params* mprm, const sepb_pnp_params* pprm,
                             const int32_t* custom_samples, int custom_h, sepb_verify_out* out, bool wait) {
  if (!ctx || !from || !to || !mprm || !pprm || !out) return SEPB_E_INVALID;
  SEPB_REQUIRE(ctx, n_pairs >= 0, SEPB_E_SHAPE);
  if (n_pairs == 0) return SEPB_OK;
  SEPB_REQUIRE(ctx, pairs && out->status, SEPB_E_INVALID);
  for (int i = 0; i < n_pairs; ++i)
    SEPB_REQUIRE(ctx, pairs[2 * i] >= 0 && pairs[2 * i] < from->size && pairs[2 * i + 1] >= 0 &&
                          pairs[2 * i + 1] < to->size, SEPB_E_SHAPE);
  int rc = check_pnp_params(ctx, pprm);
  if (rc) return rc;
  std::lock_guard<std::mutex> lk(ctx->mu);
  SEPB_CUDA(ctx, cudaSetDevice(ctx->device));
  const size_t P = (size_t)n_pairs, F = (size_t)from->max_feat, MW = (F + 31) / 32;
  const size_t b_nu = align_up(P * 4, 256);  // first slice: n_unique_to (used by verify_pairs_dev)
  const size_t b_pairs = align_up(P * 8, 256), b_i = align_up(P * 4, 256), b_pose = align_up(P * 48, 256),
               b_T = align_up(P * 48, 256), b_m = align_up(P * F * 2, 256), b_mask = align_up(P * MW * 4, 256),
               b_cs = custom_samples ? align_up(P * custom_h * 5 * 4, 256) : 0;
  const size_t total = b_nu + b_pairs + 5 * b_i + b_pose + b_T + 2 * b_m + b_mask + b_cs;
  uint8_t* s = (uint8_t*)sepb_scratch(ctx, total);
  if (!s) return (int)cudaErrorMemoryAllocation;
  s += b_nu;
  int32_t* d_pairs = (int32_t*)s; s += b_pairs;
  sepb_verify_out d{};
  d.status = (int32_t*)s; s += b_i;
  d.n_matches = (int32_t*)s; s += b_i;
  d.n_inliers = (int32_t*)s; s += b_i;
  d.best_hyp = (int32_t*)s; s += b_i;
  d.iters_run = (int32_t*)s; s += b_i;
  d.pose = (double*)s; s += b_pose;
  d.transform = (float*)s; s += b_T;
  d.match_from = (uint16_t*)s; s += b_m;
  d.match_to = (uint16_t*)s; s += b_m;
  d.inlier_mask = (uint32_t*)s; s += b_mask;
  int32_t* d_cs = nullptr;
  if (custom_samples) {
    d_cs = (int32_t*)s;
    SEPB_CUDA(ctx, cudaMemcpyAsync(d_cs, custom_samples, P * custom_h * 5 * 4, cudaMemcpyHostToDevice, ctx->stream));
  }
  SEPB_CUDA(ctx, cudaMemcpyAsync(d_pairs, pairs, P * 8, cudaMemcpyHostToDevice, ctx->stream));
  rc = sepb_verify_pairs_dev(ctx, from, to, d_pairs, n_pairs, mprm, pprm, d_cs, custom_h, &d);
  if (rc) return rc;
#define D2H(field, bytes) \
  if (out->field) SEPB_CUDA(ctx, cudaMemcpyAsync(out->field, d.field, (bytes), cudaMemcpyDeviceToHost, ctx->stream))
  D2H(status, P * 4);
  D2H(n_matches, P * 4);
  D2H(n_inliers, P * 4);
  D2H(best_hyp, P * 4);
  D2H(iters_run, P * 4);
  D2H(pose, P * 48);
  D2H(transform, P * 48);
  D2H(match_from, P * F * 2);
  D2H(match_to, P * F * 2);
  D2H(inlier_mask, P * MW * 4);
#undef D2H
  if (wait) SEPB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return SEPB_OK;
}

int sepb_verify_pairs(sepb_ctx* ctx, const sepb_frames* from, const sepb_frames* to, const int32_t* pairs, int n_pairs,
                      const sepb_match_params* mprm, const sepb_pnp_params* pprm, const int32_t* custom_samples,
                      int custom_h, sepb_verify_out* out) {
  return verify_pairs_impl(ctx, from, to, pairs, n_pairs, mprm, pprm, custom_samples, custom_h, out, true);
}

int sepb_verify_pairs_async(sepb_ctx* ctx, const sepb_frames* from, const sepb_frames* to, const int32_t* pairs,
                            int n_pairs, const sepb_match_params* mprm, const sepb_pnp_params* pprm,
                            const int32_t* custom_samples, int custom_h, sepb_verify_out* out) {
  return verify_pairs_impl(ctx, from, to, pairs, n_pairs, mprm, pprm, custom_samples, custom_h, out, false);
}

}  // extern "C"
