// C ABI: K3 (PnP-RANSAC) and the fused K2 -> K3 verification entry points.
// Compiled with -fmad=false like pnp_ransac.cu? No: only host-side parameter preparation here; the
// Rodrigues of the guess is computed with the same routine the kernel uses (cvm::), on the host.
#include "common.cuh"
#include "pnp_geom.cuh"

#include <vector>

namespace {

int ensure_table(sepb_ctx* ctx, int max_count, int H, int mp) {
  uint16_t*& tab = mp == 4 ? ctx->table4 : ctx->table;
  int& tm = mp == 4 ? ctx->table4_m : ctx->table_m;
  int& th = mp == 4 ? ctx->table4_h : ctx->table_h;
  if (tab && tm >= max_count && th >= H) return SEPB_OK;
  const int m = max_count > tm ? max_count : tm;
  const int h = H > th ? H : th;
  if (tab) {
    SEPB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    cudaFree(tab);
    tab = nullptr;
  }
  SEPB_CUDA(ctx, cudaMalloc(&tab, (size_t)(m + 1) * h * mp * sizeof(uint16_t)));
  tm = m;
  th = h;
  return sepb_launch_sample_table(ctx, tab, m, h, mp);
}

inline int model_points_of(const sepb_pnp_params* p) { return p->flags == SEPB_PNP_ITERATIVE ? 5 : 4; }

int check_pnp_params(sepb_ctx* ctx, const sepb_pnp_params* p) {
  SEPB_REQUIRE(ctx, p->min_inliers >= 4, SEPB_E_INVALID);  // util3d::solvePnPRansac: max(minInliers, 4)
  SEPB_REQUIRE(ctx, p->iterations >= 1 && p->iterations <= 65536, SEPB_E_INVALID);
  SEPB_REQUIRE(ctx, p->confidence > 0 && p->confidence < 1 && p->reproj_error > 0, SEPB_E_INVALID);
  SEPB_REQUIRE(ctx, p->fx > 0 && p->fy > 0, SEPB_E_INVALID);
  SEPB_REQUIRE(ctx, p->flags == SEPB_PNP_ITERATIVE || p->flags == SEPB_PNP_P3P || p->flags == SEPB_PNP_AP3P, SEPB_E_INVALID);
  SEPB_REQUIRE(ctx, p->refine_iterations >= 0 && p->refine_iterations <= 64, SEPB_E_INVALID);
  // RTAB-Map's refine loop calls solvePnP(inliers, flags): P3P / AP3P cannot take more than four points there
  SEPB_REQUIRE(ctx, p->refine_iterations == 0 || (p->flags == SEPB_PNP_ITERATIVE && p->refine_sigma > 0), SEPB_E_INVALID);
  SEPB_REQUIRE(ctx, p->minimal_model == 0 || p->minimal_model == 1, SEPB_E_INVALID);
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
  p->flags = SEPB_PNP_ITERATIVE;
  p->refine_iterations = 0;
  p->refine_sigma = 3.0;
  p->minimal_model = 0;
  p->reserved_ = 0;
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

static int pnp_ransac_impl(sepb_ctx* ctx, const sepb_frames* from, const sepb_frames* to, const int32_t* d_pairs,
                           int n_pairs, const int32_t* d_n_matches, const uint16_t* d_match_from,
                           const uint16_t* d_match_to, const int32_t* d_n_unique_to, const sepb_pnp_params* prm,
                           const int32_t* d_custom_samples, int custom_h, const double* d_guess,
                           const int32_t* d_only_if_ok, int32_t* d_status, int32_t* d_n_inliers, double* d_pose,
                           float* d_transform, uint32_t* d_inlier_mask, int mask_words, int32_t* d_best_hyp,
                           int32_t* d_iters_run) {
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
    rc = ensure_table(ctx, from->max_feat, prm->iterations, model_points_of(prm));
    if (rc) return rc;
  }
  const bool four = model_points_of(prm) == 4;
  FramesUse use(ctx, from, to);
  // host-bound stores: pull the key points of the pairs that will run PnP (From 3-D points, To pixels)
  rc = sepb_launch_frames_fetch(ctx, const_cast<sepb_frames*>(from), const_cast<sepb_frames*>(to), d_pairs, n_pairs, d_n_matches,
                                prm->min_inliers, nullptr, 2, 1);
  if (rc) return rc;
  return sepb_launch_pnp_ransac(ctx, view_of(from), view_of(to), d_pairs, n_pairs, d_n_matches, d_match_from,
                                d_match_to, d_n_unique_to, four ? ctx->table4 : ctx->table, four ? ctx->table4_h : ctx->table_h,
                                d_custom_samples, custom_h, d_guess,
                                d_only_if_ok, prm, d_status, d_n_inliers, d_pose, d_transform, d_inlier_mask, mask_words,
                                d_best_hyp, d_iters_run);
}

int sepb_pnp_ransac_dev(sepb_ctx* ctx, const sepb_frames* from, const sepb_frames* to, const int32_t* d_pairs,
                        int n_pairs, const int32_t* d_n_matches, const uint16_t* d_match_from,
                        const uint16_t* d_match_to, const int32_t* d_n_unique_to, const sepb_pnp_params* prm,
                        const int32_t* d_custom_samples, int custom_h, const double* d_guess, int32_t* d_status,
                        int32_t* d_n_inliers, double* d_pose, float* d_transform, uint32_t* d_inlier_mask,
                        int mask_words, int32_t* d_best_hyp, int32_t* d_iters_run) {
  return pnp_ransac_impl(ctx, from, to, d_pairs, n_pairs, d_n_matches, d_match_from, d_match_to, d_n_unique_to, prm,
                         d_custom_samples, custom_h, d_guess, nullptr, d_status, d_n_inliers, d_pose, d_transform,
                         d_inlier_mask, mask_words, d_best_hyp, d_iters_run);
}

// gprm == nullptr: pass 1 only (sepb_verify_pairs*); otherwise the two-pass driver of estimate_transformation
static int verify_dev_impl(sepb_ctx* ctx, const sepb_frames* from, const sepb_frames* to, const int32_t* d_pairs,
                           int n_pairs, const sepb_match_params* mprm, const sepb_pnp_params* pprm,
                           const sepb_guess_params* gprm, const int32_t* d_custom_samples, int custom_h,
                           const sepb_verify_out* o, int32_t* d_status_pass1) {
  if (!ctx || !from || !to || !mprm || !pprm || !o) return SEPB_E_INVALID;
  SEPB_REQUIRE(ctx, n_pairs >= 0, SEPB_E_SHAPE);
  if (n_pairs == 0) return SEPB_OK;
  SEPB_REQUIRE(ctx, o->status && o->n_matches && o->n_inliers && o->pose && o->transform && o->match_from &&
                        o->match_to, SEPB_E_INVALID);
  SEPB_CUDA(ctx, cudaSetDevice(ctx->device));
  if (gprm) SEPB_REQUIRE(ctx, gprm->image_width > 1 && gprm->image_height > 1 && gprm->win_size > 0, SEPB_E_INVALID);
  // n_unique_to (and the pass-2 guesses) live in the head of the context's grow-only scratch
  const size_t b_nu = align_up((size_t)n_pairs * 4, 256), b_g = gprm ? align_up((size_t)n_pairs * 48, 256) : 0;
  uint8_t* s0 = (uint8_t*)sepb_scratch(ctx, b_nu + b_g);
  if (!s0) return (int)cudaErrorMemoryAllocation;
  int32_t* d_nu = (int32_t*)s0;
  double* d_guess = (double*)(s0 + b_nu);
  const int mw = (from->max_feat + 31) / 32;
  // released when this function returns, i.e. after the last kernel that reads the stores (covariance, K2G) was enqueued
  FramesUse use(ctx, from, to);
  // host-bound stores: the matcher pulls the key points of the pairs that will reach PnP while it runs
  const LazyPull lp = sepb_lazy_pull_of(const_cast<sepb_frames*>(from), const_cast<sepb_frames*>(to), pprm->min_inliers);
  int rc = sepb_launch_match(ctx, view_of(from), view_of(to), d_pairs, n_pairs, mprm->nndr, mprm->nndr_strict,
                                     nullptr, nullptr, o->n_matches, o->match_from, o->match_to, d_nu, &lp);
  if (rc) return rc;
  rc = pnp_ransac_impl(ctx, from, to, d_pairs, n_pairs, o->n_matches, o->match_from, o->match_to, d_nu, pprm,
                       d_custom_samples, custom_h, nullptr, nullptr, o->status, o->n_inliers, o->pose, o->transform,
                       o->inlier_mask, mw, o->best_hyp, o->iters_run);
  const bool ba = ctx->ba.type == 1;
  if (ba) SEPB_REQUIRE(ctx, o->inlier_mask && o->covariance, SEPB_E_INVALID);
  // myRegistrationVis.cpp:1192-1370: the PnP result, its covariance and the inliers go through the 2-view bundle adjustment
  auto bundle = [&]() -> int {
    if (!ba) return SEPB_OK;
    int rcf = sepb_launch_frames_fetch(ctx, const_cast<sepb_frames*>(from), const_cast<sepb_frames*>(to), d_pairs, n_pairs, nullptr,
                                       0, o->status, 3, 3);
    if (rcf) return rcf;
    return sepb_launch_bundle_adjust(ctx, view_of(from), view_of(to), d_pairs, n_pairs, o->n_matches, o->match_from,
                                     o->match_to, o->covariance, pprm, &ctx->ba, o->status, o->n_inliers, o->pose,
                                     o->transform, o->inlier_mask, mw, nullptr);
  };
  auto covariance = [&]() -> int {
    if (!o->covariance) return SEPB_OK;
    SEPB_REQUIRE(ctx, o->inlier_mask != nullptr, SEPB_E_INVALID);
    // the 3-D residuals need the To frame's 3-D points of the verified pairs
    int rcf = sepb_launch_frames_fetch(ctx, const_cast<sepb_frames*>(from), const_cast<sepb_frames*>(to), d_pairs, n_pairs, nullptr,
                                       0, o->status, 2, 2);
    if (rcf) return rcf;
    return sepb_launch_pose_cov(ctx, view_of(from), view_of(to), d_pairs, n_pairs, o->status, o->n_matches, o->match_from,
                                o->match_to, o->inlier_mask, mw, o->transform, o->covariance);
  };
  if (rc) return rc;
  if (!gprm || ba) {
    rc = covariance();
    if (rc) return rc;
    rc = bundle();
    if (rc || !gprm) return rc;
  }
  // ---- pass 2 (stereoCamGeometricTools.cpp:153-164): guess = first result; only where the first pass succeeded
  if (d_status_pass1)
    SEPB_CUDA(ctx, cudaMemcpyAsync(d_status_pass1, o->status, (size_t)n_pairs * 4, cudaMemcpyDeviceToDevice, ctx->stream));
  // the window matcher reads pixels, 3-D points and octaves of both frames of every pair that passed the first call
  rc = sepb_launch_frames_fetch(ctx, const_cast<sepb_frames*>(from), const_cast<sepb_frames*>(to), d_pairs, n_pairs, nullptr, 0,
                                o->status, 7, 7);
  if (rc) return rc;
  rc = sepb_launch_guess_match(ctx, view_of(from), view_of(to), d_pairs, n_pairs, o->status, o->transform, pprm,
                               gprm->image_width, gprm->image_height, gprm->win_size, mprm->nndr, o->n_matches,
                               o->match_from, o->match_to, d_nu, d_guess);
  if (rc) return rc;
  rc = pnp_ransac_impl(ctx, from, to, d_pairs, n_pairs, o->n_matches, o->match_from, o->match_to, d_nu, pprm, nullptr, 0,
                       d_guess, o->status, o->status, o->n_inliers, o->pose, o->transform, o->inlier_mask, mw, o->best_hyp,
                       o->iters_run);
  if (rc) return rc;
  rc = covariance();
  if (rc) return rc;
  return bundle();
}

int sepb_verify_pairs_dev(sepb_ctx* ctx, const sepb_frames* from, const sepb_frames* to, const int32_t* d_pairs,
                          int n_pairs, const sepb_match_params* mprm, const sepb_pnp_params* pprm,
                          const int32_t* d_custom_samples, int custom_h, const sepb_verify_out* o) {
  return verify_dev_impl(ctx, from, to, d_pairs, n_pairs, mprm, pprm, nullptr, d_custom_samples, custom_h, o, nullptr);
}

int sepb_estimate_transformations_dev(sepb_ctx* ctx, const sepb_frames* from, const sepb_frames* to,
                                      const int32_t* d_pairs, int n_pairs, const sepb_match_params* mprm,
                                      const sepb_pnp_params* pprm, const sepb_guess_params* gprm,
                                      const sepb_verify_out* d_out, int32_t* d_status_pass1) {
  if (!gprm) return SEPB_E_INVALID;
  return verify_dev_impl(ctx, from, to, d_pairs, n_pairs, mprm, pprm, gprm, nullptr, 0, d_out, d_status_pass1);
}

static int verify_pairs_impl(sepb_ctx* ctx, const sepb_frames* from, const sepb_frames* to, const int32_t* pairs,
                             int n_pairs, const sepb_match_params* mprm, const sepb_pnp_params* pprm,
                             const int32_t* custom_samples, int custom_h, sepb_verify_out* out, bool wait,
                             const sepb_guess_params* gprm = nullptr, int32_t* status_pass1 = nullptr) {
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
  // head of the scratch: n_unique_to and pass-2 guesses (used by verify_dev_impl)
  const size_t b_nu = align_up(P * 4, 256) + (gprm ? align_up(P * 48, 256) : 0);
  const size_t b_pairs = align_up(P * 8, 256), b_i = align_up(P * 4, 256), b_pose = align_up(P * 48, 256),
               b_T = align_up(P * 48, 256), b_m = align_up(P * F * 2, 256), b_mask = align_up(P * MW * 4, 256),
               b_cs = custom_samples ? align_up(P * custom_h * model_points_of(pprm) * 4, 256) : 0;
  const bool want_cov = out->covariance || ctx->ba.type == 1;
  const size_t b_cov = want_cov ? align_up(P * 288, 256) : 0;
  const size_t total = b_nu + b_pairs + 6 * b_i + b_pose + b_T + 2 * b_m + b_mask + b_cs + b_cov;
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
  int32_t* d_st1 = (int32_t*)s; s += b_i;
  d.pose = (double*)s; s += b_pose;
  d.transform = (float*)s; s += b_T;
  d.match_from = (uint16_t*)s; s += b_m;
  d.match_to = (uint16_t*)s; s += b_m;
  d.inlier_mask = (uint32_t*)s; s += b_mask;
  if (want_cov) { d.covariance = (double*)s; s += b_cov; }
  int32_t* d_cs = nullptr;
  if (custom_samples) {
    d_cs = (int32_t*)s;
    SEPB_CUDA(ctx, cudaMemcpyAsync(d_cs, custom_samples, P * custom_h * model_points_of(pprm) * 4, cudaMemcpyHostToDevice, ctx->stream));
  }
  SEPB_CUDA(ctx, cudaMemcpyAsync(d_pairs, pairs, P * 8, cudaMemcpyHostToDevice, ctx->stream));
  rc = verify_dev_impl(ctx, from, to, d_pairs, n_pairs, mprm, pprm, gprm, d_cs, custom_h, &d, gprm ? d_st1 : nullptr);
  if (rc) return rc;
  if (gprm && status_pass1)
    SEPB_CUDA(ctx, cudaMemcpyAsync(status_pass1, d_st1, P * 4, cudaMemcpyDeviceToHost, ctx->stream));
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
  D2H(covariance, P * 288);
#undef D2H
  if (wait) SEPB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return SEPB_OK;
}

int sepb_verify_pairs(sepb_ctx* ctx, const sepb_frames* from, const sepb_frames* to, const int32_t* pairs, int n_pairs,
                      const sepb_match_params* mprm, const sepb_pnp_params* pprm, const int32_t* custom_samples,
                      int custom_h, sepb_verify_out* out) {
  return verify_pairs_impl(ctx, from, to, pairs, n_pairs, mprm, pprm, custom_samples, custom_h, out, true);
}

int sepb_estimate_transformations(sepb_ctx* ctx, const sepb_frames* from, const sepb_frames* to, const int32_t* pairs,
                                  int n_pairs, const sepb_match_params* mprm, const sepb_pnp_params* pprm,
                                  const sepb_guess_params* gprm, sepb_verify_out* out, int32_t* status_pass1) {
  if (!gprm) return SEPB_E_INVALID;
  return verify_pairs_impl(ctx, from, to, pairs, n_pairs, mprm, pprm, nullptr, 0, out, true, gprm, status_pass1);
}

// K2G alone on host buffers: pass-2 correspondences for explicit first-pass transforms (parity tests feed the oracle's)
int sepb_guess_match_pairs(sepb_ctx* ctx, const sepb_frames* from, const sepb_frames* to, const int32_t* pairs,
                           int n_pairs, const float* transforms, const sepb_pnp_params* pprm,
                           const sepb_guess_params* gprm, float nndr, int32_t* n_matches, uint16_t* match_from,
                           uint16_t* match_to, double* guess_out) {
  if (!ctx || !from || !to || !pprm || !gprm) return SEPB_E_INVALID;
  SEPB_REQUIRE(ctx, n_pairs >= 0, SEPB_E_SHAPE);
  if (n_pairs == 0) return SEPB_OK;
  SEPB_REQUIRE(ctx, pairs && transforms && n_matches && match_from && match_to, SEPB_E_INVALID);
  SEPB_REQUIRE(ctx, gprm->image_width > 1 && gprm->image_height > 1 && gprm->win_size > 0, SEPB_E_INVALID);
  for (int i = 0; i < n_pairs; ++i)
    SEPB_REQUIRE(ctx, pairs[2 * i] >= 0 && pairs[2 * i] < from->size && pairs[2 * i + 1] >= 0 &&
                          pairs[2 * i + 1] < to->size, SEPB_E_SHAPE);
  std::lock_guard<std::mutex> lk(ctx->mu);
  SEPB_CUDA(ctx, cudaSetDevice(ctx->device));
  const size_t P = (size_t)n_pairs, F = (size_t)from->max_feat;
  const size_t b_pairs = align_up(P * 8, 256), b_T = align_up(P * 48, 256), b_i = align_up(P * 4, 256),
               b_m = align_up(P * F * 2, 256), b_g = align_up(P * 48, 256);
  uint8_t* s = (uint8_t*)sepb_scratch(ctx, b_pairs + b_T + 2 * b_i + 2 * b_m + b_g);
  if (!s) return (int)cudaErrorMemoryAllocation;
  int32_t* d_pairs = (int32_t*)s; s += b_pairs;
  float* d_T = (float*)s; s += b_T;
  int32_t* d_nm = (int32_t*)s; s += b_i;
  int32_t* d_nu = (int32_t*)s; s += b_i;
  uint16_t* d_mf = (uint16_t*)s; s += b_m;
  uint16_t* d_mt = (uint16_t*)s; s += b_m;
  double* d_g = (double*)s;
  SEPB_CUDA(ctx, cudaMemcpyAsync(d_pairs, pairs, P * 8, cudaMemcpyHostToDevice, ctx->stream));
  SEPB_CUDA(ctx, cudaMemcpyAsync(d_T, transforms, P * 48, cudaMemcpyHostToDevice, ctx->stream));
  int rc;
  {
    FramesUse use(ctx, from, to);
    rc = sepb_launch_guess_match(ctx, view_of(from), view_of(to), d_pairs, n_pairs, nullptr, d_T, pprm, gprm->image_width,
                                 gprm->image_height, gprm->win_size, nndr, d_nm, d_mf, d_mt, d_nu, d_g);
  }
  if (rc) return rc;
  SEPB_CUDA(ctx, cudaMemcpyAsync(n_matches, d_nm, P * 4, cudaMemcpyDeviceToHost, ctx->stream));
  SEPB_CUDA(ctx, cudaMemcpyAsync(match_from, d_mf, P * F * 2, cudaMemcpyDeviceToHost, ctx->stream));
  SEPB_CUDA(ctx, cudaMemcpyAsync(match_to, d_mt, P * F * 2, cudaMemcpyDeviceToHost, ctx->stream));
  if (guess_out) SEPB_CUDA(ctx, cudaMemcpyAsync(guess_out, d_g, P * 48, cudaMemcpyDeviceToHost, ctx->stream));
  SEPB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return SEPB_OK;
}

void sepb_ba_params_init(sepb_ba_params* b, double baseline) {
  if (!b) return;
  b->type = 1;
  b->iterations = 20;
  b->pixel_variance = 1.0;
  b->robust_kernel_delta = 8.0;
  b->baseline = baseline;
  b->variance_ignored = 0;
  b->reserved_ = 0;
}

static int check_ba_params(sepb_ctx* ctx, const sepb_ba_params* b) {
  SEPB_REQUIRE(ctx, b->type == 1, SEPB_E_INVALID);   // 2 = cvsba: not built
  SEPB_REQUIRE(ctx, b->iterations >= 1 && b->iterations <= 1000 && b->pixel_variance > 0, SEPB_E_INVALID);
  return SEPB_OK;
}

int sepb_ctx_set_bundle_adjustment(sepb_ctx* ctx, const sepb_ba_params* b) {
  if (!ctx) return SEPB_E_INVALID;
  std::lock_guard<std::mutex> lk(ctx->mu);
  if (!b || b->type == 0) {
    ctx->ba = sepb_ba_params{};
    return SEPB_OK;
  }
  int rc = check_ba_params(ctx, b);
  if (rc) return rc;
  ctx->ba = *b;
  return SEPB_OK;
}

int sepb_bundle_adjust_dev(sepb_ctx* ctx, const sepb_frames* from, const sepb_frames* to, const int32_t* d_pairs,
                           int n_pairs, const int32_t* d_n_matches, const uint16_t* d_match_from,
                           const uint16_t* d_match_to, const double* d_covariance, const sepb_pnp_params* pprm,
                           const sepb_ba_params* bprm, int32_t* d_status, int32_t* d_n_inliers, double* d_pose,
                           float* d_transform, uint32_t* d_inlier_mask, int mask_words, double* d_stats) {
  if (!ctx || !from || !to || !pprm || !bprm) return SEPB_E_INVALID;
  SEPB_REQUIRE(ctx, n_pairs >= 0, SEPB_E_SHAPE);
  if (n_pairs == 0) return SEPB_OK;
  SEPB_REQUIRE(ctx, d_pairs && d_n_matches && d_match_from && d_match_to && d_status && d_n_inliers && d_pose &&
                        d_transform && d_inlier_mask, SEPB_E_INVALID);
  SEPB_REQUIRE(ctx, mask_words * 32 >= from->max_feat, SEPB_E_SHAPE);
  int rc = check_pnp_params(ctx, pprm);
  if (rc) return rc;
  rc = check_ba_params(ctx, bprm);
  if (rc) return rc;
  SEPB_CUDA(ctx, cudaSetDevice(ctx->device));
  FramesUse use(ctx, from, to);
  rc = sepb_launch_frames_fetch(ctx, const_cast<sepb_frames*>(from), const_cast<sepb_frames*>(to), d_pairs, n_pairs, nullptr, 0,
                                d_status, 3, 3);
  if (rc) return rc;
  return sepb_launch_bundle_adjust(ctx, view_of(from), view_of(to), d_pairs, n_pairs, d_n_matches, d_match_from, d_match_to,
                                   d_covariance, pprm, bprm, d_status, d_n_inliers, d_pose, d_transform, d_inlier_mask,
                                   mask_words, d_stats);
}

int sepb_bundle_adjust_pairs(sepb_ctx* ctx, const sepb_frames* from, const sepb_frames* to, const int32_t* pairs,
                             int n_pairs, const int32_t* n_matches, const uint16_t* match_from, const uint16_t* match_to,
                             const double* covariance, const sepb_pnp_params* pprm, const sepb_ba_params* bprm,
                             int32_t* status, int32_t* n_inliers, double* pose, float* transform, uint32_t* inlier_mask,
                             int mask_words, double* stats) {
  if (!ctx || !from || !to || !pprm || !bprm) return SEPB_E_INVALID;
  SEPB_REQUIRE(ctx, n_pairs >= 0, SEPB_E_SHAPE);
  if (n_pairs == 0) return SEPB_OK;
  SEPB_REQUIRE(ctx, pairs && n_matches && match_from && match_to && status && n_inliers && pose && transform &&
                        inlier_mask, SEPB_E_INVALID);
  for (int i = 0; i < n_pairs; ++i)
    SEPB_REQUIRE(ctx, pairs[2 * i] >= 0 && pairs[2 * i] < from->size && pairs[2 * i + 1] >= 0 &&
                          pairs[2 * i + 1] < to->size && n_matches[i] >= 0 && n_matches[i] <= from->max_feat, SEPB_E_SHAPE);
  std::lock_guard<std::mutex> lk(ctx->mu);
  SEPB_CUDA(ctx, cudaSetDevice(ctx->device));
  const size_t P = (size_t)n_pairs, F = (size_t)from->max_feat, MW = (size_t)mask_words;
  const size_t b_pairs = align_up(P * 8, 256), b_i = align_up(P * 4, 256), b_pose = align_up(P * 48, 256),
               b_T = align_up(P * 48, 256), b_m = align_up(P * F * 2, 256), b_mask = align_up(P * MW * 4, 256),
               b_cov = align_up(P * 288, 256), b_st = align_up(P * 32, 256);
  uint8_t* s = (uint8_t*)sepb_scratch(ctx, b_pairs + 3 * b_i + b_pose + b_T + 2 * b_m + b_mask + b_cov + b_st);
  if (!s) return (int)cudaErrorMemoryAllocation;
  int32_t* d_pairs = (int32_t*)s; s += b_pairs;
  int32_t* d_nm = (int32_t*)s; s += b_i;
  int32_t* d_st = (int32_t*)s; s += b_i;
  int32_t* d_ni = (int32_t*)s; s += b_i;
  double* d_pose = (double*)s; s += b_pose;
  float* d_T = (float*)s; s += b_T;
  uint16_t* d_mf = (uint16_t*)s; s += b_m;
  uint16_t* d_mt = (uint16_t*)s; s += b_m;
  uint32_t* d_mask = (uint32_t*)s; s += b_mask;
  double* d_cov = (double*)s; s += b_cov;
  double* d_stats = (double*)s;
#define H2D(dst, src, bytes) SEPB_CUDA(ctx, cudaMemcpyAsync((dst), (src), (bytes), cudaMemcpyHostToDevice, ctx->stream))
  H2D(d_pairs, pairs, P * 8);
  H2D(d_nm, n_matches, P * 4);
  H2D(d_st, status, P * 4);
  H2D(d_ni, n_inliers, P * 4);
  H2D(d_pose, pose, P * 48);
  H2D(d_T, transform, P * 48);
  H2D(d_mf, match_from, P * F * 2);
  H2D(d_mt, match_to, P * F * 2);
  H2D(d_mask, inlier_mask, P * MW * 4);
  if (covariance) H2D(d_cov, covariance, P * 288);
#undef H2D
  SEPB_CUDA(ctx, cudaMemsetAsync(d_stats, 0, P * 32, ctx->stream));
  int rc = sepb_bundle_adjust_dev(ctx, from, to, d_pairs, n_pairs, d_nm, d_mf, d_mt, covariance ? d_cov : nullptr, pprm, bprm,
                                  d_st, d_ni, d_pose, d_T, d_mask, mask_words, d_stats);
  if (rc) return rc;
#define D2H(dst, src, bytes) SEPB_CUDA(ctx, cudaMemcpyAsync((dst), (src), (bytes), cudaMemcpyDeviceToHost, ctx->stream))
  D2H(status, d_st, P * 4);
  D2H(n_inliers, d_ni, P * 4);
  D2H(pose, d_pose, P * 48);
  D2H(transform, d_T, P * 48);
  D2H(inlier_mask, d_mask, P * MW * 4);
  if (stats) D2H(stats, d_stats, P * 32);
#undef D2H
  SEPB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return SEPB_OK;
}

void sepb_guess_params_init(sepb_guess_params* g, int image_width, int image_height) {
  if (!g) return;
  g->image_width = image_width;
  g->image_height = image_height;
  g->win_size = 20.f;   // Vis/CorGuessWinSize (RTAB-Map <= 0.20; 40 in later versions)
}

int sepb_verify_pairs_async(sepb_ctx* ctx, const sepb_frames* from, const sepb_frames* to, const int32_t* pairs,
                            int n_pairs, const sepb_match_params* mprm, const sepb_pnp_params* pprm,
                            const int32_t* custom_samples, int custom_h, sepb_verify_out* out) {
  return verify_pairs_impl(ctx, from, to, pairs, n_pairs, mprm, pprm, custom_samples, custom_h, out, false);
}

}  // extern "C"
