/* sepb200.h -- C ABI of the B200-native separator-detection hot path.
 *
 * Drop-in boundary for DOOR-SLAM's multi_robot_separators front-end.
 * Every entry point cites the reference interface it replaces (paths under
 * /root/reference/ros_ws/src/multi_robot_separators/, abbreviated R/).
 *
 * Conventions
 *  - plain C, caller-owned buffers, opaque handles; no torch / OpenCV / ROS types.
 *  - return value: 0 ok, <0 invalid argument / shape (SEPB_E_*), >0 CUDA error code
 *    (cudaError_t); text via sepb_last_error().  Nothing here aborts the process (the
 *    reference UASSERTs on shape mismatch, R/src/myRegistrationVis.cpp:859-860).
 *  - "_dev" entry points take DEVICE pointers, enqueue on the context's stream and do not
 *    synchronise; the others take HOST pointers, copy H2D/D2H on that stream and return when
 *    results are in the caller's buffers.
 *  - there is no CPU fallback: every call fails with a CUDA error if no sm_100 device is present.
 */
#ifndef SEPB200_H
#define SEPB200_H

#include <stdint.h>

#if defined(__GNUC__)
#define SEPB_API __attribute__((visibility("default")))
#else
#define SEPB_API
#endif

#ifdef __cplusplus
extern "C" {
#endif

#define SEPB_OK 0
#define SEPB_E_INVALID (-1)   /* null pointer / negative size */
#define SEPB_E_SHAPE (-2)     /* inconsistent shapes or capacity exceeded */
#define SEPB_E_STATE (-3)     /* handle used before it was filled / NCCL not available for a multi-device group */
#define SEPB_E_NCCL (-4)      /* an NCCL call failed; text via sepb_group_last_error() */

#define SEPB_DESC_BYTES 32    /* 256-bit binary descriptors (ORB / BRIEF-32), R/msg/Descriptors.msg */
#define SEPB_MAX_FEATURES 4096

typedef struct sepb_ctx sepb_ctx;        /* one device + one stream + scratch */
typedef struct sepb_frames sepb_frames;  /* device-resident keyframe feature store of one robot */
typedef struct sepb_db sepb_db;          /* device-resident NetVLAD descriptor database */
typedef struct sepb_group sepb_group;    /* the devices of one box + their NCCL communicator */

/* ---- context -------------------------------------------------------------------------- */
SEPB_API int sepb_ctx_create(int device, sepb_ctx** out);
/* Same, with the streams at the device's highest priority: a context for short, latency-bound work (the place-recognition
 * search of the NEXT request round) whose thread blocks are dispatched ahead of the pending blocks of a long verification
 * batch running on another context of the same device. */
SEPB_API int sepb_ctx_create_priority(int device, sepb_ctx** out);
SEPB_API void sepb_ctx_destroy(sepb_ctx* ctx);
SEPB_API const char* sepb_last_error(const sepb_ctx* ctx);
SEPB_API void* sepb_ctx_stream(const sepb_ctx* ctx);           /* cudaStream_t */
SEPB_API int sepb_ctx_sync(sepb_ctx* ctx);
/* number of kernels this library launched on ctx since creation (bench.py "gpu_launches") */
SEPB_API int64_t sepb_ctx_launch_count(const sepb_ctx* ctx);

/* options (sepb_ctx_set_option): which kernel runs the Hamming matcher.  Both produce identical results.
 *   SEPB_MATCH_POPC   hamming_knn.cu  -- INT/XU-pipe kernel (__popc, 128-bit shared-memory broadcasts)
 *   SEPB_MATCH_TENSOR hamming_mma.cu  -- the same distances as an exact int8 tcgen05 GEMM
 *   SEPB_MATCH_AUTO   tensor kernel whenever the frame sizes fit its shared-memory budget (default) */
#define SEPB_OPT_MATCH_KERNEL 1
#define SEPB_MATCH_AUTO (-1)
#define SEPB_MATCH_POPC 0
#define SEPB_MATCH_TENSOR 1
/* how PnP-RANSAC is scheduled (identical results):
 *   SEPB_PNP_SCHED_PERSISTENT  (default) one persistent CTA per SM; warps pop pairs from a device work list and, once it
 *                              is empty, compute chunks of 32 hypotheses for the pairs still running (the chunks of a
 *                              pair are independent; only the scan over their inlier counts is sequential)
 *   SEPB_PNP_SCHED_PER_PAIR    one one-warp CTA per pair, every chunk by that warp (reference scheduling) */
#define SEPB_OPT_PNP_SCHEDULER 2
#define SEPB_PNP_SCHED_PERSISTENT 0
#define SEPB_PNP_SCHED_PER_PAIR 1
/* accumulator tile width of the tensor matcher: 0 (default) = 256 columns whenever the To frames fit its shared-memory
 * budget (To frames up to ~1400 features), else 128; 128 forces the narrow geometry.  Identical results. */
#define SEPB_OPT_MATCH_TILE 3
/* how the tensor matcher walks a batch (identical results):
 *   SEPB_MATCH_SCHED_PERSISTENT  (default) one CTA per SM runs through its share of the pairs; pair p's tail (ratio test,
 *                                uniqueness, list compaction) overlaps pair p + 1's first tiles
 *   SEPB_MATCH_SCHED_PER_PAIR    one CTA per pair */
#define SEPB_OPT_MATCH_SCHEDULER 4
#define SEPB_MATCH_SCHED_PERSISTENT 0
#define SEPB_MATCH_SCHED_PER_PAIR 1
SEPB_API int sepb_ctx_set_option(sepb_ctx* ctx, int option, int value);
SEPB_API int sepb_ctx_get_option(const sepb_ctx* ctx, int option, int* value);

/* ---- keyframe feature store ------------------------------------------------------------
 * Mirrors what GetFeatsAndDesc / FindMatches carry per keyframe (R/srv/GetFeatsAndDesc.srv:4-6,
 * R/msg/Descriptors.msg, R/msg/KeyPointVec.msg, R/msg/KeyPoint3DVec.msg): descriptors rows x 32
 * u8 (row-major, as MsgConversion.cpp:113-116 aliases them), keypoint pixel positions, 3-D
 * points in the robot base frame (NaN/inf = stereo failure, dropped like
 * util3d::isFinite), octaves.  Fixed row stride max_feat, per-frame count <= max_feat. */
SEPB_API int sepb_frames_create(sepb_ctx* ctx, int capacity, int max_feat, sepb_frames** out);
SEPB_API void sepb_frames_destroy(sepb_frames* f);
SEPB_API int sepb_frames_size(const sepb_frames* f);
/* upload n frames starting at slot `first`; arrays are [n][max_feat][..] host memory; octave may be NULL */
SEPB_API int sepb_frames_upload(sepb_ctx* ctx, sepb_frames* f, int first, int n, const uint8_t* desc,
                       const float* kp2d, const float* kp3d, const int32_t* octave, const int32_t* count);

/* Same, without waiting: the copies run on the context's copy stream, concurrently with kernels already enqueued
 * (the next round's keyframes travel while the current round is verified).  The host arrays must stay valid (and
 * should be pinned) until sepb_frames_wait() or any later synchronising call that uses the store returns.  Kernels
 * enqueued afterwards wait for the upload on the device; an upload waits for kernels that still read the store. */
SEPB_API int sepb_frames_upload_async(sepb_ctx* ctx, sepb_frames* f, int first, int n, const uint8_t* desc,
                                      const float* kp2d, const float* kp3d, const int32_t* octave, const int32_t* count);
SEPB_API int sepb_frames_wait(sepb_ctx* ctx, sepb_frames* f);
/* Host-resident keyframes with lazy key points.  Like sepb_frames_upload_async, but only descriptors and counts are
 * copied: kp2d / kp3d / octave (which MUST be pinned host memory -- cudaHostAlloc or cudaHostRegister -- covering
 * [n][max_feat][..]) stay where they are and the device pulls, inside the verification calls, just the rows of the
 * frames whose pairs reach PnP (>= min_inliers matches; myRegistrationVis.cpp:1069-1070) -- 32 instead of 52 bytes per
 * feature cross the bus for every frame, the remaining 20 only for the few percent that need them.  The arrays must
 * stay valid and unchanged until the next bind / upload of these slots.  Results are identical to a full upload.
 * One binding per store at a time (a later bind replaces it).  EstTransform.srv:1-6 is the payload this trims. */
SEPB_API int sepb_frames_bind_host_async(sepb_ctx* ctx, sepb_frames* f, int first, int n, const uint8_t* desc,
                                         const float* kp2d, const float* kp3d, const int32_t* octave, const int32_t* count);
/* bytes the device has pulled from the bound host arrays since the last bind (synchronises the context stream) */
SEPB_API int64_t sepb_frames_fetched_bytes(sepb_ctx* ctx, const sepb_frames* f);

/* ---- K2: brute-force Hamming 2-NN + NNDR + one-to-one filter ----------------------------
 * Replaces the global matching block R/src/myRegistrationVis.cpp:839-894 in brute-force mode
 * (VWDictionary::addNewWords(descriptorsTo, 2) == cv::BFMatcher(NORM_HAMMING).knnMatch(k=2),
 * also the call at :580-581 / :739-740).  For pair p = (from_frame, to_frame): every To
 * descriptor is a query, From descriptors are the train set; ties -> lowest train index.
 * nndr_strict = 0: accept iff d1 <= nndr*d2 (global path); 1: d1 < nndr*d2 (:584,:744).
 * Outputs (all optional except n_matches; device or host according to the entry point):
 *   knn_idx  [P][max_feat_to][2] i32, knn_dist [P][max_feat_to][2] i32  (-1 = missing)
 *   n_matches[P] i32; match_from/match_to [P][max_feat] u16: the PnP correspondence list,
 *   ascending From index, rows with non-finite From 3-D dropped
 *   n_unique_to[P] i32: To rows left after the uniqueness filter (:882-894). */
typedef struct sepb_match_params {
  float nndr;          /* Vis/CorNNDR: 0.8 (RTAB-Map >= 0.20) or 0.6 */
  int32_t nndr_strict; /* 0 global path, 1 guess path comparison */
} sepb_match_params;

SEPB_API int sepb_match_pairs(sepb_ctx* ctx, const sepb_frames* from, const sepb_frames* to, const int32_t* pairs /*[P][2]*/,
                     int n_pairs, const sepb_match_params* prm, int32_t* knn_idx, int32_t* knn_dist,
                     int32_t* n_matches, uint16_t* match_from, uint16_t* match_to, int32_t* n_unique_to);
SEPB_API int sepb_match_pairs_dev(sepb_ctx* ctx, const sepb_frames* from, const sepb_frames* to, const int32_t* d_pairs,
                         int n_pairs, const sepb_match_params* prm, int32_t* d_knn_idx, int32_t* d_knn_dist,
                         int32_t* d_n_matches, uint16_t* d_match_from, uint16_t* d_match_to,
                         int32_t* d_n_unique_to);

/* ---- K1: NetVLAD nearest-neighbour search against the peer database --------------------------
 * Replaces DataHandler.find_matches (R/scripts/data_handler.py:166-209): `cdist(local, received)`
 * (f64 Euclidean, :170), masks to +inf for local_kf_already_used rows (:178-179),
 * other_kf_already_used columns (:180-181) and frames_kept_pairs_ignored cells (:183-184), per-row
 * argmin (:187) -- extended to the k smallest per row -- and the greedy selection loop (:191-205).
 * Rows of the distance matrix = `queries`, columns = database rows.  Distances are bit-identical to
 * scipy's cdist (sequential f64 summation, sqrt); ties resolve to the lowest index (np.argmin).
 * The database is append-only like received_descriptors (:300-301) and takes the float64[] wire
 * payload of FindMatches.srv:1 as is. */
SEPB_API int sepb_db_create(sepb_ctx* ctx, int dim, int capacity, sepb_db** out);
SEPB_API void sepb_db_destroy(sepb_db* db);
SEPB_API int sepb_db_size(const sepb_db* db);
SEPB_API int sepb_db_clear(sepb_db* db);   /* forget all rows (node restart); capacity is kept */
SEPB_API int sepb_db_append(sepb_ctx* ctx, sepb_db* db, const double* rows, int n);  /* host rows [n][dim] */
/* row-sharded databases: global index of this shard's row 0 (indices returned are global) */
SEPB_API int sepb_db_set_base(sepb_db* db, int64_t base);
/* other_kf_already_used as a byte mask over the database rows [size]; NULL clears */
SEPB_API int sepb_db_set_col_mask(sepb_ctx* ctx, sepb_db* db, const uint8_t* mask);

/* k <= 16.  row_mask: [nq] bytes or NULL; ignored: [n_ignored][2] (query row, GLOBAL db row) or NULL.
 * out_idx [nq][k] i32 (-1 = none / masked), out_dist [nq][k] f64 (+inf = none), ascending (dist, idx).
 * mode: SEPB_NN_EXACT = fp64 scan (K1b) ; SEPB_NN_TENSOR = tcgen05 split-bf16 filter (K1) + exact
 * fp64 re-rank of the surviving candidates, with per-row fallback to the exact scan whenever the
 * filter's error bound cannot prove the candidate set complete.  Results are identical in both. */
#define SEPB_NN_EXACT 0
#define SEPB_NN_TENSOR 1
SEPB_API int sepb_netvlad_topk(sepb_ctx* ctx, const sepb_db* db, const double* queries, int nq, int k, int mode,
                               const uint8_t* row_mask, const int32_t* ignored, int n_ignored, int32_t* out_idx,
                               double* out_dist);
/* device form: d_queries [nq][dim] f64; d_ign_start [nq+1] / d_ign_col CSR of ignored cells (or NULL) */
SEPB_API int sepb_netvlad_topk_dev(sepb_ctx* ctx, const sepb_db* db, const double* d_queries, int nq, int k, int mode,
                                   const uint8_t* d_row_mask, const int32_t* d_ign_start, const int32_t* d_ign_col,
                                   int32_t* d_out_idx, double* d_out_dist);
/* Allocation-free variant of sepb_netvlad_topk_dev(mode = SEPB_NN_EXACT).  It has NO counterpart in the reference: it
 * was added for callers that cannot call cudaMalloc (code under CUDA graph
 * capture, tight serving loops, harnesses that forbid it).  Same arguments and bit-identical results; the per-split
 * partial lists of the two-pass scan live in a fixed 8 MiB device array inside the library (held by every GPU that
 * loads it until the process exits) instead of the context's grow-only scratch, and the call performs no cudaMalloc /
 * cudaFree / allocating library call.  Limit: align256(splits * nq * k * 8) + align256(splits * nq * k * 4) <= 8 MiB,
 * with splits = clamp(ceil(2 * SMs / ceil(nq / 128)), 1, min(64, ceil(db size / 32))); beyond it the call returns
 * SEPB_E_SHAPE and sepb_last_error() says so (on 148 SMs: every nq fits for k <= 9, every nq <= 6 241 for k <= 16; the
 * calls refused are 10 <= k <= 16 with some nq in 6 242 .. 37 760 -- split those into smaller batches).  The first call on a device resolves the array's device address (cudaGetSymbolAddress, no allocation) and
 * the launch attributes: make one call, nq = 0 is enough, before a capture starts.  All calls on a device share the
 * array: calls that overlap on the GPU (two streams, contexts or host threads) race -- serialise them. */
SEPB_API int sepb_netvlad_topk_exact_static_dev(sepb_ctx* ctx, const sepb_db* db, const double* d_queries, int nq, int k,
                                                const uint8_t* d_row_mask, const int32_t* d_ign_start,
                                                const int32_t* d_ign_col, int32_t* d_out_idx, double* d_out_dist);
/* Test / profiling hook of the K1 filter (netvlad_gemm.cu): the SEPB_NN_TENSOR search on host queries, plus
 * out_v [nq][db size] f32 = the raw filter values ||d||^2 - 2 q.d of the tcgen05 GEMM (NULL to skip) and
 * stats[4] = {candidates re-ranked in f64, rows proven complete by the filter, rows sent to the exact scan, 0}.
 * Lets the tests measure the GEMM's real error against the bound E the filter relies on (data_handler.py:170). */
SEPB_API int sepb_netvlad_filter_debug(sepb_ctx* ctx, const sepb_db* db, const double* queries, int nq, int k,
                                       float* out_v, int32_t* out_idx, double* out_dist, int32_t* stats);
/* merge `lists` per-shard results (layout [lists][nq][k], e.g. after an all-gather) -> [nq][k] */
SEPB_API int sepb_topk_merge_dev(sepb_ctx* ctx, const int32_t* d_in_idx, const double* d_in_dist, int lists, int nq,
                                 int k, int32_t* d_out_idx, double* d_out_dist);
/* the selection loop of data_handler.py:191-205 over per-row minima (host, O(nq log nq)):
 * writes up to max_matches_nb (query_row, db_row) pairs, returns their number (or <0). */
SEPB_API int sepb_select_matches(const int32_t* row_min_idx, const double* row_min_dist, int nq,
                                 double netvlad_distance, int max_matches_nb, int32_t* out_pairs);

/* ---- K3: batched PnP-RANSAC + refit ------------------------------------------------------
 * Replaces the PnP block R/src/myRegistrationVis.cpp:1055-1112: RTAB-Map
 * util3d::estimateMotion3DTo2D -> cv::solvePnPRansac(useExtrinsicGuess=true, iterations,
 * reprojError, confidence 0.99, flags=ITERATIVE => 5-point EPnP minimal sets) -> final
 * ITERATIVE refit on the inliers seeded by the caller's guess, then T = (local * pnp)^-1.
 * Correspondences come from K2's match lists (ascending From index).  Rectified cameras only
 * (zero distortion), like the stereo model the reference builds (stereoCamGeometricTools.cpp:77). */
typedef struct sepb_pnp_params {
  double fx, fy, cx, cy;  /* intrinsics of the To camera, cam_.left() (stereoCamGeometricTools.cpp:142-143) */
  int32_t iterations;     /* Vis/Iterations, RTAB-Map default 300 */
  int32_t min_inliers;    /* Vis/MinInliers = separators_min_inliers (launch default 5); must be >= 4 */
  double reproj_error;    /* Vis/PnPReprojError, default 2.0 px */
  double confidence;      /* 0.99, hard-coded in util3d::solvePnPRansac */
  double guess_rvec[3];   /* (guess * localTransform)^-1 as Rodrigues vector ... */
  double guess_tvec[3];   /* ... and translation; sepb_pnp_params_init sets the identity guess of pass 1 */
  double local_R[9];      /* camera localTransform rotation (optical -> base), row-major */
  double local_t[3];
  int32_t flags;          /* Vis/PnPFlags (R/src/myRegistrationVis.cpp:103,1084): SEPB_PNP_ITERATIVE (0, default) = 5-point
                           * EPnP minimal sets + ITERATIVE refit from the guess; SEPB_PNP_P3P (2) / SEPB_PNP_AP3P (5) =
                           * 4-point P3P minimal sets + EPnP refit over the consensus set (what cv::solvePnPRansac does
                           * for these flags).  Exactly four correspondences always take the single P3P solve. */
  int32_t refine_iterations; /* Vis/PnPRefineIterations (:104,1085; 0 in the pinned RTAB-Map, 1 in older ones): rounds of
                           * RTAB-Map's refineModel loop after RANSAC -- refit on the inliers, re-select the inliers of ALL
                           * points with threshold min(reproj_error, refine_sigma * std of the errors) -- until the set stops
                           * changing.  SEPB_PNP_ITERATIVE only. */
  double refine_sigma;    /* 3.0, hard-coded in util3d::solvePnPRansac */
  int32_t minimal_model;  /* 0 (default): OpenCV >= 3.3 semantics, the refit over the inliers is returned; 1: OpenCV 3.2
                           * semantics (the reference image's version), the winning minimal-set model is returned */
  int32_t reserved_;
} sepb_pnp_params;
#define SEPB_PNP_ITERATIVE 0
#define SEPB_PNP_P3P 2
#define SEPB_PNP_AP3P 5

/* defaults above + optical->base localTransform + identity guess (stereoCamGeometricTools.cpp:126) */
SEPB_API void sepb_pnp_params_init(sepb_pnp_params* p, double fx, double fy, double cx, double cy);
/* set guess_rvec/guess_tvec from a 3x4 row-major transform `guess` (From-base <- To-base), like
 * estimateMotion3DTo2D: guessCameraFrame = (guess * localTransform)^-1 */
SEPB_API void sepb_pnp_set_guess(sepb_pnp_params* p, const double* guess3x4);

/* status per pair */
#define SEPB_PNP_TOO_FEW 0     /* fewer than min_inliers (or than a minimal set of) correspondences: null transform, no RANSAC */
#define SEPB_PNP_OK 1          /* transform valid (success=true in EstTransform.srv) */
#define SEPB_PNP_NO_MODEL 2    /* RANSAC found no hypothesis with more inliers than a minimal set has points */
#define SEPB_PNP_FEW_INLIERS 3 /* inliers < min_inliers (after RANSAC, or after bundle adjustment dropped its outliers) */
#define SEPB_PNP_BA_FAILED 4   /* bundle adjustment diverged (NaN / huge chi2): null transform, R/src/myRegistrationVis.cpp:1367-1370 */

/* Outputs (device or host according to the entry point):
 *   status[P] i32, n_inliers[P] i32, pose[P][6] f64 (rvec, tvec of the To camera w.r.t. From base),
 *   transform[P][12] f32 (3x4 row-major T, From-base <- To-base; zeros when not SEPB_PNP_OK, like
 *   MsgConversion.cpp:77-80), inlier_mask[P][mask_words] u32 over match-list positions (optional),
 *   best_hyp[P], iters_run[P] i32 (optional; index of the winning hypothesis, iterations executed).
 * custom_samples: optional i32 [P][custom_h][mp] minimal-sample indices (into the match list; mp = 5, or 4 with the
 *   P3P flags) replacing the built-in replay of cv::RNG(-1) -- "feed the reference's own indices". */
SEPB_API int sepb_pnp_ransac_dev(sepb_ctx* ctx, const sepb_frames* from, const sepb_frames* to, const int32_t* d_pairs,
                                 int n_pairs, const int32_t* d_n_matches, const uint16_t* d_match_from,
                                 const uint16_t* d_match_to, const int32_t* d_n_unique_to, const sepb_pnp_params* prm,
                                 const int32_t* d_custom_samples, int custom_h, const double* d_guess,
                                 int32_t* d_status, int32_t* d_n_inliers, double* d_pose, float* d_transform,
                                 uint32_t* d_inlier_mask, int mask_words, int32_t* d_best_hyp, int32_t* d_iters_run);

/* ---- fused verification: K2 -> K3 on the device, nothing but ids and results cross PCIe ----
 * The batch form of StereoCamGeometricTools::estimateTransformation pass 1
 * (R/src/stereoCamGeometricTools.cpp:122-152): pairs[p] = (from_frame, to_frame).
 * Host pointers; any output except status may be NULL. */
typedef struct sepb_verify_out {
  int32_t* status;       /* [P] */
  int32_t* n_matches;    /* [P] */
  int32_t* n_inliers;    /* [P] */
  double* pose;          /* [P][6] */
  float* transform;      /* [P][12] */
  uint16_t* match_from;  /* [P][max_feat] */
  uint16_t* match_to;    /* [P][max_feat] */
  uint32_t* inlier_mask; /* [P][(max_feat+31)/32] */
  int32_t* best_hyp;     /* [P] */
  int32_t* iters_run;    /* [P] */
  double* covariance;    /* [P][36] row-major 6x6 of info.covariance (needs inlier_mask): I6 with the translation block
                          * scaled by 2.1981 * lower-quartile squared 3-D residual and the rotation block by the same
                          * of the ray angles (RTAB-Map estimateMotion3DTo2D), diagonal clamped to >= 1e-9
                          * (R/src/myRegistration.cpp:279-295); I6 when there is no transform */
} sepb_verify_out;

SEPB_API int sepb_verify_pairs(sepb_ctx* ctx, const sepb_frames* from, const sepb_frames* to, const int32_t* pairs,
                               int n_pairs, const sepb_match_params* mprm, const sepb_pnp_params* pprm,
                               const int32_t* custom_samples, int custom_h, sepb_verify_out* out);
/* same call, returning as soon as the work (H2D of the pair list, K2, K3, D2H of the results) is enqueued: `pairs`
 * and the arrays inside `out` must be pinned host memory and stay untouched until sepb_ctx_sync(ctx) returns.  Lets
 * a caller start the next round's sepb_frames_upload_async behind it (double-buffered stores). */
SEPB_API int sepb_verify_pairs_async(sepb_ctx* ctx, const sepb_frames* from, const sepb_frames* to, const int32_t* pairs,
                                     int n_pairs, const sepb_match_params* mprm, const sepb_pnp_params* pprm,
                                     const int32_t* custom_samples, int custom_h, sepb_verify_out* out);
/* same, everything (pairs and the arrays inside `out`) in device memory, asynchronous on the ctx stream */
SEPB_API int sepb_verify_pairs_dev(sepb_ctx* ctx, const sepb_frames* from, const sepb_frames* to, const int32_t* d_pairs,
                                   int n_pairs, const sepb_match_params* mprm, const sepb_pnp_params* pprm,
                                   const int32_t* d_custom_samples, int custom_h, const sepb_verify_out* d_out);

/* ---- the full estimate_transformation service call, batched ----------------------------------------
 * StereoCamGeometricTools::estimateTransformation (R/src/stereoCamGeometricTools.cpp:122-178) calls
 * computeTransformationFromFeats TWICE: with the identity guess (:141-152 -> global matching, sepb_verify_pairs) and
 * again with the first result as guess (:153-164).  When the first result is valid the second call matches inside a
 * window around the projected guess (R/src/myRegistrationVis.cpp:477-825: projectPoints with (guess*local)^-1, To
 * keypoints within Vis/CorGuessWinSize px, equal octaves, 2-NN Hamming with the STRICT ratio test d1 < nndr*d2,
 * first-come ownership) and PnP-RANSAC is seeded with that guess; when it is null the second call repeats the first.
 * `out` then holds the SECOND call's result (what the service returns: success = !result2.isNull()); match lists and
 * inlier masks are those of the second call where it ran.  status_pass1 (optional, [P]) receives the first call's.
 * The reference's radius search is FLANN's approximate randomised kd-tree (:672-680, not reproducible); this is the
 * exact search it approximates. */
typedef struct sepb_guess_params {
  int32_t image_width, image_height;  /* cameraModel.imageSize() (image_size_ is 0x0 in the node, :469-472) */
  float win_size;                     /* Vis/CorGuessWinSize, pixels (20) */
} sepb_guess_params;
SEPB_API void sepb_guess_params_init(sepb_guess_params* g, int image_width, int image_height);
/* K2G alone (host buffers): the second call's correspondence lists for explicit first-call transforms
 * [P][12] f32 (3x4 row-major, From-base <- To-base); guess_out [P][6] f64 (optional) = rvec, tvec of
 * (guess * localTransform)^-1, the seed handed to PnP. */
SEPB_API int sepb_guess_match_pairs(sepb_ctx* ctx, const sepb_frames* from, const sepb_frames* to, const int32_t* pairs,
                                    int n_pairs, const float* transforms, const sepb_pnp_params* pprm,
                                    const sepb_guess_params* gprm, float nndr, int32_t* n_matches, uint16_t* match_from,
                                    uint16_t* match_to, double* guess_out);
SEPB_API int sepb_estimate_transformations(sepb_ctx* ctx, const sepb_frames* from, const sepb_frames* to,
                                           const int32_t* pairs, int n_pairs, const sepb_match_params* mprm,
                                           const sepb_pnp_params* pprm, const sepb_guess_params* gprm,
                                           sepb_verify_out* out, int32_t* status_pass1);
SEPB_API int sepb_estimate_transformations_dev(sepb_ctx* ctx, const sepb_frames* from, const sepb_frames* to,
                                               const int32_t* d_pairs, int n_pairs, const sepb_match_params* mprm,
                                               const sepb_pnp_params* pprm, const sepb_guess_params* gprm,
                                               const sepb_verify_out* d_out, int32_t* d_status_pass1);

/* ---- 2-view bundle adjustment after PnP (Vis/BundleAdjustment = 1) ---------------------------------
 * R/src/myRegistrationVis.cpp:1192-1370: when a pose was found, RTAB-Map's OptimizerG2O::optimizeBA refines it over
 * two cameras (From fixed, To free) and the inliers' 3-D points (free, starting at the From frame's points), with the
 * pixels of both frames as stereo (u, v, u - baseline * fx / depth) or mono edges under a Huber kernel and the PnP
 * result as a camera-to-camera edge weighted by covariance^-1; 5 + `iterations` Levenberg-Marquardt iterations, edges
 * whose chi2 exceeds the kernel's delta after each stage drop their inlier; the refined pose replaces the transform,
 * fewer than min_inliers remaining inliers (SEPB_PNP_FEW_INLIERS) or a diverged optimisation (SEPB_PNP_BA_FAILED) null
 * it.  g2o and RTAB-Map are not vendored by the reference: the algorithm is restated from their published sources
 * (oracle/ba.py), parity against the reference's binaries is unpinned.  cvsba (type 2) is not built.  Stores with
 * max_feat > 2048 are refused (SEPB_E_SHAPE). */
typedef struct sepb_ba_params {
  int32_t type;                /* Vis/BundleAdjustment: 0 = off, 1 = g2o-style (the only one built) */
  int32_t iterations;          /* Optimizer/Iterations of the second stage, 20 */
  double pixel_variance;       /* g2o/PixelVariance, 1.0 */
  double robust_kernel_delta;  /* g2o/RobustKernelDelta, 8.0; <= 0: no kernel, one stage of `iterations`, no outliers */
  double baseline;             /* what VertexCam::setKcam receives: stereo baseline [m] * fy / fx (:1247-1252, Tx = -baseline * fy);
                                * <= 0: every edge is a mono edge */
  int32_t variance_ignored;    /* Optimizer/VarianceIgnored: 1 = identity information on the camera-to-camera edge */
  int32_t reserved_;
} sepb_ba_params;
SEPB_API void sepb_ba_params_init(sepb_ba_params* b, double baseline);
/* consulted by sepb_verify_pairs* / sepb_estimate_transformations*: with type == 1 every PnP of those calls (both
 * passes of the two-call driver) is followed by its covariance and the bundle adjustment, as in the reference where
 * the block sits inside computeTransformationImpl.  The _dev forms then need inlier_mask and covariance in d_out.
 * NULL or type == 0 switches it off (the default). */
SEPB_API int sepb_ctx_set_bundle_adjustment(sepb_ctx* ctx, const sepb_ba_params* b);
/* the block alone, device buffers, asynchronous on the ctx stream.  In: pairs, match lists, covariance [P][36] (may be
 * NULL: identity), and the PnP results it updates IN PLACE for the pairs with status SEPB_PNP_OK: status, n_inliers,
 * pose, transform, inlier_mask.  d_stats (optional, [P][4] f64): iterations run, final robust chi2, inliers dropped,
 * final lambda.  Key points of both frames must be resident (host-bound stores: bind + verify first). */
SEPB_API int sepb_bundle_adjust_dev(sepb_ctx* ctx, const sepb_frames* from, const sepb_frames* to, const int32_t* d_pairs,
                                    int n_pairs, const int32_t* d_n_matches, const uint16_t* d_match_from,
                                    const uint16_t* d_match_to, const double* d_covariance, const sepb_pnp_params* pprm,
                                    const sepb_ba_params* bprm, int32_t* d_status, int32_t* d_n_inliers, double* d_pose,
                                    float* d_transform, uint32_t* d_inlier_mask, int mask_words, double* d_stats);
/* same on host buffers (synchronous) */
SEPB_API int sepb_bundle_adjust_pairs(sepb_ctx* ctx, const sepb_frames* from, const sepb_frames* to, const int32_t* pairs,
                                      int n_pairs, const int32_t* n_matches, const uint16_t* match_from,
                                      const uint16_t* match_to, const double* covariance, const sepb_pnp_params* pprm,
                                      const sepb_ba_params* bprm, int32_t* status, int32_t* n_inliers, double* pose,
                                      float* transform, uint32_t* inlier_mask, int mask_words, double* stats);

/* ---- multi-GPU: row-sharded NetVLAD database over NVLink / NVSwitch ---------------------------------
 * SURVEY.md 8e / BASELINE.json configs[2]: the peer database (received_descriptors, R/scripts/data_handler.py:300-301,
 * searched whole on every call, :168-170) is partitioned by rows over the GPUs of one box; queries are replicated,
 * every device searches its shard (sepb_db_set_base gives the shard's rows their GLOBAL indices), the per-shard
 * [nq][k] lists travel with ONE all-gather (ncclAllGather: 12 bytes per candidate and shard) and are merged on every
 * device in the (distance, index) order of the single-device search -- results identical to one database.
 * Verification pairs need no collective: they are independent units (sepb_verify_pairs on each device's share).
 *
 *   sepb_group_create(ndev, devices)         one process drives ndev devices: a context per device and ncclCommInitAll
 *                                            (devices == NULL: 0 .. ndev-1).  ndev == 1 needs no NCCL.
 *   sepb_group_unique_id + _create_rank      one process per device (e.g. under torchrun): rank 0 makes the id
 *                                            (ncclGetUniqueId), the launcher hands it to every rank, each wraps ITS
 *                                            context (ncclCommInitRank); the context stays owned by the caller.
 * NCCL is loaded at run time (libnccl.so.2); without it multi-device groups fail with SEPB_E_STATE. */
SEPB_API int sepb_group_create(int ndev, const int* devices, sepb_group** out);
SEPB_API int sepb_group_unique_id(uint8_t* id128);
SEPB_API int sepb_group_create_rank(sepb_ctx* ctx, const uint8_t* id128, int rank, int nranks, sepb_group** out);
SEPB_API void sepb_group_destroy(sepb_group* g);
SEPB_API int sepb_group_local_size(const sepb_group* g);   /* devices driven by this process */
SEPB_API int sepb_group_nranks(const sepb_group* g);       /* devices in the communicator = shards */
SEPB_API sepb_ctx* sepb_group_ctx(sepb_group* g, int i);   /* context of local device i (create databases / stores on it) */
SEPB_API const char* sepb_group_last_error(const sepb_group* g);
SEPB_API int sepb_group_nccl_version(void);                /* e.g. 22809; 0 = NCCL not available */
SEPB_API int64_t sepb_group_last_gather_bytes(const sepb_group* g);   /* bytes received by this process in the last all-gather */
/* sharded search, device form: dbs[i] = the shard on local device i (created on sepb_group_ctx(g, i));
 * d_queries[i] [nq][dim] f64, d_out_idx[i] [nq][k] i32, d_out_dist[i] [nq][k] f64 on that device.  Asynchronous: shard
 * search, all-gather and merge are enqueued on each context's stream.  Every process of the communicator must call it
 * with the same nq / k.  mode as sepb_netvlad_topk. */
SEPB_API int sepb_group_netvlad_topk_dev(sepb_group* g, sepb_db* const* dbs, const double* const* d_queries, int nq, int k,
                                         int mode, int32_t* const* d_out_idx, double* const* d_out_dist);
/* host form: queries [nq][dim] f64 are copied to local device 0 and broadcast over NVLink to the other local devices;
 * out_idx / out_dist receive the merged lists (identical on every device); returns when they are in place. */
SEPB_API int sepb_group_netvlad_topk(sepb_group* g, sepb_db* const* dbs, const double* queries, int nq, int k, int mode,
                                     int32_t* out_idx, double* out_dist);

/* ---- measurement support (roofline denominators not in MEASURED_PEAKS.json) ----------------
 * Microbenchmarks, best of `reps`.  kind 0: POPC instructions/s (K2's bounding pipe); 1: FP64 FMA/s (K3's);
 * 2 / 3: TMEM bytes/s read by tcgen05.ld (plain / .pack::16b); 4 / 5: int8 op/s of a bare tcgen05.mma kind::i8 issue
 * loop, M128 x N128 / N256 x K32, both operands in shared memory (what K2T's instruction shape sustains per GPU);
 * 6: bf16 FLOP/s of the same loop with kind::f16 M128 x N256 x K16; 7: kind 4 with producer-like shared-memory
 * stores running beside it; 8 / 9: the loop issued by CTA pairs (cta_group::2, M256 x N256), int8 K32 / bf16 K16. */
SEPB_API int sepb_measure_peak(sepb_ctx* ctx, int kind, int reps, double* ops_per_s);
/* developer hook (K3): cycle counters {EPnP solve, Rodrigues, scoring, refit, chunk-solves, 12x12 Jacobi, Jacobi
 * sweeps, 0} summed over warps; all zero and return 0 unless the library was built with SEPB_K3_PROFILE=1. */
SEPB_API int sepb_debug_k3_prof(unsigned long long* out8, int reset);

#ifdef __cplusplus
}
#endif
#endif /* SEPB200_H */
