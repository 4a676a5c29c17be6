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
#define SEPB_E_STATE (-3)     /* handle used before it was filled */

#define SEPB_DESC_BYTES 32    /* 256-bit binary descriptors (ORB / BRIEF-32), R/msg/Descriptors.msg */
#define SEPB_MAX_FEATURES 4096

typedef struct sepb_ctx sepb_ctx;        /* one device + one stream + scratch */
typedef struct sepb_frames sepb_frames;  /* device-resident keyframe feature store of one robot */
typedef struct sepb_db sepb_db;          /* device-resident NetVLAD descriptor database */

/* ---- context -------------------------------------------------------------------------- */
SEPB_API int sepb_ctx_create(int device, sepb_ctx** out);
SEPB_API void sepb_ctx_destroy(sepb_ctx* ctx);
SEPB_API const char* sepb_last_error(const sepb_ctx* ctx);
SEPB_API void* sepb_ctx_stream(const sepb_ctx* ctx);           /* cudaStream_t */
SEPB_API int sepb_ctx_sync(sepb_ctx* ctx);
/* number of kernels this library launched on ctx since creation (bench.py "gpu_launches") */
SEPB_API int64_t sepb_ctx_launch_count(const sepb_ctx* ctx);

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

/* ---- measurement support (roofline denominators not in MEASURED_PEAKS.json) ----------------
 * kind 0: POPC instructions/s (K2's bounding pipe), kind 1: FP64 FMA/s (K3's).  Best of reps. */
SEPB_API int sepb_measure_peak(sepb_ctx* ctx, int kind, int reps, double* ops_per_s);

#ifdef __cplusplus
}
#endif
#endif /* SEPB200_H */
