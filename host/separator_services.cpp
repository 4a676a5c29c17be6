#include "separator_services.hpp"

#include <algorithm>
#include <cmath>
#include <cstring>

namespace sepb_host {

// ---- Device --------------------------------------------------------------------------------------
Device::Device(int device) {
  if (sepb_ctx_create(device, &ctx_) != SEPB_OK) ctx_ = nullptr;   // no CPU fallback: the services then answer failure
}
Device::~Device() { sepb_ctx_destroy(ctx_); }
std::string Device::lastError() const { return ctx_ ? sepb_last_error(ctx_) : "no CUDA context (sm_100 device required)"; }

// ---- message helpers -------------------------------------------------------------------------------
// Eigen::Quaterniond(rotation) (trace-based) + the w < 0 flip of tf::poseEigenToMsg
void transformToPoseMsg(const float* T, bool is_null, msgs::Pose& msg) {
  msg = msgs::Pose();
  if (is_null) return;   // MsgConversion.cpp:77-80: zero pose for a null transform
  const double m00 = T[0], m01 = T[1], m02 = T[2], m10 = T[4], m11 = T[5], m12 = T[6], m20 = T[8], m21 = T[9], m22 = T[10];
  double x, y, z, w;
  double t = m00 + m11 + m22;
  if (t > 0) {
    t = std::sqrt(t + 1.0);
    w = 0.5 * t;
    t = 0.5 / t;
    x = (m21 - m12) * t;
    y = (m02 - m20) * t;
    z = (m10 - m01) * t;
  } else {
    const double m[3][3] = {{m00, m01, m02}, {m10, m11, m12}, {m20, m21, m22}};
    int i = 0;
    if (m11 > m00) i = 1;
    if (m22 > m[i][i]) i = 2;
    const int j = (i + 1) % 3, k = (j + 1) % 3;
    t = std::sqrt(m[i][i] - m[j][j] - m[k][k] + 1.0);
    double q[3];
    q[i] = 0.5 * t;
    t = 0.5 / t;
    w = (m[k][j] - m[j][k]) * t;
    q[j] = (m[j][i] + m[i][j]) * t;
    q[k] = (m[k][i] + m[i][k]) * t;
    x = q[0]; y = q[1]; z = q[2];
  }
  if (w < 0) { x = -x; y = -y; z = -z; w = -w; }
  msg.position.x = T[3];
  msg.position.y = T[7];
  msg.position.z = T[11];
  msg.orientation.x = x; msg.orientation.y = y; msg.orientation.z = z; msg.orientation.w = w;
}

// ---- GeometricTools --------------------------------------------------------------------------------
GeometricTools::GeometricTools(Device& dev, const CameraModel& cam, int min_inliers, int max_features)
    : dev_(dev), max_feat_(max_features) {
  sepb_pnp_params_init(&pnp_, cam.fx, cam.fy, cam.cx, cam.cy);
  pnp_.min_inliers = std::max(min_inliers, 4);   // util3d::solvePnPRansac: max(minInliers, 4); four points take the P3P solve
  sepb_guess_params_init(&guess_, cam.width, cam.height);
  match_.nndr = 0.8f;
  match_.nndr_strict = 0;
  if (dev_.ok()) {
    if (sepb_frames_create(dev_.ctx(), 1, max_feat_, &from_) != SEPB_OK) from_ = nullptr;
    if (sepb_frames_create(dev_.ctx(), 1, max_feat_, &to_) != SEPB_OK) to_ = nullptr;
  }
  desc_.resize((size_t)max_feat_ * 32);
  kp2_.resize((size_t)max_feat_ * 2);
  kp3_.resize((size_t)max_feat_ * 3);
  oct_.resize(max_feat_);
}

bool GeometricTools::setBundleAdjustment(int type, double stereo_baseline, int iterations, double pixel_variance,
                                         double robust_kernel_delta) {
  if (!dev_.ok()) return false;
  if (type == 0) return sepb_ctx_set_bundle_adjustment(dev_.ctx(), nullptr) == SEPB_OK;
  sepb_ba_params b;
  // myRegistrationVis.cpp:1247-1252: Tx = -baseline * fy, and VertexCam::setKcam gets -Tx / fx
  sepb_ba_params_init(&b, stereo_baseline * pnp_.fy / pnp_.fx);
  b.type = type;
  b.iterations = iterations;
  b.pixel_variance = pixel_variance;
  b.robust_kernel_delta = robust_kernel_delta;
  return sepb_ctx_set_bundle_adjustment(dev_.ctx(), &b) == SEPB_OK;
}

GeometricTools::~GeometricTools() {
  sepb_frames_destroy(from_);
  sepb_frames_destroy(to_);
}

// descriptorsFromROS / keypointsFromROS / keypoints3DFromROS (MsgConversion.cpp:8-18,97-100,113-116) into one store slot
bool GeometricTools::upload(sepb_frames* f, const msgs::Descriptors& d, const msgs::KeyPointVec& k,
                            const msgs::KeyPoint3DVec& k3) {
  const size_t n = d.rows;
  if (d.cols != SEPB_DESC_BYTES && n > 0) return false;         // binary 256-bit descriptors only
  if (n > (size_t)max_feat_ || d.data.size() != n * d.cols) return false;
  if (k.kptsVec.size() != n || k3.kpts3DVec.size() != n) return false;   // the reference UASSERTs (:482-483)
  std::fill(desc_.begin(), desc_.end(), 0);
  if (n) std::memcpy(desc_.data(), d.data.data(), n * 32);
  for (size_t i = 0; i < n; ++i) {
    kp2_[2 * i] = k.kptsVec[i].pt.x;
    kp2_[2 * i + 1] = k.kptsVec[i].pt.y;
    oct_[i] = k.kptsVec[i].octave;
    kp3_[3 * i] = k3.kpts3DVec[i].x;
    kp3_[3 * i + 1] = k3.kpts3DVec[i].y;
    kp3_[3 * i + 2] = k3.kpts3DVec[i].z;
  }
  const int32_t count = (int32_t)n;
  return sepb_frames_upload(dev_.ctx(), f, 0, 1, desc_.data(), kp2_.data(), kp3_.data(), oct_.data(), &count) == SEPB_OK;
}

bool GeometricTools::estimateTransformation(const msgs::EstTransform::Request& req, msgs::EstTransform::Response& res) {
  res = msgs::EstTransform::Response();
  for (int i = 0; i < 6; ++i) res.poseWithCov.covariance[7 * i] = 1.0;   // info.covariance defaults to I6
  last_inliers_ = last_matches_ = 0;
  if (!dev_.ok() || !from_ || !to_) return true;
  if (!upload(from_, req.descriptorsFrom, req.kptsFrom, req.kptsFrom3D) ||
      !upload(to_, req.descriptorsTo, req.kptsTo, req.kptsTo3D))
    return true;
  // both calls of :141-164 in one batched launch sequence (P = 1)
  const int32_t pair[2] = {0, 0};
  int32_t status = 0, n_matches = 0, n_inliers = 0;
  double pose[6];
  float T[12];
  std::vector<uint16_t> mf(max_feat_), mt(max_feat_);
  std::vector<uint32_t> mask((max_feat_ + 31) / 32);
  sepb_verify_out out{};
  out.status = &status;
  out.n_matches = &n_matches;
  out.n_inliers = &n_inliers;
  out.pose = pose;
  out.transform = T;
  out.match_from = mf.data();
  out.match_to = mt.data();
  out.inlier_mask = mask.data();
  out.covariance = res.poseWithCov.covariance;
  if (sepb_estimate_transformations(dev_.ctx(), from_, to_, pair, 1, &match_, &pnp_, &guess_, &out, nullptr) != SEPB_OK) {
    for (int i = 0; i < 36; ++i) res.poseWithCov.covariance[i] = (i % 7 == 0) ? 1.0 : 0.0;
    return true;
  }
  last_inliers_ = n_inliers;
  last_matches_ = n_matches;
  res.success = status == SEPB_PNP_OK;                             // !result2.isNull() (:168-175)
  transformToPoseMsg(T, !res.success, res.poseWithCov.pose);
  return true;
}

// ---- PlaceRecognition ------------------------------------------------------------------------------
PlaceRecognition::PlaceRecognition(Device& dev, int dim, double distance, int max_nb, int capacity)
    : dev_(dev), dim_(dim), max_nb_(max_nb), capacity_(capacity), distance_(distance) {
  if (dev_.ok() && sepb_db_create(dev_.ctx(), dim_, capacity_, &db_) != SEPB_OK) db_ = nullptr;
}
PlaceRecognition::~PlaceRecognition() { sepb_db_destroy(db_); }
int PlaceRecognition::receivedCount() const { return sepb_db_size(db_); }

static constexpr int kInt16Max = 32767;   // int16[] ids of FindMatches.srv:3-5, ReceiveSeparators.srv:3-6

bool PlaceRecognition::addKeyframe(const std::vector<double>& netvlad, const msgs::Descriptors& d,
                                   const msgs::KeyPoint3DVec& k3, const msgs::KeyPointVec& k, int16_t kf_id) {
  if (localCount() > kInt16Max) {
    err_ = "addKeyframe: frames-kept index exceeds the int16 ids of FindMatches.srv";
    return false;
  }
  // data_handler.py:157-158: the first netvlad_dimensions values, no re-normalisation
  for (int i = 0; i < dim_; ++i) local_descriptors_.push_back(i < (int)netvlad.size() ? netvlad[i] : 0.0);
  feats_desc_.push_back(d);
  feats_k3_.push_back(k3);
  feats_k_.push_back(k);
  kf_ids_of_frames_kept_.push_back(kf_id);
  return true;
}

std::vector<std::pair<int, int>> PlaceRecognition::findMatches() {
  std::vector<std::pair<int, int>> matches;
  const int nq = (int)(local_descriptors_.size() / dim_), n = receivedCount();
  if (!db_ || nq == 0 || n == 0) return matches;
  // masks of :178-184
  std::vector<uint8_t> row_mask(nq, 0), col_mask(n, 0);
  for (int r : local_kf_already_used_)
    if (r >= 0 && r < nq) row_mask[r] = 1;
  for (int c : other_kf_already_used_)
    if (c >= 0 && c < n) col_mask[c] = 1;
  std::vector<int32_t> ign;
  for (auto& p : frames_kept_pairs_ignored_)
    if (p.first >= 0 && p.first < nq && p.second >= 0 && p.second < n) {
      ign.push_back(p.first);
      ign.push_back(p.second);
    }
  if (sepb_db_set_col_mask(dev_.ctx(), db_, other_kf_already_used_.empty() ? nullptr : col_mask.data()) != SEPB_OK)
    return matches;
  std::vector<int32_t> idx(nq);
  std::vector<double> dist(nq);
  if (sepb_netvlad_topk(dev_.ctx(), db_, local_descriptors_.data(), nq, 1, SEPB_NN_TENSOR,
                        local_kf_already_used_.empty() ? nullptr : row_mask.data(), ign.empty() ? nullptr : ign.data(),
                        (int)(ign.size() / 2), idx.data(), dist.data()) != SEPB_OK)
    return matches;
  std::vector<int32_t> out(2 * (size_t)std::max(max_nb_, 1));
  const int m = sepb_select_matches(idx.data(), dist.data(), nq, distance_, max_nb_, out.data());   // :191-205
  for (int i = 0; i < m; ++i) matches.emplace_back(out[2 * i], out[2 * i + 1]);
  return matches;
}

bool PlaceRecognition::findMatchesCompute(const msgs::FindMatches::Request& req, msgs::FindMatches::Response& res) {
  res = msgs::FindMatches::Response();
  if (!db_) return true;
  // received_descriptors.extend(reshape(-1, netvlad_dimensions)) (:300-301)
  const int n_new = (int)(req.new_netvlad_descriptors.size() / dim_);
  if (n_new > 0 && sepb_db_append(dev_.ctx(), db_, req.new_netvlad_descriptors.data(), n_new) != SEPB_OK) return true;
  if (receivedCount() == 0 || local_descriptors_.empty()) return true;   // empty response (:311)
  const auto matches = findMatches();
  for (auto& m : matches) {
    const int local = m.first;
    if (local < 0 || local >= (int)feats_desc_.size()) continue;        // `if not resp_feats_and_descs: continue` (:319-320)
    if (local > kInt16Max || m.second > kInt16Max) {   // not representable on the wire: absent from the response
      ++dropped_int16_;
      err_ = "find_matches_compute: match (" + std::to_string(local) + ", " + std::to_string(m.second) +
             ") dropped, id exceeds int16 (FindMatches.srv:3-5)";
      continue;
    }
    res.frames_kept_ids_computing_robot.push_back((int16_t)local);
    res.frames_kept_ids_querying_robot.push_back((int16_t)m.second);
    res.descriptors_vec.push_back(feats_desc_[local]);
    res.kpts3D_vec.push_back(feats_k3_[local]);
    res.kpts_vec.push_back(feats_k_[local]);
    res.kf_ids_computing_robot.push_back(kf_ids_of_frames_kept_[local]);   // get_kf_ids_from_frames_kept_ids (:327-328)
  }
  return true;   // pose_estimates stay empty: send_estimates_of_poses needs the external get_pose_estimates service
}

bool PlaceRecognition::receiveSeparators(const msgs::ReceiveSeparators::Request& req, msgs::ReceiveSeparators::Response& res) {
  const size_t n = req.kf_ids_from.size();
  for (size_t i = 0; i < n; ++i) {
    const bool ok = i < req.transform_est_success.size() && req.transform_est_success[i];
    const int to = i < req.frames_kepts_ids_to.size() ? req.frames_kepts_ids_to[i] : -1;
    const int from = i < req.frames_kepts_ids_from.size() ? req.frames_kepts_ids_from[i] : -1;
    if (ok) {
      local_kf_already_used_.push_back(to);       // :402-403
      other_kf_already_used_.push_back(from);     // :404-405
    } else {
      frames_kept_pairs_ignored_.emplace_back(to, from);   // :407-408
    }
  }
  res.success = true;
  return true;
}

}  // namespace sepb_host
