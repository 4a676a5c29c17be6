// C++ host side above the C ABI (include/sepb200.h): the two services of the reference's hot path with the
// reference's own callback signatures `bool cb(Request&, Response&)`, written against the POD message
// mirrors of sepb_msgs.hpp.  A ROS build wraps them 1:1 (ros_node.cpp, needs roscpp); here they are
// driven by tests through c_shim.cpp.
//
//   GeometricTools::estimateTransformation    <- StereoCamGeometricTools::estimateTransformation
//                                                (R/src/stereoCamGeometricTools.cpp:122-178)
//   PlaceRecognition::findMatchesCompute      <- DataHandler.find_matches_service
//                                                (R/scripts/data_handler.py:297-337; moved to C++)
//   PlaceRecognition::receiveSeparators       <- DataHandler.receive_separators_service (:373-419)
//   PlaceRecognition::addKeyframe             <- the state DataHandler keeps per kept frame
//                                                (local_descriptors :160, geometric_feats :287, kf ids :287)
#pragma once
#include <cstdint>
#include <string>
#include <utility>
#include <vector>

#include "../include/sepb200.h"
#include "sepb_msgs.hpp"

namespace sepb_host {


struct CameraModel {                  // cam_.left(): rectified intrinsics + image size (stereoCamGeometricTools.cpp:56-77)
  double fx = 0, fy = 0, cx = 0, cy = 0;
  int width = 0, height = 0;
};

// Owns one sepb context.  Both service classes can share it (one GPU per robot process).
class Device {
 public:
  explicit Device(int device = 0);
  ~Device();
  Device(const Device&) = delete;
  Device& operator=(const Device&) = delete;
  sepb_ctx* ctx() const { return ctx_; }
  bool ok() const { return ctx_ != nullptr; }
  std::string lastError() const;

 private:
  sepb_ctx* ctx_ = nullptr;
};

class GeometricTools {
 public:
  // min_inliers = separators_min_inliers (launch default 5, stereoCamGeometricTools.cpp:87); the other Vis/*
  // parameters are RTAB-Map's compiled defaults, exposed for the versions that differ
  GeometricTools(Device& dev, const CameraModel& cam, int min_inliers = 5, int max_features = 2048);
  ~GeometricTools();
  void setNNDR(float v) { match_.nndr = v; }                 // Vis/CorNNDR
  void setGuessWinSize(float px) { guess_.win_size = px; }   // Vis/CorGuessWinSize
  void setIterations(int n) { pnp_.iterations = n; }         // Vis/Iterations
  void setReprojError(double px) { pnp_.reproj_error = px; } // Vis/PnPReprojError
  // cam_.localTransform(): pose of the (left, optical) camera in the robot base frame -- tf frame_id <- camera frame
  // (stereoCamGeometricTools.cpp:42-60).  Row-major 3x3 rotation and translation; also re-derives the identity guess.
  void setLocalTransform(const double* R9, const double* t3) {
    for (int i = 0; i < 9; ++i) pnp_.local_R[i] = R9[i];
    for (int i = 0; i < 3; ++i) pnp_.local_t[i] = t3[i];
    const double I[12] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0};
    sepb_pnp_set_guess(&pnp_, I);
  }
  void setPnPFlags(int flags) { pnp_.flags = flags; }        // Vis/PnPFlags: 0 ITERATIVE, 2 P3P, 5 AP3P (myRegistrationVis.cpp:103)
  void setPnPRefineIterations(int n) { pnp_.refine_iterations = n; }   // Vis/PnPRefineIterations (:104)
  void setOpenCV32Semantics(bool on) { pnp_.minimal_model = on ? 1 : 0; }   // return the minimal-set model (OpenCV 3.2)
  // Vis/BundleAdjustment (myRegistrationVis.cpp:113,1192-1370): 0 = off (default here), 1 = the g2o-style 2-view
  // adjustment after every PnP of the call; `stereo_baseline` [m] is StereoCameraModel::baseline() (right camera_info:
  // -P[3] / P[0]).  The setting lives on the shared device context.  False for an unsupported type (2 = cvsba).
  bool setBundleAdjustment(int type, double stereo_baseline, int iterations = 20, double pixel_variance = 1.0,
                           double robust_kernel_delta = 8.0);

  // Always returns true (the ROS convention: the response is valid); failure of the computation is in-band:
  // res.success = false with a zero pose (MsgConversion.cpp:77-80).  Never throws, never aborts: malformed
  // requests (the reference UASSERTs, myRegistrationVis.cpp:482-483) also answer success = false.
  bool estimateTransformation(const msgs::EstTransform::Request& req, msgs::EstTransform::Response& res);

  int lastInliers() const { return last_inliers_; }   // info_.inliers / info_.matches of the second call
  int lastMatches() const { return last_matches_; }

 private:
  bool upload(sepb_frames* f, const msgs::Descriptors& d, const msgs::KeyPointVec& k, const msgs::KeyPoint3DVec& k3);
  Device& dev_;
  sepb_frames* from_ = nullptr;
  sepb_frames* to_ = nullptr;
  int max_feat_;
  sepb_match_params match_{};
  sepb_pnp_params pnp_{};
  sepb_guess_params guess_{};
  int last_inliers_ = 0, last_matches_ = 0;
  std::vector<uint8_t> desc_;
  std::vector<float> kp2_, kp3_;
  std::vector<int32_t> oct_;
};

class PlaceRecognition {
 public:
  // netvlad_distance / dimensions / max_matches_nb: launch parameters (multi_robot_separators.launch:19-22)
  PlaceRecognition(Device& dev, int netvlad_dimensions = 128, double netvlad_distance = 0.13, int max_matches_nb = 20,
                   int capacity = 65536);
  ~PlaceRecognition();

  // one kept keyframe of the local robot: NetVLAD descriptor (first netvlad_dimensions values), features, kf id.
  // false (nothing stored, lastError() set) once the frames-kept index would no longer fit the int16 ids of
  // FindMatches.srv:3-5 / ReceiveSeparators.srv:3-6 (rospy refuses to serialise such a response).
  bool addKeyframe(const std::vector<double>& netvlad, const msgs::Descriptors& d, const msgs::KeyPoint3DVec& k3,
                   const msgs::KeyPointVec& k, int16_t kf_id);

  bool findMatchesCompute(const msgs::FindMatches::Request& req, msgs::FindMatches::Response& res);
  bool receiveSeparators(const msgs::ReceiveSeparators::Request& req, msgs::ReceiveSeparators::Response& res);

  // DataHandler.find_matches (data_handler.py:166-209): [(idx_local, idx_other)]
  std::vector<std::pair<int, int>> findMatches();
  int receivedCount() const;
  int localCount() const { return (int)kf_ids_of_frames_kept_.size(); }
  // matches left out of responses because an id exceeded the int16 wire types (in-band: the match is simply absent,
  // like "no match"); lastError() says why
  int droppedForInt16() const { return dropped_int16_; }
  const std::string& lastError() const { return err_; }

 private:
  Device& dev_;
  sepb_db* db_ = nullptr;
  int dim_, max_nb_, capacity_;
  double distance_;
  std::vector<double> local_descriptors_;            // [n_local][dim]
  std::vector<msgs::Descriptors> feats_desc_;        // geometric_feats
  std::vector<msgs::KeyPoint3DVec> feats_k3_;
  std::vector<msgs::KeyPointVec> feats_k_;
  std::vector<int16_t> kf_ids_of_frames_kept_;
  std::vector<int> local_kf_already_used_, other_kf_already_used_;
  std::vector<std::pair<int, int>> frames_kept_pairs_ignored_;
  int dropped_int16_ = 0;
  std::string err_;
};

// rtabmap::Transform (f32 3x4, row-major) -> geometry_msgs/Pose like tf::poseEigenToMsg(transform.toEigen3d())
void transformToPoseMsg(const float* T12, bool is_null, msgs::Pose& msg);

}  // namespace sepb_host
