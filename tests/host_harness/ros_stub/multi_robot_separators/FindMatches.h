#pragma once
#include "msgs.h"
namespace multi_robot_separators {
struct FindMatches {       // srv/FindMatches.srv
  struct Request {
    std::vector<double> new_netvlad_descriptors;
  } request;
  struct Response {
    std::vector<int16_t> kf_ids_computing_robot, frames_kept_ids_computing_robot, frames_kept_ids_querying_robot;
    std::vector<Descriptors> descriptors_vec;
    std::vector<KeyPoint3DVec> kpts3D_vec;
    std::vector<KeyPointVec> kpts_vec;
    std::vector<geometry_msgs::PoseWithCovariance> pose_estimates;
  } response;
};
}  // namespace multi_robot_separators
