#pragma once
#include "msgs.h"
namespace multi_robot_separators {
struct ReceiveSeparators { // srv/ReceiveSeparators.srv
  struct Request {
    int8_t robot_from_id = 0, robot_to_id = 0;
    std::vector<int16_t> kf_ids_from, kf_ids_to, frames_kepts_ids_from, frames_kepts_ids_to;
    std::vector<geometry_msgs::PoseWithCovariance> pose_estimates_from, pose_estimates_to;
    std::vector<uint8_t> transform_est_success;
    std::vector<geometry_msgs::PoseWithCovariance> separators;
  } request;
  struct Response {
    uint8_t success = 0;
  } response;
};
}  // namespace multi_robot_separators
