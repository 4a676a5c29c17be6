#pragma once
#include "msgs.h"
namespace multi_robot_separators {
struct EstTransform {      // srv/EstTransform.srv
  struct Request {
    Descriptors descriptorsFrom, descriptorsTo;
    KeyPoint3DVec kptsFrom3D, kptsTo3D;
    KeyPointVec kptsFrom, kptsTo;
  } request;
  struct Response {
    geometry_msgs::PoseWithCovariance poseWithCov;
    uint8_t success = 0;
  } response;
};
}  // namespace multi_robot_separators
