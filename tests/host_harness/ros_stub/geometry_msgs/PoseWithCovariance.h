// TEST-ONLY stand-in for geometry_msgs/PoseWithCovariance (see ros/ros.h in this directory).
#pragma once
#include <array>

namespace geometry_msgs {
struct Point { double x = 0, y = 0, z = 0; };
struct Quaternion { double x = 0, y = 0, z = 0, w = 0; };
struct Pose { Point position; Quaternion orientation; };
struct PoseWithCovariance {
  Pose pose;
  std::array<double, 36> covariance{};
};
}  // namespace geometry_msgs
