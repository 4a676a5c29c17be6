// TEST-ONLY stand-in for sensor_msgs/CameraInfo (see ros/ros.h in this directory).
#pragma once
#include <array>
#include <cstdint>
#include <string>

#include <ros/ros.h>

namespace sensor_msgs {
struct Header {
  ros::Time stamp;
  std::string frame_id;
};
struct CameraInfo {
  Header header;
  uint32_t height = 0, width = 0;
  std::array<double, 9> K{};
  std::array<double, 12> P{};
};
using CameraInfoConstPtr = boost::shared_ptr<CameraInfo const>;
}  // namespace sensor_msgs
