// TEST-ONLY stand-ins for the generated headers of R/msg/*.msg (field for field) and rtabmap_ros/{KeyPoint,Point3f}.
#pragma once
#include <cstdint>
#include <vector>

#include <geometry_msgs/PoseWithCovariance.h>

namespace rtabmap_ros {
struct Point2f { float x = 0, y = 0; };
struct Point3f { float x = 0, y = 0, z = 0; };
struct KeyPoint {
  Point2f pt;
  float size = 0, angle = 0, response = 0;
  int32_t octave = 0, class_id = 0;
};
}  // namespace rtabmap_ros

namespace multi_robot_separators {
struct Descriptors {       // msg/Descriptors.msg
  uint16_t rows = 0, cols = 0;
  std::vector<uint8_t> data;
};
struct KeyPoint3DVec {     // msg/KeyPoint3DVec.msg
  int16_t size = 0;
  std::vector<rtabmap_ros::Point3f> kpts3DVec;
};
struct KeyPointVec {       // msg/KeyPointVec.msg
  int16_t size = 0;
  std::vector<rtabmap_ros::KeyPoint> kptsVec;
};
}  // namespace multi_robot_separators
