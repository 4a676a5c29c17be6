// POD mirrors of the ROS message / service types of multi_robot_separators, field for field, so that
// the service handlers in separator_services.hpp read like the reference's and a ROS build only has to
// copy fields (ros_node.cpp).  Sources (paths under ros_ws/src/multi_robot_separators/):
//   msg/Descriptors.msg:1-3, msg/KeyPointVec.msg:1-2, msg/KeyPoint3DVec.msg:1-2,
//   srv/EstTransform.srv:1-9, srv/FindMatches.srv:1-9, srv/ReceiveSeparators.srv:1-12,
//   rtabmap_ros/KeyPoint.msg, rtabmap_ros/Point3f.msg, geometry_msgs/PoseWithCovariance.
// The int16 ids / sizes of the wire format are kept (ids above 32767 need the batch C ABI, which is int32).
#pragma once
#include <cstdint>
#include <vector>

// (their own namespace: a ROS build also sees the generated multi_robot_separators:: types, ros_node.cpp copies between the two)
namespace sepb_host {
namespace msgs {

struct Descriptors {             // msg/Descriptors.msg
  uint16_t rows = 0;
  uint16_t cols = 0;
  std::vector<uint8_t> data;     // rows x cols, row-major, CV_8U (MsgConversion.cpp:113-116 aliases it)
};

struct Point2f { float x = 0, y = 0; };
struct Point3f { float x = 0, y = 0, z = 0; };   // rtabmap_ros/Point3f

struct KeyPoint {                // rtabmap_ros/KeyPoint == cv::KeyPoint (MsgConversion.cpp:34-42)
  Point2f pt;
  float size = 0, angle = 0, response = 0;
  int32_t octave = 0, class_id = 0;
};

struct KeyPointVec {             // msg/KeyPointVec.msg
  int16_t size = 0;
  std::vector<KeyPoint> kptsVec;
};

struct KeyPoint3DVec {           // msg/KeyPoint3DVec.msg
  int16_t size = 0;
  std::vector<Point3f> kpts3DVec;
};

struct Point { double x = 0, y = 0, z = 0; };
struct Quaternion { double x = 0, y = 0, z = 0, w = 0; };
struct Pose { Point position; Quaternion orientation; };
struct PoseWithCovariance {      // geometry_msgs/PoseWithCovariance
  Pose pose;
  double covariance[36] = {0};
};

struct EstTransform {            // srv/EstTransform.srv
  struct Request {
    Descriptors descriptorsFrom, descriptorsTo;
    KeyPoint3DVec kptsFrom3D, kptsTo3D;
    KeyPointVec kptsFrom, kptsTo;
  };
  struct Response {
    PoseWithCovariance poseWithCov;
    bool success = false;
  };
};

struct FindMatches {             // srv/FindMatches.srv
  struct Request {
    std::vector<double> new_netvlad_descriptors;
  };
  struct Response {
    std::vector<int16_t> kf_ids_computing_robot;
    std::vector<int16_t> frames_kept_ids_computing_robot;
    std::vector<int16_t> frames_kept_ids_querying_robot;
    std::vector<Descriptors> descriptors_vec;
    std::vector<KeyPoint3DVec> kpts3D_vec;
    std::vector<KeyPointVec> kpts_vec;
    std::vector<PoseWithCovariance> pose_estimates;
  };
};

struct ReceiveSeparators {       // srv/ReceiveSeparators.srv
  struct Request {
    int8_t robot_from_id = 0, robot_to_id = 0;
    std::vector<int16_t> kf_ids_from, kf_ids_to, frames_kepts_ids_from, frames_kepts_ids_to;
    std::vector<PoseWithCovariance> pose_estimates_from, pose_estimates_to;
    std::vector<uint8_t> transform_est_success;   // bool[]
    std::vector<PoseWithCovariance> separators;
  };
  struct Response {
    bool success = false;
  };
};

}  // namespace msgs
}  // namespace sepb_host
