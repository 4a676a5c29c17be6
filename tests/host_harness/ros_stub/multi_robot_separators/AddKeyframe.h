#pragma once
#include "msgs.h"
namespace multi_robot_separators {
struct AddKeyframe {       // srv/AddKeyframe.srv -- the ONE new .srv the drop-in adds
  struct Request {
    std::vector<double> netvlad_descriptor;   // float64[]  first netvlad_dimensions values are kept
    int16_t kf_id = 0;                        // int16      id among all odometry keyframes (kf_ids_of_frames_kept)
    Descriptors descriptors;                  // the GetFeatsAndDesc.srv response of this keyframe
    KeyPoint3DVec kpts3D;
    KeyPointVec kpts;
  } request;
  struct Response {
    uint8_t success = 0;
    int16_t frames_kept_id = -1;              // index of the keyframe among the frames kept (-1 on failure)
  } response;
};
}  // namespace multi_robot_separators
