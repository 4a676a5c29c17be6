// ROS1 wrapper of the service handlers (needs roscpp + the generated multi_robot_separators / rtabmap_ros message
// headers; not built in this repository's image, which has no ROS).  Same service names and .srv types as the reference:
//   estimate_transformation   (R/src/stereoCamGeometricTools.cpp:210)
//   find_matches_compute      (R/scripts/find_separators.py:31-32, served from C++ here)
//   receive_separators_py     (R/scripts/find_separators.py:33-34)
// Build: catkin, add this file + separator_services.cpp to the package and link libsepb200.so.
#ifdef SEPB_WITH_ROS
#include <ros/ros.h>

#include "multi_robot_separators/EstTransform.h"
#include "multi_robot_separators/FindMatches.h"
#include "multi_robot_separators/ReceiveSeparators.h"
#include "separator_services.hpp"

namespace mrs = multi_robot_separators;
namespace pod = sepb_host::msgs;

static void toPod(const mrs::Descriptors& m, pod::Descriptors& d) { d.rows = m.rows; d.cols = m.cols; d.data = m.data; }
static void toPod(const mrs::KeyPointVec& m, pod::KeyPointVec& k) {
  k.size = m.size;
  k.kptsVec.resize(m.kptsVec.size());
  for (size_t i = 0; i < m.kptsVec.size(); ++i) {
    k.kptsVec[i].pt.x = m.kptsVec[i].pt.x; k.kptsVec[i].pt.y = m.kptsVec[i].pt.y;
    k.kptsVec[i].size = m.kptsVec[i].size; k.kptsVec[i].angle = m.kptsVec[i].angle;
    k.kptsVec[i].response = m.kptsVec[i].response; k.kptsVec[i].octave = m.kptsVec[i].octave;
    k.kptsVec[i].class_id = m.kptsVec[i].class_id;
  }
}
static void toPod(const mrs::KeyPoint3DVec& m, pod::KeyPoint3DVec& k) {
  k.size = m.size;
  k.kpts3DVec.resize(m.kpts3DVec.size());
  for (size_t i = 0; i < m.kpts3DVec.size(); ++i) { k.kpts3DVec[i].x = m.kpts3DVec[i].x; k.kpts3DVec[i].y = m.kpts3DVec[i].y; k.kpts3DVec[i].z = m.kpts3DVec[i].z; }
}
static void fromPod(const pod::Descriptors& d, mrs::Descriptors& m) { m.rows = d.rows; m.cols = d.cols; m.data = d.data; }
static void fromPod(const pod::KeyPointVec& k, mrs::KeyPointVec& m) {
  m.size = k.size;
  m.kptsVec.resize(k.kptsVec.size());
  for (size_t i = 0; i < k.kptsVec.size(); ++i) {
    m.kptsVec[i].pt.x = k.kptsVec[i].pt.x; m.kptsVec[i].pt.y = k.kptsVec[i].pt.y; m.kptsVec[i].size = k.kptsVec[i].size;
    m.kptsVec[i].angle = k.kptsVec[i].angle; m.kptsVec[i].response = k.kptsVec[i].response;
    m.kptsVec[i].octave = k.kptsVec[i].octave; m.kptsVec[i].class_id = k.kptsVec[i].class_id;
  }
}
static void fromPod(const pod::KeyPoint3DVec& k, mrs::KeyPoint3DVec& m) {
  m.size = k.size;
  m.kpts3DVec.resize(k.kpts3DVec.size());
  for (size_t i = 0; i < k.kpts3DVec.size(); ++i) { m.kpts3DVec[i].x = k.kpts3DVec[i].x; m.kpts3DVec[i].y = k.kpts3DVec[i].y; m.kpts3DVec[i].z = k.kpts3DVec[i].z; }
}

int main(int argc, char** argv) {
  ros::init(argc, argv, "separators_b200");
  ros::NodeHandle n;
  sepb_host::CameraModel cam;   // filled from the camera_info / tf lookups of stereoCamGeometricTools.cpp:19-77
  n.getParam("fx", cam.fx); n.getParam("fy", cam.fy); n.getParam("cx", cam.cx); n.getParam("cy", cam.cy);
  n.getParam("image_width", cam.width); n.getParam("image_height", cam.height);
  int min_inliers = 5, dims = 128, max_nb = 20;
  double distance = 0.13;
  n.getParam("separators_min_inliers", min_inliers);
  n.getParam("netvlad_dimensions", dims);
  n.getParam("netvlad_distance", distance);
  n.getParam("netvlad_max_matches_nb", max_nb);
  sepb_host::Device dev(0);
  sepb_host::GeometricTools geo(dev, cam, min_inliers);
  sepb_host::PlaceRecognition place(dev, dims, distance, max_nb);

  boost::function<bool(mrs::EstTransform::Request&, mrs::EstTransform::Response&)> est =
      [&](mrs::EstTransform::Request& req, mrs::EstTransform::Response& res) {
        pod::EstTransform::Request r;
        pod::EstTransform::Response o;
        toPod(req.descriptorsFrom, r.descriptorsFrom); toPod(req.descriptorsTo, r.descriptorsTo);
        toPod(req.kptsFrom, r.kptsFrom); toPod(req.kptsTo, r.kptsTo);
        toPod(req.kptsFrom3D, r.kptsFrom3D); toPod(req.kptsTo3D, r.kptsTo3D);
        geo.estimateTransformation(r, o);
        res.success = o.success;
        res.poseWithCov.pose.position.x = o.poseWithCov.pose.position.x;
        res.poseWithCov.pose.position.y = o.poseWithCov.pose.position.y;
        res.poseWithCov.pose.position.z = o.poseWithCov.pose.position.z;
        res.poseWithCov.pose.orientation.x = o.poseWithCov.pose.orientation.x;
        res.poseWithCov.pose.orientation.y = o.poseWithCov.pose.orientation.y;
        res.poseWithCov.pose.orientation.z = o.poseWithCov.pose.orientation.z;
        res.poseWithCov.pose.orientation.w = o.poseWithCov.pose.orientation.w;
        for (int i = 0; i < 36; ++i) res.poseWithCov.covariance[i] = o.poseWithCov.covariance[i];
        return true;
      };
  boost::function<bool(mrs::FindMatches::Request&, mrs::FindMatches::Response&)> fm =
      [&](mrs::FindMatches::Request& req, mrs::FindMatches::Response& res) {
        pod::FindMatches::Request r;
        pod::FindMatches::Response o;
        r.new_netvlad_descriptors = req.new_netvlad_descriptors;
        place.findMatchesCompute(r, o);
        res.kf_ids_computing_robot = o.kf_ids_computing_robot;
        res.frames_kept_ids_computing_robot = o.frames_kept_ids_computing_robot;
        res.frames_kept_ids_querying_robot = o.frames_kept_ids_querying_robot;
        res.descriptors_vec.resize(o.descriptors_vec.size());
        res.kpts3D_vec.resize(o.kpts3D_vec.size());
        res.kpts_vec.resize(o.kpts_vec.size());
        for (size_t i = 0; i < o.descriptors_vec.size(); ++i) {
          fromPod(o.descriptors_vec[i], res.descriptors_vec[i]);
          fromPod(o.kpts3D_vec[i], res.kpts3D_vec[i]);
          fromPod(o.kpts_vec[i], res.kpts_vec[i]);
        }
        return true;
      };
  boost::function<bool(mrs::ReceiveSeparators::Request&, mrs::ReceiveSeparators::Response&)> rs =
      [&](mrs::ReceiveSeparators::Request& req, mrs::ReceiveSeparators::Response& res) {
        pod::ReceiveSeparators::Request r;
        pod::ReceiveSeparators::Response o;
        r.kf_ids_from = req.kf_ids_from; r.kf_ids_to = req.kf_ids_to;
        r.frames_kepts_ids_from = req.frames_kepts_ids_from; r.frames_kepts_ids_to = req.frames_kepts_ids_to;
        r.transform_est_success.assign(req.transform_est_success.begin(), req.transform_est_success.end());
        place.receiveSeparators(r, o);
        res.success = o.success;
        return true;
      };
  ros::ServiceServer s1 = n.advertiseService("estimate_transformation", est);
  ros::ServiceServer s2 = n.advertiseService("find_matches_compute", fm);
  ros::ServiceServer s3 = n.advertiseService("receive_separators_py", rs);
  ros::spin();   // single-threaded like the reference node (stereoCamGeometricTools.cpp:212)
  return 0;
}
#endif  // SEPB_WITH_ROS
