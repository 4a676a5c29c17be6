// ROS1 node of the B200 separator path (needs roscpp, tf and the generated multi_robot_separators / rtabmap_ros message
// headers; this repository's image has no ROS, so the file is compiled here only against the stub headers of
// tests/host_harness/ros_stub -- tests/test_host_services.py::test_ros_node_compiles).  Same service names and .srv
// types as the reference:
//   estimate_transformation   (R/src/stereoCamGeometricTools.cpp:210)
//   find_matches_compute      (R/scripts/find_separators.py:31-32, served from C++ here)
//   receive_separators_py     (R/scripts/find_separators.py:33-34), forwarding the kept separators to
//                             add_separators_pose_graph like R/scripts/data_handler.py:410-415
// plus ONE new service, add_keyframe (srv/AddKeyframe.srv): the keyframe feed.  The reference fills
// local_descriptors / geometric_feats / kf_ids_of_frames_kept inside the Python process that also answers
// find_matches_compute (data_handler.py:143-164, :287); with the search served from C++, the Python front-end (NetVLAD
// CNN + get_features_and_descriptor, both out of this path) hands each kept keyframe over through this call instead.
// Camera intrinsics and the camera's localTransform come from camera_info + tf exactly as the reference node reads them
// (R/src/stereoCamGeometricTools.cpp:38-77, main :180-215).
// Build: catkin, add this file + separator_services.cpp to the package and link libsepb200.so.
#ifdef SEPB_WITH_ROS
#include <ros/ros.h>
#include <sensor_msgs/CameraInfo.h>
#include <tf/transform_listener.h>

#include "multi_robot_separators/AddKeyframe.h"

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

static void toRos(const pod::PoseWithCovariance& p, geometry_msgs::PoseWithCovariance& m) {
  m.pose.position.x = p.pose.position.x; m.pose.position.y = p.pose.position.y; m.pose.position.z = p.pose.position.z;
  m.pose.orientation.x = p.pose.orientation.x; m.pose.orientation.y = p.pose.orientation.y;
  m.pose.orientation.z = p.pose.orientation.z; m.pose.orientation.w = p.pose.orientation.w;
  for (int i = 0; i < 36; ++i) m.covariance[i] = p.covariance[i];
}

int main(int argc, char** argv) {
  ros::init(argc, argv, "separators_b200");
  ros::NodeHandle n;
  // ---- parameters of the reference launch file (R/launch/multi_robot_separators.launch:19-23,52-53)
  std::string frame_id;
  int min_inliers = 5, dims = 128, max_nb = 20, device = 0;
  int pnp_flags = 0, pnp_refine_iterations = 0;
  bool opencv32_semantics = false;
  int bundle_adjustment = 0;   // RTAB-Map's default is 1 when it was built with g2o; opt in here (parity unpinned)
  double distance = 0.13;
  if (!n.getParam("frame_id", frame_id)) ROS_ERROR("Couldn't find frame ID");
  if (!n.getParam("separators_min_inliers", min_inliers)) ROS_ERROR("Couldn't find separators_min_inliers param");
  n.getParam("netvlad_dimensions", dims);
  n.getParam("netvlad_distance", distance);
  n.getParam("netvlad_max_matches_nb", max_nb);
  n.getParam("sepb_device", device);
  n.getParam("Vis/PnPFlags", pnp_flags);                         // R/src/myRegistrationVis.cpp:103
  n.getParam("Vis/PnPRefineIterations", pnp_refine_iterations);  // :104
  n.getParam("sepb_opencv32_semantics", opencv32_semantics);
  n.getParam("Vis/BundleAdjustment", bundle_adjustment);        // :113

  // ---- camera model: left camera_info (rectified: P holds fx, fy, cx, cy) + tf frame_id <- camera frame, as
  //      stereoCameraModelFromROS builds cam_.left() and its localTransform (stereoCamGeometricTools.cpp:38-77)
  sensor_msgs::CameraInfoConstPtr info_l = ros::topic::waitForMessage<sensor_msgs::CameraInfo>("left/camera_info", n);
  sensor_msgs::CameraInfoConstPtr info_r = ros::topic::waitForMessage<sensor_msgs::CameraInfo>("right/camera_info", n);
  if (!info_l || !info_r) {
    ROS_ERROR("no camera_info");
    return 1;
  }
  sepb_host::CameraModel cam;
  cam.fx = info_l->P[0]; cam.fy = info_l->P[5]; cam.cx = info_l->P[2]; cam.cy = info_l->P[6];
  cam.width = (int)info_l->width; cam.height = (int)info_l->height;
  double local_R[9] = {0, 0, 1, -1, 0, 0, 0, -1, 0}, local_t[3] = {0, 0, 0};   // optical -> base if tf has nothing
  {
    tf::TransformListener listener;
    tf::StampedTransform local_tf;
    const ros::Time stamp = info_l->header.stamp > info_r->header.stamp ? info_l->header.stamp : info_r->header.stamp;
    try {
      listener.waitForTransform(frame_id, info_l->header.frame_id, stamp, ros::Duration(3.0));
      listener.lookupTransform(frame_id, info_l->header.frame_id, stamp, local_tf);
      for (int r = 0; r < 3; ++r) {
        for (int c = 0; c < 3; ++c) local_R[3 * r + c] = local_tf.getBasis()[r][c];
        local_t[r] = local_tf.getOrigin()[r];
      }
    } catch (tf::TransformException& ex) {
      ROS_ERROR("%s", ex.what());
      ros::Duration(1.0).sleep();
    }
  }

  sepb_host::Device dev(device);
  if (!dev.ok()) ROS_ERROR("separators_b200: no sm_100 device (%s); every request will answer success = false", dev.lastError().c_str());
  sepb_host::GeometricTools geo(dev, cam, min_inliers);
  geo.setLocalTransform(local_R, local_t);
  geo.setPnPFlags(pnp_flags);
  geo.setPnPRefineIterations(pnp_refine_iterations);
  geo.setOpenCV32Semantics(opencv32_semantics);
  if (bundle_adjustment != 0) {
    // StereoCameraModel::baseline() = -right.Tx / right.fx (right camera_info: P[3] = Tx)
    const double baseline = info_r->P[0] != 0 ? -info_r->P[3] / info_r->P[0] : 0.0;
    if (!geo.setBundleAdjustment(bundle_adjustment, baseline))
      ROS_ERROR("separators_b200: Vis/BundleAdjustment = %d is not available (1 = g2o-style only); running without", bundle_adjustment);
  }
  sepb_host::PlaceRecognition place(dev, dims, distance, max_nb);
  ros::ServiceClient add_seps = n.serviceClient<mrs::ReceiveSeparators>("add_separators_pose_graph");

  boost::function<bool(mrs::EstTransform::Request&, mrs::EstTransform::Response&)> est =
      [&](mrs::EstTransform::Request& req, mrs::EstTransform::Response& res) {
        pod::EstTransform::Request r;
        pod::EstTransform::Response o;
        toPod(req.descriptorsFrom, r.descriptorsFrom); toPod(req.descriptorsTo, r.descriptorsTo);
        toPod(req.kptsFrom, r.kptsFrom); toPod(req.kptsTo, r.kptsTo);
        toPod(req.kptsFrom3D, r.kptsFrom3D); toPod(req.kptsTo3D, r.kptsTo3D);
        geo.estimateTransformation(r, o);
        res.success = o.success;
        toRos(o.poseWithCov, res.poseWithCov);
        return true;
      };
  // keyframe feed: what data_handler.py appends per kept keyframe (:143-164 descriptors, :287 geometric_feats and
  // kf_ids_of_frames_kept)
  boost::function<bool(mrs::AddKeyframe::Request&, mrs::AddKeyframe::Response&)> addkf =
      [&](mrs::AddKeyframe::Request& req, mrs::AddKeyframe::Response& res) {
        pod::Descriptors d;
        pod::KeyPoint3DVec k3;
        pod::KeyPointVec k;
        toPod(req.descriptors, d); toPod(req.kpts3D, k3); toPod(req.kpts, k);
        res.success = place.addKeyframe(req.netvlad_descriptor, d, k3, k, req.kf_id);
        res.frames_kept_id = res.success ? (int16_t)(place.localCount() - 1) : (int16_t)-1;
        if (!res.success) ROS_ERROR("add_keyframe: %s", place.lastError().c_str());
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
        // pose_estimates stay empty here: they come from the back-end's get_pose_estimates service, which the caller
        // (find_separators.py) queries itself when send_estimates_of_poses is set (data_handler.py:322-335)
        if (place.droppedForInt16()) ROS_ERROR("find_matches_compute: %s", place.lastError().c_str());
        return true;
      };
  boost::function<bool(mrs::ReceiveSeparators::Request&, mrs::ReceiveSeparators::Response&)> rs =
      [&](mrs::ReceiveSeparators::Request& req, mrs::ReceiveSeparators::Response& res) {
        pod::ReceiveSeparators::Request r;
        pod::ReceiveSeparators::Response o;
        r.robot_from_id = req.robot_from_id; r.robot_to_id = req.robot_to_id;
        r.kf_ids_from = req.kf_ids_from; r.kf_ids_to = req.kf_ids_to;
        r.frames_kepts_ids_from = req.frames_kepts_ids_from; r.frames_kepts_ids_to = req.frames_kepts_ids_to;
        r.transform_est_success.assign(req.transform_est_success.begin(), req.transform_est_success.end());
        place.receiveSeparators(r, o);      // masks: local / other frames already used, pairs to ignore
        // ---- data_handler.py:373-415: the separators whose estimation succeeded go on to the pose graph
        mrs::ReceiveSeparators fwd;
        fwd.request.robot_from_id = req.robot_from_id;
        fwd.request.robot_to_id = req.robot_to_id;
        const bool with_estimates = req.pose_estimates_from.size() == req.kf_ids_from.size() &&
                                    req.pose_estimates_to.size() == req.kf_ids_from.size();
        for (size_t i = 0; i < req.kf_ids_from.size(); ++i) {
          if (i >= req.transform_est_success.size() || !req.transform_est_success[i]) continue;
          if (i >= req.kf_ids_to.size() || i >= req.frames_kepts_ids_from.size() || i >= req.frames_kepts_ids_to.size() ||
              i >= req.separators.size())
            continue;
          fwd.request.kf_ids_from.push_back(req.kf_ids_from[i]);
          fwd.request.kf_ids_to.push_back(req.kf_ids_to[i]);
          fwd.request.frames_kepts_ids_from.push_back(req.frames_kepts_ids_from[i]);
          fwd.request.frames_kepts_ids_to.push_back(req.frames_kepts_ids_to[i]);
          fwd.request.separators.push_back(req.separators[i]);
          fwd.request.transform_est_success.push_back(req.transform_est_success[i]);
          if (with_estimates) {
            fwd.request.pose_estimates_from.push_back(req.pose_estimates_from[i]);
            fwd.request.pose_estimates_to.push_back(req.pose_estimates_to[i]);
          }
        }
        if (!add_seps.call(fwd)) ROS_ERROR("Service call add sep to pose graph failed");
        res.success = o.success;
        return true;
      };
  ros::ServiceServer s1 = n.advertiseService("estimate_transformation", est);
  ros::ServiceServer s2 = n.advertiseService("find_matches_compute", fm);
  ros::ServiceServer s3 = n.advertiseService("receive_separators_py", rs);
  ros::ServiceServer s4 = n.advertiseService("add_keyframe", addkf);
  ROS_INFO("separators_b200 ready");
  ros::spin();   // single-threaded like the reference node (stereoCamGeometricTools.cpp:212)
  return 0;
}
#endif  // SEPB_WITH_ROS
