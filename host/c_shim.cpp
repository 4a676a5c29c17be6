// Test shim: plain-C entry points that build the POD requests, call the C++ service handlers and flatten the
// responses, so tests/test_host_services_gpu.py can drive them with numpy arrays through ctypes.
#include <cstring>

#include "separator_services.hpp"

using namespace sepb_host;

static void fill(msgs::Descriptors& d, msgs::KeyPointVec& k, msgs::KeyPoint3DVec& k3, const uint8_t* desc, int n,
                 const float* kp2, const int32_t* oct, const float* kp3) {
  d.rows = (uint16_t)n;
  d.cols = 32;
  d.data.assign(desc, desc + (size_t)n * 32);
  k.size = (int16_t)n;
  k.kptsVec.resize(n);
  k3.size = (int16_t)n;
  k3.kpts3DVec.resize(n);
  for (int i = 0; i < n; ++i) {
    k.kptsVec[i].pt.x = kp2[2 * i];
    k.kptsVec[i].pt.y = kp2[2 * i + 1];
    k.kptsVec[i].octave = oct ? oct[i] : 0;
    k3.kpts3DVec[i].x = kp3[3 * i];
    k3.kpts3DVec[i].y = kp3[3 * i + 1];
    k3.kpts3DVec[i].z = kp3[3 * i + 2];
  }
}

extern "C" {

void* sh_device_create(int device) {
  Device* d = new Device(device);
  return d;
}
int sh_device_ok(void* d) { return static_cast<Device*>(d)->ok() ? 1 : 0; }
void sh_device_destroy(void* d) { delete static_cast<Device*>(d); }

void* sh_geo_create(void* dev, double fx, double fy, double cx, double cy, int w, int h, int min_inliers, int max_feat) {
  CameraModel cam{fx, fy, cx, cy, w, h};
  return new GeometricTools(*static_cast<Device*>(dev), cam, min_inliers, max_feat);
}
void sh_geo_destroy(void* g) { delete static_cast<GeometricTools*>(g); }
// Vis/PnPFlags, Vis/PnPRefineIterations, the OpenCV 3.2 return convention, the camera's localTransform (NULL keeps it)
void sh_geo_configure(void* g, int pnp_flags, int refine_iterations, int opencv32, const double* local_R9, const double* local_t3) {
  GeometricTools* gt = static_cast<GeometricTools*>(g);
  if (local_R9 && local_t3) gt->setLocalTransform(local_R9, local_t3);
  gt->setPnPFlags(pnp_flags);
  gt->setPnPRefineIterations(refine_iterations);
  gt->setOpenCV32Semantics(opencv32 != 0);
}

// Vis/BundleAdjustment (0 off, 1 g2o-style) with the stereo baseline in metres; 1 on success
int sh_geo_bundle_adjustment(void* g, int type, double stereo_baseline) {
  return static_cast<GeometricTools*>(g)->setBundleAdjustment(type, stereo_baseline) ? 1 : 0;
}

// returns res.success; pose7 = position xyz + orientation xyzw
int sh_geo_estimate(void* g, const uint8_t* descF, int nF, const float* kp2F, const int32_t* octF, const float* kp3F,
                    const uint8_t* descT, int nT, const float* kp2T, const int32_t* octT, const float* kp3T, double* pose7,
                    double* cov36, int* inliers, int* matches) {
  msgs::EstTransform::Request req;
  msgs::EstTransform::Response res;
  fill(req.descriptorsFrom, req.kptsFrom, req.kptsFrom3D, descF, nF, kp2F, octF, kp3F);
  fill(req.descriptorsTo, req.kptsTo, req.kptsTo3D, descT, nT, kp2T, octT, kp3T);
  GeometricTools* gt = static_cast<GeometricTools*>(g);
  if (!gt->estimateTransformation(req, res)) return -1;
  const msgs::Pose& p = res.poseWithCov.pose;
  const double v[7] = {p.position.x, p.position.y, p.position.z, p.orientation.x, p.orientation.y, p.orientation.z, p.orientation.w};
  std::memcpy(pose7, v, sizeof v);
  std::memcpy(cov36, res.poseWithCov.covariance, 36 * sizeof(double));
  if (inliers) *inliers = gt->lastInliers();
  if (matches) *matches = gt->lastMatches();
  return res.success ? 1 : 0;
}

void* sh_pr_create(void* dev, int dim, double distance, int max_nb, int capacity) {
  return new PlaceRecognition(*static_cast<Device*>(dev), dim, distance, max_nb, capacity);
}
void sh_pr_destroy(void* p) { delete static_cast<PlaceRecognition*>(p); }

void sh_pr_add_keyframe(void* p, const double* netvlad, int n_values, const uint8_t* desc, int n, const float* kp2,
                        const int32_t* oct, const float* kp3, int kf_id) {
  msgs::Descriptors d;
  msgs::KeyPointVec k;
  msgs::KeyPoint3DVec k3;
  fill(d, k, k3, desc, n, kp2, oct, kp3);
  static_cast<PlaceRecognition*>(p)->addKeyframe(std::vector<double>(netvlad, netvlad + n_values), d, k3, k, (int16_t)kf_id);
}
int sh_pr_dropped_int16(void* p) { return static_cast<PlaceRecognition*>(p)->droppedForInt16(); }
const char* sh_pr_last_error(void* p) { return static_cast<PlaceRecognition*>(p)->lastError().c_str(); }

// returns the number of matches; ids into the three int16 arrays (capacity max_out); feature rows of the first match
int sh_pr_find_matches_compute(void* p, const double* new_desc, int n_values, int16_t* kf_ids, int16_t* local_ids,
                               int16_t* other_ids, int max_out, int* first_match_rows) {
  msgs::FindMatches::Request req;
  msgs::FindMatches::Response res;
  req.new_netvlad_descriptors.assign(new_desc, new_desc + n_values);
  if (!static_cast<PlaceRecognition*>(p)->findMatchesCompute(req, res)) return -1;
  const int n = (int)res.frames_kept_ids_computing_robot.size();
  for (int i = 0; i < n && i < max_out; ++i) {
    kf_ids[i] = res.kf_ids_computing_robot[i];
    local_ids[i] = res.frames_kept_ids_computing_robot[i];
    other_ids[i] = res.frames_kept_ids_querying_robot[i];
  }
  if (first_match_rows) *first_match_rows = n ? res.descriptors_vec[0].rows : 0;
  return n;
}

int sh_pr_receive_separators(void* p, int n, const int16_t* frames_from, const int16_t* frames_to, const uint8_t* success) {
  msgs::ReceiveSeparators::Request req;
  msgs::ReceiveSeparators::Response res;
  for (int i = 0; i < n; ++i) {
    req.kf_ids_from.push_back(frames_from[i]);
    req.kf_ids_to.push_back(frames_to[i]);
    req.frames_kepts_ids_from.push_back(frames_from[i]);
    req.frames_kepts_ids_to.push_back(frames_to[i]);
    req.transform_est_success.push_back(success[i]);
  }
  static_cast<PlaceRecognition*>(p)->receiveSeparators(req, res);
  return res.success ? 1 : 0;
}

}  // extern "C"
