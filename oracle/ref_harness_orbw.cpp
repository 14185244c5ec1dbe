// ref_harness_orbw.cpp -- C entry point over the ORB-SLAM3 wrapper's own landmark query, compiled from /root/reference where it
// lies into oracle/_ref/libfitslam_ref_orbw.so (oracle/Makefile): ORBSLAM3Interface::mapPointsVisibleFromPose, both overloads
// (ORBW/src/orb_slam3_interface.cpp:174-317), the service handler RgbdSlamNode::getMapPointsInViewServer
// (ORBW/src/rgbd/rgbd-slam-node.cpp:406-429), WrapperTypeConversions::{se3ROSToORB, vector3fORBToROS, poseToAffine,
// rotationORBToEulerROS, transformPointWithReference<Vector3f>} (ORBW/src/type_conversion.cpp) and Pinhole::project
// (ORB/src/CameraModels/Pinhole.cpp:43-49), against stand-in headers (oracle/ref_shims_orbw, oracle/ref_shims): ORB_SLAM3's
// atlas classes, Sophus::SE3f kept as (R, t) and the fixed-size Eigen stand-in.  What this pins is the reference's control flow
// and sequence of operations (prefilter per keyframe, the maxLandmarks break, each map point once, the image test, the transform
// into the query-pose frame); the arithmetic INSIDE Sophus (quaternion storage) and Eigen's Affine rotation() SVD is the
// stand-ins' (R, t) model.  Test infrastructure only (tests/test_ref_pin.py pins rows F1, F2 and the keyframe mode of 8f-1).
#include <cstdint>
#include <cstring>
#include <iostream>
#include <memory>
#include <sstream>
#include <string>
#include <vector>

#include "rgbd-slam-node.hpp"

namespace {
// std::cout of the reference's functions goes into a string for the length of a call ("Number of keyframes checked: N" is the
// only place the reference reports that number)
struct CoutCapture {
  std::ostringstream text;
  std::streambuf *old;
  CoutCapture() : old(std::cout.rdbuf(text.rdbuf())) {}
  ~CoutCapture() { std::cout.rdbuf(old); }
};
long number_after(const std::string &s, const char *label) {
  const size_t at = s.rfind(label);
  if (at == std::string::npos) return -1;
  return std::strtol(s.c_str() + at + std::strlen(label), nullptr, 10);
}
}  // namespace

extern "C" {

// pose7: x, y, z, qx, qy, qz, qw (doubles, geometry_msgs/Pose).  xyz: P map points in the ROS map frame (their ORB world
// coordinates are the axis permutation (-y, -z, x)).  Keyframe k: GetPoseInverse() = (Rwc row-major, twc) in kf_pose[12 k ..],
// map points point_index[kf_first[k] .. + kf_count[k]).  cam: fx, fy, cx, cy.
// max_landmarks == 1000 goes through the service handler (which passes the literal 1000); any other value calls the interface
// with it and then runs the handler's per-point statements (rgbd-slam-node.cpp:414-427) on the same reference functions.
// out_xyz: capacity * 3 doubles (the response's float64 positions).  Returns the number of landmarks; *n_checked = the number the
// reference prints as "Number of keyframes checked".
int ref_orbw_landmarks_in_view(const double *pose7, const float *xyz, uint64_t P, const float *kf_pose, const uint32_t *kf_first,
                               const uint32_t *kf_count, uint32_t n_kf, const uint32_t *point_index, int max_landmarks,
                               float max_distance, int exhaustive, const float *cam, double *out_xyz, uint32_t capacity,
                               int *n_checked) {
  using namespace ORB_SLAM3_Wrapper;
  RgbdSlamNode node;
  ORB_SLAM3::Atlas &atlas = node.interface_->atlas_;
  ORB_SLAM3::Pinhole camera(std::vector<float>{cam[0], cam[1], cam[2], cam[3]});
  atlas.cameras.push_back(&camera);
  std::vector<ORB_SLAM3::MapPoint> points;
  points.reserve(P);
  for (uint64_t i = 0; i < P; ++i) points.emplace_back(Eigen::Vector3f(-xyz[3 * i + 1], -xyz[3 * i + 2], xyz[3 * i]));
  std::vector<ORB_SLAM3::KeyFrame> keyframes(n_kf);
  for (uint32_t k = 0; k < n_kf; ++k) {
    const float *r = kf_pose + 12 * (size_t)k;
    Eigen::Matrix3f Rwc;
    for (int a = 0; a < 3; ++a)
      for (int b = 0; b < 3; ++b) Rwc(a, b) = r[3 * a + b];
    keyframes[k].Twc = Sophus::SE3f(Rwc, Eigen::Vector3f(r[9], r[10], r[11]));
    for (uint32_t e = kf_first[k]; e < kf_first[k] + kf_count[k]; ++e) keyframes[k].points.push_back(&points[point_index[e]]);
    atlas.map.keyframes.push_back(&keyframes[k]);
  }
  geometry_msgs::msg::Pose pose;
  pose.position.x = pose7[0];
  pose.position.y = pose7[1];
  pose.position.z = pose7[2];
  pose.orientation.x = pose7[3];
  pose.orientation.y = pose7[4];
  pose.orientation.z = pose7[5];
  pose.orientation.w = pose7[6];

  CoutCapture capture;
  std::vector<slam_msgs::msg::MapPoint> landmarks;
  if (max_landmarks == 1000) {
    auto request = std::make_shared<slam_msgs::srv::GetLandmarksInView::Request>();
    auto response = std::make_shared<slam_msgs::srv::GetLandmarksInView::Response>();
    request->pose = pose;
    request->max_dist_pose_observation = max_distance;
    request->max_angle_pose_observation = 0.0f;  // passed on as maxAngle, which the reference never reads
    request->exhaustive_search = exhaustive != 0;
    node.getMapPointsInViewServer(std::make_shared<rmw_request_id_t>(), request, response);
    landmarks = response->map_points;
  } else {
    std::vector<ORB_SLAM3::MapPoint *> found;
    node.interface_->mapPointsVisibleFromPose(pose, found, max_landmarks, max_distance, 0.0f, exhaustive != 0);
    auto conv = node.interface_->getTypeConversionPtr();
    auto affineMapToPos = conv->poseToAffine(pose);
    auto affinePosToMap = affineMapToPos.inverse();
    for (auto &point : found) {
      slam_msgs::msg::MapPoint landmark;
      Eigen::Vector3f landmark_position = point->GetWorldPos();
      auto position = conv->vector3fORBToROS(landmark_position);
      position = conv->transformPointWithReference<Eigen::Vector3f>(affinePosToMap, position);
      landmark.position.x = position.x();
      landmark.position.y = position.y();
      landmark.position.z = position.z();
      landmarks.push_back(landmark);
    }
  }
  if (n_checked) *n_checked = (int)number_after(capture.text.str(), "Number of keyframes checked: ");
  for (size_t i = 0; i < landmarks.size() && i < capacity; ++i) {
    out_xyz[3 * i] = landmarks[i].position.x;
    out_xyz[3 * i + 1] = landmarks[i].position.y;
    out_xyz[3 * i + 2] = landmarks[i].position.z;
  }
  return (int)landmarks.size();
}

// WrapperTypeConversions::rotationORBToEulerROS on a row-major 3 x 3 (known answers for the Euler convention)
void ref_orbw_euler_ros(const float *R9, float *out3) {
  Eigen::Matrix3f R;
  for (int a = 0; a < 3; ++a)
    for (int b = 0; b < 3; ++b) R(a, b) = R9[3 * a + b];
  ORB_SLAM3_Wrapper::WrapperTypeConversions conv;
  const Eigen::Vector3f e = conv.rotationORBToEulerROS(R);
  out3[0] = e(0);
  out3[1] = e(1);
  out3[2] = e(2);
}

}  // extern "C"
