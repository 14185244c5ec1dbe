// Stand-in for orb_slam3_ros2_wrapper/orb_slam3_interface.hpp and for the ORB_SLAM3 classes mapPointsVisibleFromPose walks
// (ORB_SLAM3 needs OpenCV, Pangolin, DBoW2, g2o: not in this image).  The bodies compiled against it are the reference's own:
// ORBSLAM3Interface::mapPointsVisibleFromPose, both overloads (ORBW/src/orb_slam3_interface.cpp:174-317) and
// Pinhole::project(const Eigen::Vector3f &) (ORB/src/CameraModels/Pinhole.cpp:43-49).  The atlas is what the harness fills in:
// keyframes in the caller's order (Map::GetAllKeyFrames() of ORB_SLAM3 walks a std::set ordered by pointer), each with its
// GetPoseInverse() and its map points in the caller's order (KeyFrame::GetMapPoints() returns a std::set<MapPoint *>).
// Tracking::externalKFAdded() answers true and LoopClosing::loopDetected() false, so neither wait loop spins.
// Test infrastructure only.
#pragma once
#include <chrono>
#include <iostream>
#include <memory>
#include <stdexcept>
#include <thread>
#include <unordered_set>
#include <vector>
#include <Eigen/Core>
#include <Eigen/Geometry>
#include <geometry_msgs/msg/pose.hpp>
#include <sophus/se3.hpp>
#include "orb_slam3_ros2_wrapper/type_conversion.hpp"

namespace ORB_SLAM3 {
class MapPoint {
public:
  explicit MapPoint(const Eigen::Vector3f &world) : world_(world) {}
  Eigen::Vector3f GetWorldPos() { return world_; }
  float GetMaxDistanceInvariance() { return 0.0f; }  // read into locals the reference never uses (its range test is commented out)
  float GetMinDistanceInvariance() { return 0.0f; }

private:
  Eigen::Vector3f world_;
};
class KeyFrame {
public:
  Sophus::SE3f GetPoseInverse() { return Twc; }
  std::vector<MapPoint *> GetMapPoints() { return points; }
  Sophus::SE3f Twc;
  std::vector<MapPoint *> points;
};
class GeometricCamera {
public:
  virtual ~GeometricCamera() {}
  virtual Eigen::Vector2f project(const Eigen::Vector3f &v3D) = 0;
};
class Pinhole : public GeometricCamera {
public:
  explicit Pinhole(const std::vector<float> &parameters) : mvParameters(parameters) {}
  Eigen::Vector2f project(const Eigen::Vector3f &v3D) override;

protected:
  std::vector<float> mvParameters;  // fx, fy, cx, cy
};
class Map {
public:
  std::vector<KeyFrame *> GetAllKeyFrames() { return keyframes; }
  std::vector<KeyFrame *> keyframes;
};
class Atlas {
public:
  std::vector<GeometricCamera *> GetAllCameras() { return cameras; }
  Map *GetCurrentMap() { return &map; }
  std::vector<GeometricCamera *> cameras;
  Map map;
};
class Tracking {
public:
  void needExternalKF() {}
  bool externalKFAdded() { return true; }
};
class LoopClosing {
public:
  bool loopDetected() { return false; }
};
class System {
public:
  Tracking *getTracking() { return &tracking_; }
  LoopClosing *GetLoopClosing() { return &loop_closing_; }

private:
  Tracking tracking_;
  LoopClosing loop_closing_;
};
}  // namespace ORB_SLAM3

namespace ORB_SLAM3_Wrapper {
class ORBSLAM3Interface {
public:
  ORBSLAM3Interface() : mSLAM_(std::make_shared<ORB_SLAM3::System>()), typeConversions_(std::make_shared<WrapperTypeConversions>()) {}
  void mapPointsVisibleFromPose(geometry_msgs::msg::Pose cameraPose, std::vector<ORB_SLAM3::MapPoint *> &points, int maxLandmarks,
                                float maxDistance, float maxAngle, bool exhaustive_search);
  void mapPointsVisibleFromPose(Sophus::SE3f &cameraPose, std::vector<ORB_SLAM3::MapPoint *> &points, int maxLandmarks,
                                float maxDistance, float maxAngle, bool exhaustive_search);
  std::shared_ptr<WrapperTypeConversions> getTypeConversionPtr() { return typeConversions_; }

  std::shared_ptr<ORB_SLAM3::System> mSLAM_;
  std::shared_ptr<WrapperTypeConversions> typeConversions_;
  ORB_SLAM3::Atlas atlas_, *orbAtlas_ = &atlas_;
  bool loopClosing_ = true;
  Eigen::Affine3f robotBase_to_cameraLink_;  // identity: initialRobotPose_ of SURVEY.md 8d
};
}  // namespace ORB_SLAM3_Wrapper
