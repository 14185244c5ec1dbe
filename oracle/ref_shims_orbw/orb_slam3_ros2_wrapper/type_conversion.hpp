// Stand-in for orb_slam3_ros2_wrapper/type_conversion.hpp: the declarations of the member functions oracle/Makefile cuts out of
// ORBW/src/type_conversion.cpp (se3ROSToORB :74-100, vector3fORBToROS :102-121, poseToAffine :129-141, rotationORBToEulerROS
// :143-153, transformPointWithReference<Eigen::Vector3f> :296-301), without the ROS / pcl / OpenCV includes of the real header.
// Test infrastructure only.
#pragma once
#include <cmath>
#include <Eigen/Core>
#include <Eigen/Geometry>
#include <geometry_msgs/msg/pose.hpp>
#include <sophus/se3.hpp>

namespace ORB_SLAM3_Wrapper {
class WrapperTypeConversions {
public:
  Sophus::SE3f se3ROSToORB(const Eigen::Affine3f &affineMatrix);
  Eigen::Vector3f vector3fORBToROS(const Eigen::Vector3f &s);
  Eigen::Affine3f poseToAffine(const geometry_msgs::msg::Pose &pose);
  Eigen::Vector3f rotationORBToEulerROS(const Eigen::Matrix3f &rotation);
  template <typename T>
  T transformPointWithReference(Eigen::Affine3f &referencePose, Eigen::Vector3f &s);
};
}  // namespace ORB_SLAM3_Wrapper
