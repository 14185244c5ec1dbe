// Stand-in for sophus/se3.hpp (Sophus is not in this image): SE3f kept as (rotation matrix, translation).  Sophus itself stores a
// unit quaternion: its constructor converts the matrix, rotationMatrix() converts back, and a product multiplies quaternions, so
// the real values differ from these by a few float roundings; that round trip is NOT modelled (the same
// model as oracle.cpp's).  What oracle/_ref/libfitslam_ref_orbw.so pins is therefore the reference's control flow and sequence
// of operations on top of this model, not Sophus's arithmetic.  Test infrastructure only.
#pragma once
#include <Eigen/Core>
#include <Eigen/Geometry>

namespace Sophus {
struct SE3f {
  Eigen::Matrix3f R;
  Eigen::Vector3f t;
  SE3f() : R(Eigen::Matrix3f::Identity()) {}
  SE3f(const Eigen::Matrix3f &rotation, const Eigen::Vector3f &translation) : R(rotation), t(translation) {}
  SE3f inverse() const {
    const Eigen::Matrix3f Rt = R.transpose();
    return SE3f(Rt, -(Rt * t));
  }
  SE3f operator*(const SE3f &o) const { return SE3f(R * o.R, R * o.t + t); }
  Eigen::Matrix3f rotationMatrix() const { return R; }
  Eigen::Vector3f &translation() { return t; }
  const Eigen::Vector3f &translation() const { return t; }
};
}  // namespace Sophus
