// Stand-in for tf2's Quaternion / Matrix3x3 (tf2/LinearMath, ROS 2 Humble): setRPY, normalize and getRPY restated from
// the published source (half-angle products; getRPY = the first solution of getEulerYPR).
#pragma once
// tf2/LinearMath/Scalar.h includes <math.h> and <stdlib.h>: with libstdc++ those put every std::abs overload into the global
// namespace, which is what makes the reference's unqualified abs(double) (FrontierSearch.hpp:107-108) the double one
#include <math.h>
#include <stdlib.h>
#include <cmath>
#include "tf2/LinearMath/Vector3.h"
namespace tf2 {
class Quaternion {
public:
  Quaternion() : v_{0, 0, 0, 1} {}
  Quaternion(double x, double y, double z, double w) : v_{x, y, z, w} {}
  void setRPY(double roll, double pitch, double yaw) {
    const double halfYaw = yaw * 0.5, halfPitch = pitch * 0.5, halfRoll = roll * 0.5;
    const double cosYaw = std::cos(halfYaw), sinYaw = std::sin(halfYaw);
    const double cosPitch = std::cos(halfPitch), sinPitch = std::sin(halfPitch);
    const double cosRoll = std::cos(halfRoll), sinRoll = std::sin(halfRoll);
    v_[0] = sinRoll * cosPitch * cosYaw - cosRoll * sinPitch * sinYaw;
    v_[1] = cosRoll * sinPitch * cosYaw + sinRoll * cosPitch * sinYaw;
    v_[2] = cosRoll * cosPitch * sinYaw - sinRoll * sinPitch * cosYaw;
    v_[3] = cosRoll * cosPitch * cosYaw + sinRoll * sinPitch * sinYaw;
  }
  double length2() const { return v_[0] * v_[0] + v_[1] * v_[1] + v_[2] * v_[2] + v_[3] * v_[3]; }
  Quaternion &normalize() {
    const double s = 1.0 / std::sqrt(length2());
    for (double &c : v_) c *= s;
    return *this;
  }
  double x() const { return v_[0]; }
  double y() const { return v_[1]; }
  double z() const { return v_[2]; }
  double w() const { return v_[3]; }

private:
  double v_[4];
};
class Matrix3x3 {
public:
  explicit Matrix3x3(const Quaternion &q) {
    const double d = q.length2(), s = 2.0 / d;
    const double xs = q.x() * s, ys = q.y() * s, zs = q.z() * s;
    const double wx = q.w() * xs, wy = q.w() * ys, wz = q.w() * zs;
    const double xx = q.x() * xs, xy = q.x() * ys, xz = q.x() * zs;
    const double yy = q.y() * ys, yz = q.y() * zs, zz = q.z() * zs;
    m_[0][0] = 1.0 - (yy + zz); m_[0][1] = xy - wz; m_[0][2] = xz + wy;
    m_[1][0] = xy + wz; m_[1][1] = 1.0 - (xx + zz); m_[1][2] = yz - wx;
    m_[2][0] = xz - wy; m_[2][1] = yz + wx; m_[2][2] = 1.0 - (xx + yy);
  }
  void getRPY(double &roll, double &pitch, double &yaw) const {
    if (std::fabs(m_[2][0]) >= 1) {
      yaw = 0;
      const double delta = std::atan2(m_[2][1], m_[2][2]);
      if (m_[2][0] < 0) { pitch = M_PI / 2.0; roll = delta; }
      else { pitch = -M_PI / 2.0; roll = delta; }
    } else {
      pitch = -std::asin(m_[2][0]);
      roll = std::atan2(m_[2][1] / std::cos(pitch), m_[2][2] / std::cos(pitch));
      yaw = std::atan2(m_[1][0] / std::cos(pitch), m_[0][0] / std::cos(pitch));
    }
  }

private:
  double m_[3][3];
};
}  // namespace tf2
