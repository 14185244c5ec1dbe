// Stand-in for nav2_util/geometry_utils.hpp (ROS 2 Humble): orientationAroundZAxis = tf2 setRPY(0, 0, angle) as a message.
#pragma once
#include "geometry_msgs/msg/quaternion.hpp"
#include "tf2/LinearMath/Quaternion.h"
namespace nav2_util { namespace geometry_utils {
inline geometry_msgs::msg::Quaternion orientationAroundZAxis(double angle) {
  tf2::Quaternion q;
  q.setRPY(0, 0, angle);
  geometry_msgs::msg::Quaternion m;
  m.x = q.x(); m.y = q.y(); m.z = q.z(); m.w = q.w();
  return m;
}
} }
