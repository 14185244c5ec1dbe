// Stand-in for geometry_msgs/msg/quaternion.hpp (ROS 2 message struct; identity by default, as generated).
#pragma once
namespace geometry_msgs { namespace msg {
struct Quaternion {
  double x = 0.0, y = 0.0, z = 0.0, w = 1.0;
  bool operator==(const Quaternion &o) const { return x == o.x && y == o.y && z == o.z && w == o.w; }
};
} }
