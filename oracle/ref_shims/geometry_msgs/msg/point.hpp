// Stand-in for geometry_msgs/msg/point.hpp (ROS 2 message struct: three doubles, member-wise equality).
#pragma once
namespace geometry_msgs { namespace msg {
struct Point {
  double x = 0.0, y = 0.0, z = 0.0;
  bool operator==(const Point &o) const { return x == o.x && y == o.y && z == o.z; }
  bool operator!=(const Point &o) const { return !(*this == o); }
};
} }
