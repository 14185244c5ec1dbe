// Stand-in for geometry_msgs/msg/pose.hpp and pose_stamped (ROS 2 message structs).
#pragma once
#include <string>
#include "geometry_msgs/msg/point.hpp"
#include "geometry_msgs/msg/quaternion.hpp"
namespace geometry_msgs { namespace msg {
struct Pose {
  Point position;
  Quaternion orientation;
};
struct PoseStamped {
  struct { std::string frame_id; } header;
  Pose pose;
};
} }
