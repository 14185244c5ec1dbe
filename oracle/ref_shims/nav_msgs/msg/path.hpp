// Stand-in for nav_msgs/msg/path.hpp (declarations in the reference's Helpers.hpp only name the type).
#pragma once
#include <vector>
#include "geometry_msgs/msg/pose.hpp"
namespace nav_msgs { namespace msg {
struct Path {
  struct { std::string frame_id; } header;
  std::vector<geometry_msgs::msg::PoseStamped> poses;
};
} }
