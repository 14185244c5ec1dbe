// Stand-in for <rclcpp/rclcpp.hpp> (ROS 2 Humble is not in this image): the reference's frontier search only asks whether
// the process is still up (DEP/src/FrontierSearch.cpp:58,123).  Test infrastructure for oracle/_ref, never shipped.
#pragma once
// (the real header brings these standard headers in; the reference's sources rely on that)
#include <algorithm>
#include <memory>
#include <queue>
#include <string>
#include <vector>
namespace rclcpp {
inline bool ok() { return true; }
}  // namespace rclcpp
