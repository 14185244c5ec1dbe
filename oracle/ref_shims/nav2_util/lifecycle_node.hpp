// Stand-in for nav2_util/lifecycle_node.hpp.
#pragma once
#include <rclcpp/rclcpp.hpp>
namespace nav2_util { class LifecycleNode : public rclcpp::Node {}; }
