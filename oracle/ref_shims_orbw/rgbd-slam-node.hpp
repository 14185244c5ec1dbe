// Stand-in for ORBW/src/rgbd/rgbd-slam-node.hpp: the node with the one member its landmark service handler uses
// (getMapPointsInViewServer, ORBW/src/rgbd/rgbd-slam-node.cpp:406-429 up to `response->map_points = landmarks;`; the two debug
// publishers after it stay out).  Test infrastructure only.
#pragma once
#include <memory>
#include <vector>
#include <slam_msgs/srv/get_landmarks_in_view.hpp>
#include "orb_slam3_ros2_wrapper/orb_slam3_interface.hpp"

#define RCLCPP_INFO(logger, ...) do { (void)(logger); } while (0)
struct rmw_request_id_t {};

namespace ORB_SLAM3_Wrapper {
class RgbdSlamNode {
public:
  RgbdSlamNode() : interface_(std::make_shared<ORBSLAM3Interface>()) {}
  int get_logger() { return 0; }
  void getMapPointsInViewServer(std::shared_ptr<rmw_request_id_t> request_header,
                                std::shared_ptr<slam_msgs::srv::GetLandmarksInView::Request> request,
                                std::shared_ptr<slam_msgs::srv::GetLandmarksInView::Response> response);
  std::shared_ptr<ORBSLAM3Interface> interface_;
};
}  // namespace ORB_SLAM3_Wrapper
