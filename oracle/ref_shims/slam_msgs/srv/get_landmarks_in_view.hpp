// Stand-in for slam_msgs/srv/GetLandmarksInView (ORB-SLAM3-ROS2-Docker/slam_msgs/srv/GetLandmarksInView.srv): the request fields
// FisherInfoManager.cpp fills and the map_points of the response (slam_msgs/MapPoint: position float64 x y z).
#pragma once
#include <vector>
#include "geometry_msgs/msg/point.hpp"
#include "geometry_msgs/msg/pose.hpp"
namespace slam_msgs {
namespace msg { struct MapPoint { geometry_msgs::msg::Point position; }; }
namespace srv {
struct GetLandmarksInView {
  struct Request {
    geometry_msgs::msg::Pose pose;
    float max_dist_pose_observation = 0;
    float max_angle_pose_observation = 0;
    bool exhaustive_search = false;
  };
  struct Response { std::vector<slam_msgs::msg::MapPoint> map_points; };
};
} }
