// Stand-in for frontier_exploration/CostCalculator.hpp: the declaration of FrontierCostCalculator reduced to what the member
// functions compiled from the reference's CostCalculator.cpp use (oracle/Makefile cuts setArrivalInformationForFrontier,
// setMaxArrivalInformation, setPlanForFrontierEuclidean and recomputeNormalizationFactors out of the source where it lies; their
// bodies are the reference's, this header only replaces the includes the image does not have: rclcpp nodes, pcl, sensor_msgs,
// visualization_msgs, slam_msgs, the roadmap planner and the Fisher manager).  The constructor takes its three parameters and
// the costmap directly instead of reading a Costmap2DROS and the parameter server.
#pragma once
#include <algorithm>
#include <cmath>
#include <limits>
#include <memory>
#include <numeric>
#include <stdexcept>
#include <string>
#include <vector>

#include <frontier_exploration/Frontier.hpp>
#include <frontier_exploration/Helpers.hpp>
#include <frontier_exploration/util/GeometryUtils.hpp>
#include <frontier_exploration/util/logger.hpp>
#include <geometry_msgs/msg/pose.hpp>
#include <nav2_costmap_2d/costmap_2d.hpp>

namespace slam_msgs { namespace srv { struct GetMap_Response {}; } }

namespace frontier_exploration {

class FrontierCostCalculator {
public:
  FrontierCostCalculator(nav2_costmap_2d::Costmap2D *costmap, double max_camera_depth, double delta_theta, double camera_fov,
                         double robot_radius)
      : exploration_costmap_(costmap), robot_radius_(robot_radius), MAX_CAMERA_DEPTH(max_camera_depth), DELTA_THETA(delta_theta),
        CAMERA_FOV(camera_fov) {}

  void setArrivalInformationForFrontier(FrontierPtr &goal_point_w, std::vector<double> &polygon_xy_min_max);
  double setMaxArrivalInformation();
  void setPlanForFrontierEuclidean(geometry_msgs::msg::Pose start_pose_w, FrontierPtr &goal_point_w,
                                   std::shared_ptr<slam_msgs::srv::GetMap_Response> map_data, bool compute_information,
                                   bool planner_allow_unknown_);
  void recomputeNormalizationFactors(FrontierPtr &frontier);
  // the cost types this library does not carry (A* / roadmap planners, random and closest-frontier costs): assignCosts names them,
  // the harness never selects them
  void setPlanForFrontier(geometry_msgs::msg::Pose, FrontierPtr &, std::shared_ptr<slam_msgs::srv::GetMap_Response>, bool, bool) { throw std::logic_error("not in oracle/_ref"); }
  void setPlanForFrontierRoadmap(geometry_msgs::msg::Pose, FrontierPtr &, std::shared_ptr<slam_msgs::srv::GetMap_Response>, bool, bool) { throw std::logic_error("not in oracle/_ref"); }
  void setRandomMetaData(FrontierPtr &) { throw std::logic_error("not in oracle/_ref"); }
  void setClosestFrontierMetaData(geometry_msgs::msg::Pose, FrontierPtr &, std::shared_ptr<slam_msgs::srv::GetMap_Response>, bool, bool) { throw std::logic_error("not in oracle/_ref"); }

  double getMinPlanDistance() { return min_traversable_distance; }
  double getMaxPlanDistance() { return max_traversable_distance; }
  double getMinArrivalInformation() { return min_arrival_info_per_frontier; }
  double getMaxArrivalInformation() { return max_arrival_info_gt_; }
  double getMinArrivalInformationGT() { return min_arrival_info_gt_; }
  void reset() {  // as the reference's inline reset()
    min_traversable_distance = std::numeric_limits<double>::max();
    max_traversable_distance = -1.0;
    min_arrival_info_per_frontier = std::numeric_limits<double>::max();
    max_arrival_info_per_frontier = -1.0;
  }

private:
  nav2_costmap_2d::Costmap2D *exploration_costmap_;
  double min_traversable_distance = std::numeric_limits<double>::max();
  double max_traversable_distance = 0.0;
  double min_arrival_info_per_frontier = std::numeric_limits<double>::max();
  double max_arrival_info_per_frontier = 0.0;
  double robot_radius_;
  double max_arrival_info_gt_ = 0.0;
  double min_arrival_info_gt_ = 0.0;
  bool arrival_info_limits_set_ = false;
  double MAX_CAMERA_DEPTH;
  double DELTA_THETA;
  double CAMERA_FOV;
};

}  // namespace frontier_exploration
