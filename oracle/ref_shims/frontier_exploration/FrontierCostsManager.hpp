// Stand-in for frontier_exploration/FrontierCostsManager.hpp: the class declaration with the reference's member names, without
// the ROS / pcl includes; oracle/Makefile compiles findDuplicates, FrontierCostsManager::assignCosts and setFrontierBlacklist
// (DEP/src/FrontierCostsManager.cpp:25-223) against it.  The constructor takes the parameters directly.
#pragma once
#include <cmath>
#include <memory>
#include <mutex>
#include <stdexcept>
#include <string>
#include <unordered_map>
#include <vector>

#include <frontier_exploration/CostCalculator.hpp>
#include <frontier_exploration/Frontier.hpp>
#include "frontier_exploration/util/general_utils.hpp"

namespace frontier_exploration {

class FrontierCostsManager {
public:
  FrontierCostsManager(nav2_costmap_2d::Costmap2D *costmap, double max_camera_depth, double delta_theta, double camera_fov,
                       double robot_radius, double alpha, double beta, double vx_max, double wz_max)
      : costmap_(costmap), costCalculator_(std::make_shared<FrontierCostCalculator>(costmap, max_camera_depth, delta_theta, camera_fov, robot_radius)),
        planner_allow_unknown_(false), alpha_(alpha), beta_(beta), max_vx_(vx_max), max_wx_(wz_max) {}

  bool assignCosts(std::vector<FrontierPtr> &frontier_list, std::vector<double> polygon_xy_min_max, geometry_msgs::msg::Pose start_pose_w,
                   std::shared_ptr<slam_msgs::srv::GetMap_Response> map_data, std::vector<std::vector<std::string>> &costTypes);
  void setFrontierBlacklist(std::vector<FrontierPtr> &blacklist);
  std::shared_ptr<FrontierCostCalculator> getCostCalcPtr() { return costCalculator_; }

private:
  nav2_costmap_2d::Costmap2D *costmap_;
  std::shared_ptr<FrontierCostCalculator> costCalculator_;
  std::unordered_map<FrontierPtr, bool, FrontierHash, FrontierGoalPointEquality> frontier_blacklist_;
  bool planner_allow_unknown_;
  std::mutex blacklist_mutex_;
  double alpha_;
  double beta_;
  double max_vx_;
  double max_wx_;
};

}  // namespace frontier_exploration
