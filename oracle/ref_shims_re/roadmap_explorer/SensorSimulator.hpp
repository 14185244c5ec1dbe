// Stand-in for roadmap_explorer/SensorSimulator.hpp: the members SensorSimulator::markRay touches (RE/src/SensorSimulator.cpp:
// 211-246, compiled from the reference by oracle/Makefile), without the node, the tf buffer, the subscriptions and the map I/O.
#pragma once
#include <cmath>
#include <memory>
#include <mutex>
#include <nav_msgs/msg/occupancy_grid.hpp>
#include "roadmap_explorer/Parameters.hpp"
#include "roadmap_explorer/util/Logger.hpp"
namespace roadmap_explorer {
class SensorSimulator {
public:
  SensorSimulator(nav_msgs::msg::OccupancyGrid::SharedPtr latest, nav_msgs::msg::OccupancyGrid::SharedPtr explored)
      : latest_map_(latest), explored_map_(explored) {}
  void markRay(const geometry_msgs::msg::Pose &base_pose, double ray_angle);
private:
  nav_msgs::msg::OccupancyGrid::SharedPtr latest_map_;
  nav_msgs::msg::OccupancyGrid::SharedPtr explored_map_;
  std::recursive_mutex map_mutex_;
};
}  // namespace roadmap_explorer
