// Stand-in for frontier_exploration/nav2_plugins/lethal_marker.hpp (the real header needs nav2_costmap_2d::Layer, tf2_ros, rclcpp
// services and frontier_msgs, none of which is in this image).  It declares the class with the members that addNewMarkedArea,
// addNewMarkedAreaFOV, markCells and updateCosts touch; their BODIES, and the free ray / circle / triangle functions above them,
// are the reference's own (DEP/src/nav2_plugins/lethal_marker.cpp:1-148, 218-244, 305-327, cut by oracle/Makefile).  The layer's
// node, service, parameter callback and bounds bookkeeping stay out.  Test infrastructure for oracle/_ref, never shipped.
#pragma once
#include <cmath>
#include <cstdint>
#include <map>
#include <mutex>
#include <vector>
#include <nav2_costmap_2d/costmap_2d.hpp>

#ifndef LOG_INFO
#define LOG_INFO(X) do { } while (0)
#endif

namespace nav2_costmap_2d {
// the one member of LayeredCostmap the layer calls
class LayeredCostmap {
public:
  explicit LayeredCostmap(Costmap2D *master) : master_(master) {}
  Costmap2D *getCostmap() { return master_; }

private:
  Costmap2D *master_;
};
}  // namespace nav2_costmap_2d

namespace frontier_exploration {
struct WorldLocation {
  float x;
  float y;
};

class LethalMarker {
public:
  LethalMarker(nav2_costmap_2d::LayeredCostmap *layered) : layered_costmap_(layered), numAdditions_(0) {}
  void updateCosts(nav2_costmap_2d::Costmap2D &master_grid, int min_i, int min_j, int max_i, int max_j);
  void addNewMarkedArea(double center_wx, double center_wy, double radius);
  void addNewMarkedAreaFOV(double center_wx, double center_wy, double angleYaw, double radius);
  int additions() const { return numAdditions_; }

private:
  void markCells(unsigned char *master_array, const std::vector<unsigned int> &cells_to_mark);
  const std::map<int, std::vector<unsigned int>> &getIndexCache() {
    std::lock_guard<std::mutex> lock(cacheMutex_);
    return latest_cells_to_mark_index_;
  }
  nav2_costmap_2d::LayeredCostmap *layered_costmap_;
  bool current_ = false, enabled_ = true;
  std::mutex cacheMutex_;
  std::map<int, std::vector<unsigned int>> latest_cells_to_mark_index_;
  int numAdditions_;
};
}  // namespace frontier_exploration
