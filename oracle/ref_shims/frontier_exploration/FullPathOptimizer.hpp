// Stand-in for frontier_exploration/FullPathOptimizer.hpp: SortedFrontiers and the part of the class that
// FullPathOptimizer::getFilteredFrontiersN (DEP/src/FullPathOptimizer.cpp:157-227, compiled from the reference by oracle/Makefile)
// touches.  isInBlacklistedRegion has no regions to test here.
#pragma once
#include <algorithm>
#include <limits>
#include <vector>

#include <frontier_exploration/Frontier.hpp>
#include <frontier_exploration/util/logger.hpp>
#include <geometry_msgs/msg/pose.hpp>

namespace frontier_exploration {

const double LOCAL_FRONTIER_SEARCH_RADIUS = 12.0;  // FullPathOptimizer.hpp:16

struct SortedFrontiers {
  std::vector<FrontierPtr> local_frontiers;
  std::vector<FrontierPtr> global_frontiers;
  FrontierPtr closest_global_frontier;
};

class FullPathOptimizer {
public:
  void getFilteredFrontiersN(std::vector<FrontierPtr> &frontier_list, size_t n, SortedFrontiers &sortedFrontiers,
                             geometry_msgs::msg::PoseStamped &robotP);
  bool isInBlacklistedRegion(const FrontierPtr &) { return false; }
};

}  // namespace frontier_exploration
