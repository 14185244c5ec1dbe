// Stand-in for roadmap_explorer/planners/FrontierRoadmap.hpp: the members FrontierRoadMap::isConnectable touches
// (RE/src/planners/FrontierRoadmap.cpp:671-694, compiled from the reference by oracle/Makefile).
#pragma once
#include <nav2_costmap_2d/costmap_2d.hpp>
#include "roadmap_explorer/Frontier.hpp"
#include "roadmap_explorer/Helpers.hpp"
namespace roadmap_explorer {
class FrontierRoadMap {
public:
  FrontierRoadMap(nav2_costmap_2d::Costmap2D *costmap, double radius_to_decide_edges)
      : costmap_(costmap), max_connection_length_(radius_to_decide_edges * 1.5), RADIUS_TO_DECIDE_EDGES(radius_to_decide_edges) {}  // :22-32
  bool isConnectable(const FrontierPtr &f1, const FrontierPtr &f2);
private:
  nav2_costmap_2d::Costmap2D *costmap_;
  double max_connection_length_;
  double RADIUS_TO_DECIDE_EDGES;
};
}  // namespace roadmap_explorer
