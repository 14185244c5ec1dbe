// ref_harness.cpp -- C entry points over the REFERENCE's own frontier search and ray helpers, compiled from
// /root/reference where it lies (oracle/Makefile, target _ref/libfitslam_ref.so): DEP/src/Frontier.cpp,
// DEP/src/FrontierSearch.cpp and the costmap half of DEP/src/Helpers.cpp (lines 1-330: bresenham2D, getTracedCells,
// isRobotFootprintInLethal, nhood4 / nhood8, nearestFreeCell), against the stand-in headers of oracle/ref_shims/.
// Test infrastructure: it pins the restatement in oracle.cpp (tests/test_ref_pin.py) and nothing else links it.
#include <cstdint>
#include <cstring>
#include <vector>

#include <frontier_exploration/FrontierSearch.hpp>
#include <frontier_exploration/Helpers.hpp>

using frontier_exploration::FrontierSearch;
using nav2_costmap_2d::Costmap2D;

namespace {
Costmap2D make_map(const uint8_t *cells, uint32_t W, uint32_t H, double res, double ox, double oy) {
  Costmap2D m(W, H, res, ox, oy);
  std::memcpy(m.getCharMap(), cells, (size_t)W * H);
  return m;
}
}  // namespace

extern "C" {

// FrontierSearch::searchFrom + getAllFrontiers (DEP/src/FrontierSearch.cpp:21-96, 251-254).
// out_frontiers: capacity * 3 doubles (goal x, goal y, size); out_cells: cell_capacity * 2 doubles (x, y).
// Returns the number of frontiers (-1: the robot is off the map), *n_cells the number of every_frontier entries.
int ref_search_frontiers(const uint8_t *cells, uint32_t W, uint32_t H, double res, double ox, double oy, double min_size,
                         double max_size, double max_distance, int lethal_threshold, double robot_x, double robot_y,
                         double *out_frontiers, uint32_t capacity, double *out_cells, uint32_t cell_capacity, uint32_t *n_cells) {
  parameterInstance.set("frontierSearch/min_frontier_cluster_size", min_size);
  parameterInstance.set("frontierSearch/max_frontier_cluster_size", max_size);
  parameterInstance.set("frontierSearch/max_frontier_distance", max_distance);
  parameterInstance.set("frontierSearch/lethal_threshold", lethal_threshold);
  Costmap2D map = make_map(cells, W, H, res, ox, oy);
  unsigned int mx, my;
  if (!map.worldToMap(robot_x, robot_y, mx, my)) return -1;
  FrontierSearch search(map);
  search.reset();
  geometry_msgs::msg::Point p;
  p.x = robot_x;
  p.y = robot_y;
  std::vector<FrontierPtr> list = search.searchFrom(p);
  std::vector<std::vector<double>> every = search.getAllFrontiers();
  for (size_t i = 0; i < list.size() && i < capacity; ++i) {
    out_frontiers[3 * i] = list[i]->getGoalPoint().x;
    out_frontiers[3 * i + 1] = list[i]->getGoalPoint().y;
    out_frontiers[3 * i + 2] = list[i]->getSize();
  }
  for (size_t i = 0; i < every.size() && i < cell_capacity; ++i) {
    out_cells[2 * i] = every[i][0];
    out_cells[2 * i + 1] = every[i][1];
  }
  if (n_cells) *n_cells = (uint32_t)every.size();
  return (int)list.size();
}

// getTracedCells with a fresh RayTracedCells gatherer (DEP/src/Helpers.cpp:32-96, Helpers.hpp:20-111).
// out[0] = cells pushed, out[1] = hit_obstacle, out[2] = unknown cells, out[3] = all cells visited.  Returns 0 on failure.
int ref_traced_cells(const uint8_t *cells, uint32_t W, uint32_t H, double res, double ox, double oy, double sx, double sy, double wx,
                     double wy, double max_length, int obstacle_min, int obstacle_max, int trace_min, int trace_max, int32_t *out) {
  Costmap2D map = make_map(cells, W, H, res, ox, oy);
  std::vector<nav2_costmap_2d::MapLocation> traced;
  frontier_exploration::RayTracedCells gatherer(&map, traced, obstacle_min, obstacle_max, trace_min, trace_max);
  const bool ok = frontier_exploration::getTracedCells(sx, sy, wx, wy, gatherer, max_length, &map);
  out[0] = (int32_t)gatherer.getCells().size();
  out[1] = gatherer.hasHitObstacle() ? 1 : 0;
  out[2] = (int32_t)gatherer.getNumUnknown();
  out[3] = (int32_t)gatherer.getCellsSize();
  return ok ? 1 : 0;
}

// isRobotFootprintInLethal (DEP/src/Helpers.cpp:135-155); the caller keeps the disc inside the map (the reference does
// not check: SURVEY.md H5)
int ref_footprint_in_lethal(const uint8_t *cells, uint32_t W, uint32_t H, double res, double ox, double oy, uint32_t cx, uint32_t cy,
                            double radius_cells) {
  Costmap2D map = make_map(cells, W, H, res, ox, oy);
  return frontier_exploration::isRobotFootprintInLethal(&map, cx, cy, radius_cells) ? 1 : 0;
}

// nearestFreeCell (DEP/src/Helpers.cpp:285-330): returns 1 and *result, or 0
int ref_nearest_free_cell(const uint8_t *cells, uint32_t W, uint32_t H, uint32_t start, int val, uint32_t *result) {
  Costmap2D map = make_map(cells, W, H, 1.0, 0.0, 0.0);
  unsigned int r = 0;
  const bool ok = frontier_exploration::nearestFreeCell(r, start, (unsigned char)val, map);
  *result = r;
  return ok ? 1 : 0;
}

// nhood4 / nhood8 (DEP/src/Helpers.cpp:185-262): writes up to 8 indices, returns their number
int ref_nhood(uint32_t W, uint32_t H, uint32_t idx, int eight, uint32_t *out) {
  Costmap2D map(W, H, 1.0, 0.0, 0.0);
  const std::vector<unsigned int> v = eight ? frontier_exploration::nhood8(idx, map) : frontier_exploration::nhood4(idx, map);
  for (size_t i = 0; i < v.size() && i < 8; ++i) out[i] = v[i];
  return (int)v.size();
}

}  // extern "C"
