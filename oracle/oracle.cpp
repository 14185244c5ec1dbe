// oracle.cpp -- CPU restatement of the FIT-SLAM exploration hot path.
//
// *** TEST INFRASTRUCTURE ONLY. ***  Nothing under oracle/ is part of the product: only tests/,
// __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load this library.
// The product (fit-slam_b200/csrc) never links, imports or calls it.
//
// *** PARITY: PINNED FOR STAGES S, R, N AND THE CLIENT HALF OF STAGE F; FOR THE ORB-SLAM3 WRAPPER'S HALF (F1, F2, KEYFRAME MODE) THE
// CONTROL FLOW IS PINNED, THE ARITHMETIC INSIDE SOPHUS / EIGEN'S AFFINE TRANSFORMS IS NOT. ***
// Pinned: the reference's own FrontierSearch.cpp, Frontier.cpp, the costmap half of Helpers.cpp, the member functions of
// CostCalculator.cpp (setArrivalInformationForFrontier, setMaxArrivalInformation, setPlanForFrontierEuclidean,
// recomputeNormalizationFactors), FrontierCostsManager.cpp (assignCosts, setFrontierBlacklist), FullPathOptimizer.cpp
// (getFilteredFrontiersN), FisherInfoManager.cpp (whole) and the Fisher helpers of Helpers.cpp it calls compile from where they lie
// against stand-in ROS / pcl / Eigen headers (oracle/Makefile -> oracle/_ref/libfitslam_ref.so, oracle/ref_shims/), and
// tests/test_ref_pin.py holds this file to that library bit for bit: rows S1-S6, R1-R6, T4, N1-N3, F3-F7 of SURVEY.md 8a; a second
// library (oracle/ref_harness_re.cpp) does the same for FrontierRoadMap::isConnectable and SensorSimulator::markRay of the
// roadmap-explorer package (rows 8f-2, 8f-4); lethal_marker.cpp (rays, circle, triangle, addNewMarkedArea / FOV, updateCosts) and the
// threshold node's callback (row 8f-3) are in the first library.  (Eigen
// is not in the image: the fixed-size float algebra under the Fisher helpers is ref_shims/Eigen/Core's restatement of Eigen 3.4's
// evaluation order, the same reading as this file's.)
// The wrapper's half: the reference ships no golden vectors, known-answer tests or fixtures for this path (SURVEY.md section 4 /
// 8c), and the ORB-SLAM3 wrapper (mapPointsVisibleFromPose, the service handler, se3ROSToORB) needs Sophus, ORB_SLAM3's classes
// and ROS 2 nodes, none of which exist in this image.  A third library (oracle/ref_harness_orbw.cpp ->
// oracle/_ref/libfitslam_ref_orbw.so) compiles those function bodies against stand-in declarations (oracle/ref_shims_orbw/: an
// atlas the harness fills, Sophus::SE3f as (R, t)) and tests/test_ref_pin.py holds landmarksInView / landmarksInViewKeyframes to
// it coordinate for coordinate: the control flow and the order of operations are the reference's, Sophus's quaternion storage
// and Eigen's Affine rotation() are not modelled.  The only reference test that applies (ORBW/tests/typeConversionsTests.cpp:8-33, SE3 ROS<->ORB round trip) is reproduced against
// orc_se3_roundtrip() in tests/test_oracle_micro.py, next to the hand-derivable micro-cases of SURVEY.md 8c.
//
// Path shorthand (all under /root/reference):
//   DEP/  = dev_ws/src/DEPRECATED/frontier_exploration/frontier_exploration/
//   ORBW/ = ORB-SLAM3-ROS2-Docker/orb_slam3_ros2_wrapper/
//   ORB/  = ORB-SLAM3-ROS2-Docker/ORB_SLAM3/
//
// Third-party arithmetic that is not under /root/reference is restated from its published behaviour:
//   nav2_costmap_2d (ROS 2 Humble) Costmap2D::{worldToMap,mapToWorld,getIndex,indexToCells,
//   getSizeInMetersX}; Eigen 3.4 fixed-size float products (coefficient-based lazy product, redux
//   order x0 + (x1 + x2) for 3-vectors), Quaternion::toRotationMatrix, Matrix3f::inverse (cofactors);
//   tf2 orientationAroundZAxis.  glibc libm and libstdc++ std::sort / std::hash are used directly.
//
// Build: g++ -O2 -std=c++17 -ffp-contract=off -fPIC -shared  (see oracle/Makefile).

#include <algorithm>
#include <array>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <limits>
#include <queue>
#include <string>
#include <unordered_map>
#include <utility>
#include <map>
#include <vector>

#include "../include/fitslam_cuda.h"

namespace {

// ---------------------------------------------------------------------------------------------
// nav2_costmap_2d::Costmap2D (third party, ROS 2 Humble costmap_2d.cpp) -- only what the path calls.
// Call sites: DEP/src/FrontierSearch.cpp:28-36,72-73 ; DEP/src/Helpers.cpp:40,67 ;
// DEP/src/CostCalculator.cpp:47-48,76.
// ---------------------------------------------------------------------------------------------
struct Costmap2D {
  const unsigned char *costmap_;
  unsigned int size_x_, size_y_;
  double resolution_, origin_x_, origin_y_;

  const unsigned char *getCharMap() const { return costmap_; }
  unsigned int getSizeInCellsX() const { return size_x_; }
  unsigned int getSizeInCellsY() const { return size_y_; }
  double getResolution() const { return resolution_; }
  double getOriginX() const { return origin_x_; }
  double getOriginY() const { return origin_y_; }
  double getSizeInMetersX() const { return (size_x_ - 1 + 0.5) * resolution_; }
  double getSizeInMetersY() const { return (size_y_ - 1 + 0.5) * resolution_; }
  unsigned int getIndex(unsigned int mx, unsigned int my) const { return my * size_x_ + mx; }
  void indexToCells(unsigned int index, unsigned int &mx, unsigned int &my) const {
    my = index / size_x_;
    mx = index - (my * size_x_);
  }
  void mapToWorld(unsigned int mx, unsigned int my, double &wx, double &wy) const {
    wx = origin_x_ + (mx + 0.5) * resolution_;
    wy = origin_y_ + (my + 0.5) * resolution_;
  }
  bool worldToMap(double wx, double wy, unsigned int &mx, unsigned int &my) const {
    if (wx < origin_x_ || wy < origin_y_) return false;
    mx = static_cast<unsigned int>((wx - origin_x_) / resolution_);
    my = static_cast<unsigned int>((wy - origin_y_) / resolution_);
    if (mx < size_x_ && my < size_y_) return true;
    return false;
  }
  unsigned char getCost(unsigned int index) const { return costmap_[index]; }
  unsigned char getCost(unsigned int mx, unsigned int my) const { return costmap_[getIndex(mx, my)]; }
};

constexpr unsigned char NO_INFORMATION = 255;
constexpr unsigned char LETHAL_OBSTACLE = 254;

struct MapLocation {
  unsigned int x, y;
};

// ---------------------------------------------------------------------------------------------
// DEP/src/Helpers.cpp:185-255  nhood4 / nhood8 (neighbour ORDER matters: -1, +1, -W, +W, then the
// four diagonals -1-W, -1+W, +1-W, +1+W)
// ---------------------------------------------------------------------------------------------
inline int nhood4(unsigned int idx, const Costmap2D &costmap, unsigned int out[8]) {
  int n = 0;
  unsigned int size_x_ = costmap.getSizeInCellsX();
  unsigned int size_y_ = costmap.getSizeInCellsY();
  if (idx > size_x_ * size_y_ - 1) return 0;
  if (idx % size_x_ > 0) out[n++] = idx - 1;
  if (idx % size_x_ < size_x_ - 1) out[n++] = idx + 1;
  if (idx >= size_x_) out[n++] = idx - size_x_;
  if (idx < size_x_ * (size_y_ - 1)) out[n++] = idx + size_x_;
  return n;
}

inline int nhood8(unsigned int idx, const Costmap2D &costmap, unsigned int out[8]) {
  int n = nhood4(idx, costmap, out);
  unsigned int size_x_ = costmap.getSizeInCellsX();
  unsigned int size_y_ = costmap.getSizeInCellsY();
  if (idx > size_x_ * size_y_ - 1) return n;
  if (idx % size_x_ > 0 && idx >= size_x_) out[n++] = idx - 1 - size_x_;
  if (idx % size_x_ > 0 && idx < size_x_ * (size_y_ - 1)) out[n++] = idx - 1 + size_x_;
  if (idx % size_x_ < size_x_ - 1 && idx >= size_x_) out[n++] = idx + 1 - size_x_;
  if (idx % size_x_ < size_x_ - 1 && idx < size_x_ * (size_y_ - 1)) out[n++] = idx + 1 + size_x_;
  return n;
}

// DEP/src/Helpers.cpp:285-330  nearestFreeCell
bool nearestFreeCell(unsigned int &result, unsigned int start, unsigned char val, const Costmap2D &costmap) {
  const unsigned char *map = costmap.getCharMap();
  const unsigned int size_x = costmap.getSizeInCellsX();
  const unsigned int size_y = costmap.getSizeInCellsY();
  if (start >= size_x * size_y) return false;
  std::queue<unsigned int> bfs;
  std::vector<bool> visited_flag(size_x * size_y, false);
  bfs.push(start);
  visited_flag[start] = true;
  while (!bfs.empty()) {
    unsigned int idx = bfs.front();
    bfs.pop();
    if (map[idx] < val) {
      result = idx;
      return true;
    }
    unsigned int nb[8];
    int n = nhood8(idx, costmap, nb);
    for (int i = 0; i < n; ++i) {
      unsigned nbr = nb[i];
      if (!visited_flag[nbr]) {
        bfs.push(nbr);
        visited_flag[nbr] = true;
      }
    }
  }
  return false;
}

// DEP/include/frontier_exploration/util/GeometryUtils.hpp:102-105
inline double distanceBetweenPoints(double f1x, double f1y, const double x2, const double y2) {
  return sqrt(pow(f1x - x2, 2) + pow(f1y - y2, 2));
}

// ---------------------------------------------------------------------------------------------
// DEP/include/frontier_exploration/FrontierSearch.hpp:156-184  SortByMedianFunctor
// (elements carry the map index along; the comparator looks at the coordinates only)
// ---------------------------------------------------------------------------------------------
struct CellXY {
  double first, second;
  unsigned int idx;
};

class SortByMedianFunctor {
 public:
  explicit SortByMedianFunctor(std::pair<double, double> centroid) : centroid(centroid) {}
  bool operator()(const CellXY &a, const CellXY &b) const {
    auto angle_a = atan2(a.second - centroid.second, a.first - centroid.first);
    if (angle_a < 0) angle_a = angle_a + (2 * M_PI);
    auto angle_b = atan2(b.second - centroid.second, b.first - centroid.first);
    if (angle_b < 0) angle_b = angle_b + (2 * M_PI);
    if (0 <= angle_a && angle_a <= M_PI / 2 && 3 * M_PI / 2 <= angle_b && angle_b <= 2 * M_PI) return false;
    if (0 <= angle_b && angle_b <= M_PI / 2 && 3 * M_PI / 2 <= angle_a && angle_a <= 2 * M_PI) return true;
    return angle_a < angle_b;
  }

 private:
  std::pair<double, double> centroid;
};

struct FrontierOut {
  double gx, gy;
  unsigned int gidx;
  int size;
  unsigned int component, chunk, first_cell;
  double centroid_x, centroid_y;  // after the offset rule (stage dump)
};

// DEP/src/FrontierSearch.cpp (all) + FrontierSearch.hpp:85-142
class FrontierSearch {
 public:
  FrontierSearch(const Costmap2D &costmap, const fitslam_params &p) : costmap_(costmap) {
    min_frontier_cluster_size_ = p.min_frontier_cluster_size;
    max_frontier_cluster_size_ = p.max_frontier_cluster_size;
    max_frontier_distance_ = p.max_frontier_distance;
    lethal_threshold_ = (unsigned char)p.lethal_threshold;
  }

  std::vector<unsigned int> every_frontier_cells;  // getAllFrontiers(), as map indices
  std::vector<unsigned int> component_seed;        // stage dump: seed cell of every component
  unsigned int start_cell = 0;
  bool robot_out_of_bounds = false;

  // DEP/src/FrontierSearch.cpp:21-96
  std::vector<FrontierOut> searchFrom(double position_x, double position_y) {
    std::vector<FrontierOut> frontier_list;
    unsigned int mx, my;
    if (!costmap_.worldToMap(position_x, position_y, mx, my)) {
      robot_out_of_bounds = true;
      return frontier_list;
    }
    map_ = costmap_.getCharMap();
    size_x_ = costmap_.getSizeInCellsX();
    size_y_ = costmap_.getSizeInCellsY();
    std::vector<bool> frontier_flag(size_x_ * size_y_, false);
    std::vector<bool> visited_flag(size_x_ * size_y_, false);
    std::queue<unsigned int> bfs;
    unsigned int clear, pos = costmap_.getIndex(mx, my);
    if (nearestFreeCell(clear, pos, lethal_threshold_, costmap_)) {
      bfs.push(clear);
    } else {
      bfs.push(pos);
    }
    start_cell = bfs.front();
    visited_flag[bfs.front()] = true;

    while (!bfs.empty()) {
      unsigned int idx = bfs.front();
      bfs.pop();
      unsigned int nb[8];
      int n = nhood4(idx, costmap_, nb);
      for (int i = 0; i < n; ++i) {
        unsigned nbr = nb[i];
        if (map_[nbr] < LETHAL_OBSTACLE && !visited_flag[nbr]) {
          visited_flag[nbr] = true;
          unsigned int nbr_mx, nbr_my;
          double nbr_wx, nbr_wy;
          costmap_.indexToCells(nbr, nbr_mx, nbr_my);
          costmap_.mapToWorld(nbr_mx, nbr_my, nbr_wx, nbr_wy);
          if (distanceBetweenPoints(position_x, position_y, nbr_wx, nbr_wy) <
              max_frontier_distance_ + (max_frontier_cluster_size_ * costmap_.getResolution() * 1.414))
            bfs.push(nbr);
        } else if (isNewFrontierCell(nbr, frontier_flag)) {
          frontier_flag[nbr] = true;
          std::vector<FrontierOut> new_frontier = buildNewFrontier(nbr, pos, frontier_flag);
          for (auto &curr_frontier : new_frontier) {
            if (curr_frontier.size > min_frontier_cluster_size_) {
              frontier_list.push_back(curr_frontier);
            }
          }
        }
      }
    }
    return frontier_list;
  }

  // DEP/src/FrontierSearch.cpp:218-249 against the static map only (stage dump of the frontier mask)
  bool isFrontierMaskCell(unsigned int idx) {
    std::vector<bool> none;
    return isNewFrontierCellImpl(idx, nullptr);
  }
  void bindMap() {
    map_ = costmap_.getCharMap();
    size_x_ = costmap_.getSizeInCellsX();
    size_y_ = costmap_.getSizeInCellsY();
  }

 private:
  // DEP/src/FrontierSearch.cpp:98-216 (FRONTIER_POINT_MEDIAN defined: DEP/CMakeLists.txt:63-64)
  std::vector<FrontierOut> buildNewFrontier(unsigned int initial_cell, unsigned int /*reference*/,
                                            std::vector<bool> &frontier_flag) {
    int currentFrontierSize = 1;
    std::vector<FrontierOut> calculated_frontiers;
    std::vector<CellXY> frontier_cell_indices;
    unsigned int ix, iy;
    costmap_.indexToCells(initial_cell, ix, iy);
    double wix, wiy;
    costmap_.mapToWorld(ix, iy, wix, wiy);
    unsigned int component = (unsigned int)component_seed.size();
    component_seed.push_back(initial_cell);
    unsigned int chunk = 0;
    unsigned int first_cell = (unsigned int)every_frontier_cells.size();
    every_frontier_cells.push_back(initial_cell);
    frontier_cell_indices.push_back({wix, wiy, initial_cell});

    std::queue<unsigned int> bfs;
    bfs.push(initial_cell);

    while (!bfs.empty()) {
      unsigned int idx = bfs.front();
      bfs.pop();
      unsigned int nb[8];
      int n = nhood8(idx, costmap_, nb);
      for (int i = 0; i < n; ++i) {
        unsigned int nbr = nb[i];
        if (isNewFrontierCell(nbr, frontier_flag)) {
          frontier_flag[nbr] = true;
          unsigned int mx, my;
          double wx, wy;
          costmap_.indexToCells(nbr, mx, my);
          costmap_.mapToWorld(mx, my, wx, wy);
          every_frontier_cells.push_back(nbr);
          frontier_cell_indices.push_back({wx, wy, nbr});
          currentFrontierSize = currentFrontierSize + 1;
          bfs.push(nbr);
          if (currentFrontierSize > max_frontier_cluster_size_) {
            FrontierOut output = makeFrontier(frontier_cell_indices, currentFrontierSize);
            output.component = component;
            output.chunk = chunk++;
            output.first_cell = first_cell;
            first_cell = (unsigned int)every_frontier_cells.size();
            frontier_cell_indices.clear();
            calculated_frontiers.push_back(output);
            currentFrontierSize = 0;
          }
        }
      }
    }
    if (currentFrontierSize > min_frontier_cluster_size_) {
      FrontierOut output = makeFrontier(frontier_cell_indices, currentFrontierSize);
      output.component = component;
      output.chunk = chunk++;
      output.first_cell = first_cell;
      frontier_cell_indices.clear();
      calculated_frontiers.push_back(output);
    }
    return calculated_frontiers;
  }

  // DEP/src/FrontierSearch.cpp:161-165 / 194-199
  FrontierOut makeFrontier(std::vector<CellXY> &frontier_cell_indices, int currentFrontierSize) {
    FrontierOut output{};
    auto cluster_centroid = getCentroidOfCells(frontier_cell_indices, (costmap_.getResolution() * 1.414 * 2));
    SortByMedianFunctor sortFunctor(cluster_centroid);
    std::sort(frontier_cell_indices.begin(), frontier_cell_indices.end(), sortFunctor);
    auto goal_point = frontier_cell_indices[static_cast<int>(frontier_cell_indices.size() / 2)];
    output.gx = goal_point.first;
    output.gy = goal_point.second;
    output.gidx = goal_point.idx;
    output.size = currentFrontierSize;
    output.centroid_x = cluster_centroid.first;
    output.centroid_y = cluster_centroid.second;
    return output;
  }

  // DEP/include/frontier_exploration/FrontierSearch.hpp:85-126
  std::pair<double, double> getCentroidOfCells(std::vector<CellXY> &cells, double distance_to_offset) {
    double sumX = 0;
    double sumY = 0;
    for (const auto &point : cells) {
      sumX += point.first;
      sumY += point.second;
    }
    double centerX = static_cast<double>(sumX) / cells.size();
    double centerY = static_cast<double>(sumY) / cells.size();
    bool offset_centroid = false;
    double varX = 0, varY = 0;
    for (const auto &point : cells) {
      if (sqrt(pow(point.first - centerX, 2) + pow(point.second - centerY, 2)) < costmap_.getResolution() * 3) {
        offset_centroid = true;
      }
      // the reference writes unqualified abs(); with <cstdlib>/<cmath> in scope it is the double overload
      varX += std::abs(point.first - centerX);
      varY += std::abs(point.second - centerY);
    }
    if (varX > varY && offset_centroid) centerY -= distance_to_offset;
    if (varX < varY && offset_centroid) centerX -= distance_to_offset;
    return std::make_pair(centerX, centerY);
  }

  // DEP/src/FrontierSearch.cpp:218-249
  bool isNewFrontierCell(unsigned int idx, const std::vector<bool> &frontier_flag) {
    return isNewFrontierCellImpl(idx, &frontier_flag);
  }
  bool isNewFrontierCellImpl(unsigned int idx, const std::vector<bool> *frontier_flag) {
    if (!isUnknown(map_[idx]) || (frontier_flag && (*frontier_flag)[idx])) return false;
    bool has_one_free_neighbour = false;
    bool has_one_lethal_neighbour = false;
    unsigned int nb[8];
    int n = nhood4(idx, costmap_, nb);
    for (int i = 0; i < n; ++i) {
      unsigned int nbr = nb[i];
      if (isFree(map_[nbr])) has_one_free_neighbour = true;
      if (isLethal(map_[nbr])) has_one_lethal_neighbour = true;
    }
    if (has_one_lethal_neighbour)
      return false;
    else if (has_one_free_neighbour)
      return true;
    else
      return false;
  }

  // DEP/include/frontier_exploration/FrontierSearch.hpp:129-142
  inline bool isLethal(unsigned char value) { return (int)value >= lethal_threshold_ && value != NO_INFORMATION; }
  inline bool isUnknown(unsigned char value) { return value == NO_INFORMATION; }
  inline bool isFree(unsigned char value) { return (int)value < lethal_threshold_; }

  const Costmap2D &costmap_;
  const unsigned char *map_ = nullptr;
  unsigned int size_x_ = 0, size_y_ = 0;
  int min_frontier_cluster_size_;
  int max_frontier_cluster_size_;
  double max_frontier_distance_;
  unsigned char lethal_threshold_;
};

// ---------------------------------------------------------------------------------------------
// DEP/include/frontier_exploration/Helpers.hpp:20-116  RayTracedCells, sign
// ---------------------------------------------------------------------------------------------
class RayTracedCells {
 public:
  RayTracedCells(const Costmap2D *costmap, std::vector<MapLocation> &cells, int obstacle_min, int obstacle_max,
                 int trace_min, int trace_max)
      : costmap_(costmap), cells_(cells), obstacle_min_(obstacle_min), obstacle_max_(obstacle_max),
        trace_min_(trace_min), trace_max_(trace_max) {
    hit_obstacle = false;
    unknown_cells_ = 0;
    all_cells_count_ = 0;
  }
  inline void operator()(unsigned int offset) {
    MapLocation loc;
    costmap_->indexToCells(offset, loc.x, loc.y);
    bool presentflag = false;
    for (auto item : cells_) {
      if (item.x == loc.x && item.y == loc.y) presentflag = true;
    }
    if (presentflag == false) {
      ++all_cells_count_;
      auto cost = (int)costmap_->getCost(offset);
      if (cost <= trace_max_ && cost >= trace_min_ && !hit_obstacle) cells_.push_back(loc);
      if (cost >= obstacle_min_ && cost <= obstacle_max_) hit_obstacle = true;
      if (cost == 255) unknown_cells_++;
    }
  }
  std::vector<MapLocation> getCells() { return cells_; }
  bool hasHitObstacle() { return hit_obstacle; }
  size_t getNumUnknown() { return unknown_cells_; }  // RE/include/roadmap_explorer/Helpers.hpp:77-80

 private:
  const Costmap2D *costmap_;
  std::vector<MapLocation> &cells_;
  bool hit_obstacle;
  int obstacle_min_, obstacle_max_;
  int trace_min_, trace_max_;
  int unknown_cells_ = 0;
  int all_cells_count_ = 0;
};

inline int sign(int x) { return x > 0 ? 1.0 : -1.0; }

// DEP/src/Helpers.cpp:7-30
void bresenham2D(RayTracedCells &at, unsigned int abs_da, unsigned int abs_db, int error_b, int offset_a,
                 int offset_b, unsigned int offset, unsigned int max_length, int resolution_cut_factor,
                 const Costmap2D * /*exploration_costmap_*/) {
  unsigned int end = std::min(max_length, abs_da);
  for (unsigned int i = 0; i < end; ++i) {
    if (i % resolution_cut_factor == 0) at(offset);
    offset += offset_a;
    error_b += abs_db;
    if ((unsigned int)error_b >= abs_da) {
      offset += offset_b;
      error_b -= abs_da;
    }
  }
  at(offset);
}

// DEP/src/Helpers.cpp:32-96
bool getTracedCells(double sx, double sy, double wx, double wy, RayTracedCells &cell_gatherer, double max_length,
                    const Costmap2D *exploration_costmap_) {
  unsigned int min_length = 0.0;
  int resolution_cut_factor = 1;
  unsigned int x1, y1;
  unsigned int x0, y0;
  if (!exploration_costmap_->worldToMap(wx, wy, x1, y1) || !exploration_costmap_->worldToMap(sx, sy, x0, y0)) {
    return false;
  }
  int dx_full = x1 - x0;
  int dy_full = y1 - y0;
  double dist = std::hypot(dx_full, dy_full);
  if (dist < min_length) return false;
  unsigned int min_x0, min_y0;
  if (dist > 0.0) {
    min_x0 = (unsigned int)(x0 + dx_full / dist * min_length);
    min_y0 = (unsigned int)(y0 + dy_full / dist * min_length);
  } else {
    min_x0 = x0;
    min_y0 = y0;
  }
  unsigned int offset = min_y0 * exploration_costmap_->getSizeInCellsX() + min_x0;
  int dx = x1 - min_x0;
  int dy = y1 - min_y0;
  unsigned int abs_dx = abs(dx);
  unsigned int abs_dy = abs(dy);
  int offset_dx = sign(dx);
  int offset_dy = sign(dy) * exploration_costmap_->getSizeInCellsX();
  double scale = (dist == 0.0) ? 1.0 : std::min(1.0, max_length / dist);
  if (abs_dx >= abs_dy) {
    int error_y = abs_dx / 2;
    bresenham2D(cell_gatherer, abs_dx, abs_dy, error_y, offset_dx, offset_dy, offset, (unsigned int)(scale * abs_dx),
                resolution_cut_factor, exploration_costmap_);
  } else {
    int error_x = abs_dy / 2;
    bresenham2D(cell_gatherer, abs_dy, abs_dx, error_x, offset_dy, offset_dx, offset, (unsigned int)(scale * abs_dy),
                resolution_cut_factor, exploration_costmap_);
  }
  return true;
}

// DEP/src/Helpers.cpp:135-155, with the bounds check the live tree adds (RE/src/Helpers.cpp:126-149):
// DEP reads out of bounds when the disc leaves the map, which is undefined; SURVEY.md H5.
bool isRobotFootprintInLethal(const Costmap2D *costmap, unsigned int center_x, unsigned int center_y,
                              double radius_in_cells) {
  for (int dx = -radius_in_cells; dx <= radius_in_cells; ++dx) {
    for (int dy = -radius_in_cells; dy <= radius_in_cells; ++dy) {
      if (dx * dx + dy * dy <= radius_in_cells * radius_in_cells) {
        long x = (long)center_x + dx;
        long y = (long)center_y + dy;
        if (x < 0 || y < 0 || x >= (long)costmap->getSizeInCellsX() || y >= (long)costmap->getSizeInCellsY()) continue;
        unsigned int cost = costmap->getCost((unsigned int)x, (unsigned int)y);
        if (cost == 254) return true;
      }
    }
  }
  return false;
}

// ---------------------------------------------------------------------------------------------
// DEP/src/CostCalculator.cpp:23-191  FrontierCostCalculator (arrival information half)
// ---------------------------------------------------------------------------------------------
struct FrontierCostCalculator {
  const Costmap2D *exploration_costmap_;
  double MAX_CAMERA_DEPTH, DELTA_THETA, CAMERA_FOV, robot_radius_;
  double max_arrival_factor, min_arrival_factor;
  double max_arrival_info_gt_ = 0.0, min_arrival_info_gt_ = 0.0;
  bool arrival_info_limits_set_ = false;

  // :23-121.  Returns the record; info_ray / result are the reference's intermediate vectors.
  void setArrivalInformationForFrontier(double goal_x, double goal_y, int frontier_size, const double *polygon_xy_min_max,
                                        fitslam_arrival &out, std::vector<int> &information_along_ray,
                                        std::vector<int> &result) {
    std::memset(&out, 0, sizeof(out));
    out.achievable = 1;  // Frontier() constructor: is_achievable = true (DEP/src/Frontier.cpp:17)
    double sx, sy;
    double wx, wy;
    unsigned int max_length = MAX_CAMERA_DEPTH / (exploration_costmap_->getResolution());
    sx = goal_x;
    sy = goal_y;
    information_along_ray.clear();
    result.clear();
    for (double theta = 0; theta <= (2 * M_PI); theta += DELTA_THETA) {
      std::vector<MapLocation> traced_cells;
      RayTracedCells cell_gatherer(exploration_costmap_, traced_cells, 240, 254, 255, 255);
      wx = sx + (MAX_CAMERA_DEPTH * cos(theta));
      wy = sy + (MAX_CAMERA_DEPTH * sin(theta));
      wx = std::max(polygon_xy_min_max[0],
                    std::max(exploration_costmap_->getOriginX(),
                             std::min(polygon_xy_min_max[2],
                                      std::min(exploration_costmap_->getOriginX() + exploration_costmap_->getSizeInMetersX(), wx))));
      wy = std::max(polygon_xy_min_max[1],
                    std::max(exploration_costmap_->getOriginY(),
                             std::min(polygon_xy_min_max[3],
                                      std::min(exploration_costmap_->getOriginY() + exploration_costmap_->getSizeInMetersY(), wy))));
      if (!getTracedCells(sx, sy, wx, wy, cell_gatherer, max_length, exploration_costmap_)) {
        out.information = 0;   // setArrivalInformation(0.0)
        out.theta_s_star = 0;  // setGoalOrientation(0.0)
        out.max_index = 0;
        out.ray_failed = 1;
        information_along_ray.clear();
        return;
      }
      auto info_addition = cell_gatherer.getCells();
      information_along_ray.push_back(info_addition.size());
    }
    unsigned int sxm, sym;
    exploration_costmap_->worldToMap(sx, sy, sxm, sym);
    double footprintInLethalPenalty = isRobotFootprintInLethal(
        exploration_costmap_, sxm, sym, std::ceil(robot_radius_ / exploration_costmap_->getResolution()));
    out.footprint_in_lethal = footprintInLethalPenalty != 0.0;
    if (1.0 - footprintInLethalPenalty == 0.0 && frontier_size < 10.0) out.achievable = 0;

    std::vector<int> kernel(static_cast<int>(CAMERA_FOV / DELTA_THETA), 1);
    int n = information_along_ray.size();
    int k = kernel.size();
    result.assign(n - k + 1, 0);
    for (int i = 0; i < n - k + 1; ++i)
      for (int j = 0; j < k; ++j) result[i] += information_along_ray[i + j] * kernel[j];
    int maxIndex = 0;
    int maxValue = result[0];
    for (int i = 1; i < (int)result.size(); ++i) {
      if (result[i] > maxValue) {
        maxValue = result[i];
        maxIndex = i;
      }
    }
    out.information = maxValue;
    out.max_index = maxIndex;
    if ((double)maxValue < min_arrival_info_gt_) out.achievable = 0;
    out.theta_s_star = (maxIndex * DELTA_THETA) + (CAMERA_FOV / 2);
  }

  // :123-191
  double setMaxArrivalInformation() {
    if (arrival_info_limits_set_) return 0.0;
    double sx, sy;
    double wx, wy;
    unsigned int max_length = MAX_CAMERA_DEPTH / (exploration_costmap_->getResolution());
    sx = 0.0;
    sy = 0.0;
    std::vector<int> information_along_ray;
    for (double theta = 0; theta <= (2 * M_PI); theta += DELTA_THETA) {
      std::vector<MapLocation> traced_cells;
      RayTracedCells cell_gatherer(exploration_costmap_, traced_cells, 260, 260, 0, 255);
      wx = sx + (MAX_CAMERA_DEPTH * cos(theta));
      wy = sy + (MAX_CAMERA_DEPTH * sin(theta));
      if (!getTracedCells(sx, sy, wx, wy, cell_gatherer, max_length, exploration_costmap_)) return 0;
      auto info_addition = cell_gatherer.getCells();
      information_along_ray.push_back(info_addition.size());
    }
    std::vector<int> kernel(static_cast<int>(CAMERA_FOV / DELTA_THETA), 1);
    int n = information_along_ray.size();
    int k = kernel.size();
    std::vector<int> result(n - k + 1, 0);
    for (int i = 0; i < n - k + 1; ++i)
      for (int j = 0; j < k; ++j) result[i] += information_along_ray[i + j] * kernel[j];
    int maxValue = result[0];
    for (int i = 1; i < (int)result.size(); ++i)
      if (result[i] > maxValue) maxValue = result[i];
    arrival_info_limits_set_ = true;
    max_arrival_info_gt_ = maxValue * max_arrival_factor;
    min_arrival_info_gt_ = min_arrival_factor * max_arrival_info_gt_;
    return maxValue;
  }
};

FrontierCostCalculator makeCalc(const Costmap2D *cm, const fitslam_params *p) {
  FrontierCostCalculator c;
  c.exploration_costmap_ = cm;
  c.MAX_CAMERA_DEPTH = p->max_camera_depth;
  c.DELTA_THETA = p->delta_theta;
  c.CAMERA_FOV = p->camera_fov;
  c.robot_radius_ = p->robot_radius;
  c.max_arrival_factor = p->max_arrival_factor;
  c.min_arrival_factor = p->min_arrival_factor;
  return c;
}

Costmap2D makeMap(const uint8_t *map, uint32_t W, uint32_t H, double res, double ox, double oy) {
  Costmap2D cm;
  cm.costmap_ = map;
  cm.size_x_ = W;
  cm.size_y_ = H;
  cm.resolution_ = res;
  cm.origin_x_ = ox;
  cm.origin_y_ = oy;
  return cm;
}

// ---------------------------------------------------------------------------------------------
// Fixed-size float algebra in Eigen 3.4's evaluation order (Eigen itself is absent from this image).
// 3-term reductions follow redux_novec_unroller: x0 + (x1 + x2); 6-term: (x0+(x1+x2)) + (x3+(x4+x5)).
// ---------------------------------------------------------------------------------------------
struct V3 {
  float v[3];
};
struct M3 {
  float m[3][3];
};  // m[row][col]
inline float dot3(float a0, float b0, float a1, float b1, float a2, float b2) { return a0 * b0 + (a1 * b1 + a2 * b2); }
inline V3 mul(const M3 &A, const V3 &x) {
  V3 r;
  for (int i = 0; i < 3; ++i) r.v[i] = dot3(A.m[i][0], x.v[0], A.m[i][1], x.v[1], A.m[i][2], x.v[2]);
  return r;
}
inline M3 mul(const M3 &A, const M3 &B) {
  M3 r;
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) r.m[i][j] = dot3(A.m[i][0], B.m[0][j], A.m[i][1], B.m[1][j], A.m[i][2], B.m[2][j]);
  return r;
}
inline M3 transpose(const M3 &A) {
  M3 r;
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) r.m[i][j] = A.m[j][i];
  return r;
}
// Eigen::Quaternionf::toRotationMatrix (Eigen/src/Geometry/Quaternion.h)
inline M3 quatToMatrix(float w, float x, float y, float z) {
  M3 res;
  const float tx = 2.0f * x, ty = 2.0f * y, tz = 2.0f * z;
  const float twx = tx * w, twy = ty * w, twz = tz * w;
  const float txx = tx * x, txy = ty * x, txz = tz * x;
  const float tyy = ty * y, tyz = tz * y, tzz = tz * z;
  res.m[0][0] = 1.0f - (tyy + tzz);
  res.m[0][1] = txy - twz;
  res.m[0][2] = txz + twy;
  res.m[1][0] = txy + twz;
  res.m[1][1] = 1.0f - (txx + tzz);
  res.m[1][2] = tyz - twx;
  res.m[2][0] = txz - twy;
  res.m[2][1] = tyz + twx;
  res.m[2][2] = 1.0f - (txx + tyy);
  return res;
}
// Eigen compute_inverse<Matrix3f>: cofactor_3x3<i,j>, det from column 0, multiply by 1/det
inline float cofactor(const M3 &m, int i, int j) {
  int i1 = (i + 1) % 3, i2 = (i + 2) % 3, j1 = (j + 1) % 3, j2 = (j + 2) % 3;
  return m.m[i1][j1] * m.m[i2][j2] - m.m[i1][j2] * m.m[i2][j1];
}
inline M3 inverse3(const M3 &m) {
  float c0 = cofactor(m, 0, 0), c1 = cofactor(m, 1, 0), c2 = cofactor(m, 2, 0);
  float det = dot3(c0, m.m[0][0], c1, m.m[1][0], c2, m.m[2][0]);
  float invdet = 1.0f / det;
  M3 r;
  r.m[0][0] = c0 * invdet;
  r.m[0][1] = c1 * invdet;
  r.m[0][2] = c2 * invdet;
  r.m[1][0] = cofactor(m, 0, 1) * invdet;
  r.m[1][1] = cofactor(m, 1, 1) * invdet;
  r.m[1][2] = cofactor(m, 2, 1) * invdet;
  r.m[2][0] = cofactor(m, 0, 2) * invdet;
  r.m[2][1] = cofactor(m, 1, 2) * invdet;
  r.m[2][2] = cofactor(m, 2, 2) * invdet;
  return r;
}
struct Affine {
  M3 R;
  V3 t;
};
// ORBW/src/type_conversion.cpp:141-153  poseToAffine: translation * quaternion, doubles cast to float
inline Affine poseToAffine(const fitslam_pose &pose) {
  Affine a;
  a.t.v[0] = static_cast<float>(pose.x);
  a.t.v[1] = static_cast<float>(pose.y);
  a.t.v[2] = static_cast<float>(pose.z);
  a.R = quatToMatrix(static_cast<float>(pose.qw), static_cast<float>(pose.qx), static_cast<float>(pose.qy),
                     static_cast<float>(pose.qz));
  return a;
}
// Eigen Transform<float,3,Affine>::inverse(): linear().inverse(), translation = -linear_inv * t
inline Affine inverseAffine(const Affine &a) {
  Affine r;
  r.R = inverse3(a.R);
  M3 neg;
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) neg.m[i][j] = -r.R.m[i][j];
  r.t = mul(neg, a.t);
  return r;
}
inline V3 apply(const Affine &a, const V3 &p) {
  V3 lp = mul(a.R, p);
  V3 r;
  for (int i = 0; i < 3; ++i) r.v[i] = lp.v[i] + a.t.v[i];
  return r;
}
const M3 tfROSToORB = {{{0, -1, 0}, {0, 0, -1}, {1, 0, 0}}};
const M3 tfORBToROS = {{{0, 0, 1}, {-1, 0, 0}, {0, -1, 0}}};

// ORBW/src/type_conversion.cpp:74-100  se3ROSToORB (Affine::rotation() is taken as the linear part: the
// SVD Eigen runs for Affine-mode transforms returns it to within rounding for a rotation)
inline Affine se3ROSToORB(const Affine &affineMatrix) {
  M3 tfCameraRotation = affineMatrix.R;
  V3 tfCameraTranslation = affineMatrix.t;
  M3 tfCameraRotationTemp = mul(tfROSToORB, tfCameraRotation);
  V3 tfCameraTranslationTemp = mul(tfROSToORB, tfCameraTranslation);
  M3 tfCameraRotationInv = transpose(tfCameraRotationTemp);
  V3 t = mul(tfCameraRotationInv, tfCameraTranslationTemp);
  V3 tfCameraTranslationInv = {{-t.v[0], -t.v[1], -t.v[2]}};
  Affine out;
  out.R = mul(tfROSToORB, tfCameraRotationInv);
  out.t = mul(tfROSToORB, tfCameraTranslationInv);
  return out;
}
// ORBW/src/type_conversion.cpp:45-72  se3ORBToROS
inline Affine se3ORBToROS(const Affine &s) {
  M3 tfCameraRotationTemp = mul(tfORBToROS, s.R);
  V3 tfCameraTranslationTemp = mul(tfORBToROS, s.t);
  M3 tfCameraRotationInv = transpose(tfCameraRotationTemp);
  V3 t = mul(tfCameraRotationInv, tfCameraTranslationTemp);
  V3 tfCameraTranslationInv = {{-t.v[0], -t.v[1], -t.v[2]}};
  Affine out;
  out.R = mul(tfORBToROS, tfCameraRotationInv);
  out.t = mul(tfORBToROS, tfCameraTranslationInv);
  return out;
}

// ---------------------------------------------------------------------------------------------
// DEP/src/Helpers.cpp:333-449  Fisher information helpers (Eigen types restated as float arrays)
// ---------------------------------------------------------------------------------------------
// :397-422  computeJacobianForPointLocal(p)  -> J[3][6]
inline void computeJacobianForPointLocal(const V3 &p, float J[3][6]) {
  // n = p.norm() = sqrt(squaredNorm)
  const float n = std::sqrt(dot3(p.v[0], p.v[0], p.v[1], p.v[1], p.v[2], p.v[2]));
  // df_dpc = (1 / n) * I - (1 / (n*n*n)) * p * p^T        ((s * p) * p^T, coefficient-wise)
  const float a = 1 / n;
  const float s = 1 / (n * n * n);
  M3 df;
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) df.m[i][j] = a * (i == j ? 1.0f : 0.0f) - (s * p.v[i]) * p.v[j];
  // rightMat = [ -I | skew(p) ]   (skew: DEP/src/Helpers.cpp:333-340)
  float right[3][6];
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) right[i][j] = (float)((-1.0) * (i == j ? 1.0f : 0.0f));
  right[0][3] = 0;
  right[0][4] = -p.v[2];
  right[0][5] = p.v[1];
  right[1][3] = p.v[2];
  right[1][4] = 0;
  right[1][5] = -p.v[0];
  right[2][3] = -p.v[1];
  right[2][4] = p.v[0];
  right[2][5] = 0;
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 6; ++j)
      J[i][j] = dot3(df.m[i][0], right[0][j], df.m[i][1], right[1][j], df.m[i][2], right[2][j]);
}
// :424-427 computeFIM = J^T * Q^-1 * J ; :440-449 trace.   Q = Identity (FisherInfoManager.cpp:174)
inline float computeInformationOfPointLocal(const V3 &p) {
  float J[3][6];
  computeJacobianForPointLocal(p, J);
  M3 Q = {{{1, 0, 0}, {0, 1, 0}, {0, 0, 1}}};
  M3 Qi = inverse3(Q);
  float JtQ[6][3];
  for (int i = 0; i < 6; ++i)
    for (int k = 0; k < 3; ++k) JtQ[i][k] = dot3(J[0][i], Qi.m[0][k], J[1][i], Qi.m[1][k], J[2][i], Qi.m[2][k]);
  float d[6];
  for (int i = 0; i < 6; ++i) d[i] = dot3(JtQ[i][0], J[0][i], JtQ[i][1], J[1][i], JtQ[i][2], J[2][i]);
  return (d[0] + (d[1] + d[2])) + (d[3] + (d[4] + d[5]));
}

// ---------------------------------------------------------------------------------------------
// DEP/include/frontier_exploration/fisher_information/FisherInfoManager.hpp:23-123 and
// DEP/src/fisher_information/FisherInfoManager.cpp:39-324
// ---------------------------------------------------------------------------------------------
const float step_min = 0.09;
const float subSampleVoxelUntil_m = -1.0;

struct LookupKey {
  std::array<float, 3> components;
  bool operator==(const LookupKey &other) const { return components == other.components; }
};
struct LookupValue {
  double information;
  int pointCount;
  unsigned int version;
  LookupValue() : information(0.0), pointCount(0), version(std::numeric_limits<unsigned int>::max()) {}
};
struct LookupKeyHash {
  size_t operator()(const LookupKey &key) const {
    size_t result = 0;
    for (const auto &component : key.components)
      result ^= std::hash<float>{}(component) + 0x9e3779b9 + (result << 6) + (result >> 2);
    return result;
  }
};

class FisherInformationManager {
 public:
  explicit FisherInformationManager(const fitslam_params &p) : prm(p), step_max(p.voxel_step) {
    latest_version_ = 1;
    occupied_voxel_count_ = 0;
  }
  fitslam_params prm;
  const float step_max;
  std::unordered_map<LookupKey, LookupValue, LookupKeyHash> lookup_table_fi_;
  std::vector<std::pair<LookupKey, float>> records_;  // file order (16-byte records)
  unsigned int latest_version_;
  unsigned int occupied_voxel_count_;

  inline float getFactorFromNum(int num, float s) { return std::exp(1 - std::pow(num, s)); }

  // hpp:108-123
  inline bool getVoxelCoordinate(float &voxel_x, float &voxel_y, float &voxel_z, float landmark_x, float landmark_y,
                                 float landmark_z) {
    double corrected_step;
    if (std::abs(landmark_x) < subSampleVoxelUntil_m && std::abs(landmark_y) < subSampleVoxelUntil_m &&
        std::abs(landmark_z) < subSampleVoxelUntil_m)
      corrected_step = step_min;
    else
      corrected_step = step_max;
    if (corrected_step == 0) return false;
    voxel_x = std::round(landmark_x * (1 / corrected_step)) * corrected_step;
    voxel_y = std::round(landmark_y * (1 / corrected_step)) * corrected_step;
    voxel_z = std::round(landmark_z * (1 / corrected_step)) * corrected_step;
    return true;
  }

  // cpp:117-229 (file output kept in memory as records_; writeRecords() reproduces the .dat bytes)
  void generateLookupTable(float minX, float maxX, float minY, float maxY, float minZ, float maxZ) {
    records_.clear();
    float max_fi_value = -std::numeric_limits<float>::max();
    float min_fi_value = std::numeric_limits<float>::max();
    minX = std::floor(minX * (1 / step_max)) * step_max;
    minY = std::floor(minY * (1 / step_max)) * step_max;
    minZ = std::floor(minZ * (1 / step_max)) * step_max;
    maxX = std::ceil(maxX * (1 / step_max)) * step_max;
    maxY = std::ceil(maxY * (1 / step_max)) * step_max;
    maxZ = std::ceil(maxZ * (1 / step_max)) * step_max;
    std::unordered_map<LookupKey, bool, LookupKeyHash> lookupTableExistingValues;
    float increment_value = step_min;
    for (float counter_x = minX; counter_x <= maxX; counter_x += increment_value) {
      if (counter_x > subSampleVoxelUntil_m + step_max) increment_value = step_max;
      for (float counter_y = minY; counter_y <= maxY; counter_y += increment_value) {
        for (float counter_z = minZ; counter_z <= maxZ; counter_z += increment_value) {
          float x, y, z;
          if (!getVoxelCoordinate(x, y, z, counter_x, counter_y, counter_z)) continue;
          LookupKey key;
          key.components[0] = x;
          key.components[1] = y;
          key.components[2] = z;
          V3 landmark_point = {{x, y, z}};
          if (lookupTableExistingValues.find(key) != lookupTableExistingValues.end()) continue;
          lookupTableExistingValues[key] = true;
          float value = computeInformationOfPointLocal(landmark_point);
          if (std::isnan(value)) continue;
          max_fi_value = std::max(max_fi_value, value);
          min_fi_value = std::min(min_fi_value, value);
          records_.push_back({key, value});
        }
      }
    }
    LookupKey key;
    key.components[0] = 0.0;
    key.components[1] = 0.0;
    key.components[2] = 0.0;
    records_.push_back({key, max_fi_value});
  }

  bool writeRecords(const char *path) {
    std::ofstream outfile(path, std::ios::binary);
    if (!outfile) return false;
    for (auto &r : records_) {
      outfile.write(reinterpret_cast<const char *>(&r.first), sizeof(LookupKey));
      outfile.write(reinterpret_cast<const char *>(&r.second), sizeof(float));
    }
    return true;
  }

  // cpp:231-262
  bool loadLookupTable(const char *path) {
    lookup_table_fi_.clear();
    std::ifstream infile(path, std::ios::binary);
    if (!infile) return false;
    LookupKey key;
    float value;
    while (infile.read(reinterpret_cast<char *>(&key), sizeof(LookupKey)) &&
           infile.read(reinterpret_cast<char *>(&value), sizeof(float))) {
      lookup_table_fi_[key].information = value;
    }
    return true;
  }
  void loadFromRecords() {
    lookup_table_fi_.clear();
    for (auto &r : records_) lookup_table_fi_[r.first].information = r.second;
  }

  // cpp:287-324 (tf2::Vector3 overload: doubles narrowed to the float parameters of getVoxelCoordinate)
  float getInformationFromLookup(double px, double py, double pz, unsigned int latest_version) {
    LookupKey search_key;
    if (!getVoxelCoordinate(search_key.components[0], search_key.components[1], search_key.components[2], px, py, pz))
      return std::numeric_limits<float>::quiet_NaN();
    auto it = lookup_table_fi_.find(search_key);
    if (it != lookup_table_fi_.end()) {
      if (it->second.version == latest_version) {
        it->second.pointCount++;
      } else {
        it->second.pointCount = 1;
        it->second.version = latest_version;
        occupied_voxel_count_++;
      }
      return it->second.information * getFactorFromNum(it->second.pointCount, prm.crowd_exponent);
    } else {
      return std::numeric_limits<float>::quiet_NaN();
    }
  }

  // ORBW/src/orb_slam3_interface.cpp:174-317 (exhaustive search over one keyframe holding every point)
  // + ORBW/src/rgbd/rgbd-slam-node.cpp:406-429 : returns the visible points in the query-pose frame.
  // robotBase_to_cameraLink_ = identity.  Map points are given in the ROS map frame;
  // their ORB world coordinates are the exact axis permutation (-y, -z, x).
  void landmarksInView(const fitslam_pose &cameraPose, const float *xyz, uint64_t P, std::vector<V3> &map_points) {
    Affine T_mapros_base = poseToAffine(cameraPose);
    Affine T_maporb_cam = T_mapros_base;  // B^-1 * T * B with B = identity
    Affine Tcw = se3ROSToORB(T_maporb_cam);
    const M3 &mRcw = Tcw.R;
    const V3 &mtcw = Tcw.t;
    const float fx = prm.fx, fy = prm.fy, cx = prm.cx, cy = prm.cy;
    float mnMinX = 0.0f, mnMaxX = prm.img_max_u, mnMinY = 0.0f, mnMaxY = prm.img_max_v;
    Affine affineMapToPos = poseToAffine(cameraPose);
    Affine affinePosToMap = inverseAffine(affineMapToPos);
    map_points.clear();
    for (uint64_t i = 0; i < P; ++i) {
      const float rx = xyz[3 * i], ry = xyz[3 * i + 1], rz = xyz[3 * i + 2];
      V3 Pw = {{-ry, -rz, rx}};  // ROS -> ORB axes (type_conversion.cpp:80-83 applied to a point)
      V3 RP = mul(mRcw, Pw);
      V3 Pc = {{RP.v[0] + mtcw.v[0], RP.v[1] + mtcw.v[1], RP.v[2] + mtcw.v[2]}};
      const float PcZ = Pc.v[2];
      if (PcZ < 0.0f) continue;
      // ORB/src/CameraModels/Pinhole.cpp:43-49
      float u = fx * Pc.v[0] / Pc.v[2] + cx;
      float v = fy * Pc.v[1] / Pc.v[2] + cy;
      if (u < mnMinX || u > mnMaxX) continue;
      if (v < mnMinY || v > mnMaxY) continue;
      // rgbd-slam-node.cpp:419-425: vector3fORBToROS then affinePosToMap * position
      V3 position = mul(tfORBToROS, Pw);
      map_points.push_back(apply(affinePosToMap, position));
    }
  }

  // Eigen 3.4 MatrixBase::eulerAngles(2, 0, 1) (Eigen/src/Geometry/EulerAngles.h: a0 = 2, a1 = 0 => odd = 0, i = 2, j = 0, k = 1)
  // followed by WrapperTypeConversions::rotationORBToEulerROS (ORBW/src/type_conversion.cpp:143-153); only component 2 is used
  static float yawOfRotationORBToEulerROS(const M3 &R) {
    const int i = 2, j = 0, k = 1;
    float res0 = atan2f(R.m[j][k], R.m[k][k]);
    const float c2 = sqrtf(R.m[i][i] * R.m[i][i] + R.m[i][j] * R.m[i][j]);
    float res1;
    if (res0 > 0.0f) {  // (!odd) && res[0] > 0
      if (res0 > 0.0f) res0 -= (float)M_PI;
      else res0 += (float)M_PI;
      res1 = atan2f(-R.m[i][k], -c2);
    } else {
      res1 = atan2f(-R.m[i][k], c2);
    }
    (void)res1;
    const float s1 = sinf(res0), c1 = cosf(res0);
    float res2 = atan2f(s1 * R.m[k][i] - c1 * R.m[j][i], c1 * R.m[j][j] - s1 * R.m[k][j]);
    res2 = -res2;          // if (!odd) res = -res
    float e2 = -res2;      // type_conversion.cpp:147
    // :151 (the float is promoted to double for the sum and narrowed back into the Vector3f)
    e2 = e2 < 0 ? (float)((2 * M_PI) + (double)e2) : e2;
    return e2;
  }

  // ORBSLAM3Interface::mapPointsVisibleFromPose over a keyframe-structured atlas (ORBW/src/orb_slam3_interface.cpp:188-317):
  // keyframes in the caller's order, each with GetPoseInverse() and its map-point list; per-keyframe prefilter unless
  // exhaustive_search (:240-267); the maxLandmarks cutoff (:270-271); each map point once (:263-265); then the node's transform
  // into the query-pose frame (ORBW/src/rgbd/rgbd-slam-node.cpp:406-429).  rel_T_CamToKF = Tcw * Twc is formed with 3x3
  // products here (Sophus multiplies unit quaternions: the same transform up to float rounding).
  void landmarksInViewKeyframes(const fitslam_pose &cameraPose, const float *xyz, uint64_t P, const fitslam_keyframe *kfs,
                                uint32_t n_kf, const uint32_t *point_index, int maxLandmarks, float maxDistance,
                                bool exhaustive_search, std::vector<V3> &map_points, uint32_t &numKFsChecked_out) {
    Affine Tcw = se3ROSToORB(poseToAffine(cameraPose));
    const M3 &mRcw = Tcw.R;
    const V3 &mtcw = Tcw.t;
    const float fx = prm.fx, fy = prm.fy, cx = prm.cx, cy = prm.cy;
    float mnMinX = 0.0f, mnMaxX = prm.img_max_u, mnMinY = 0.0f, mnMaxY = prm.img_max_v;
    Affine affinePosToMap = inverseAffine(poseToAffine(cameraPose));
    int numVisibleMapPoints = 0;
    int numKFsChecked = 0;
    std::vector<uint8_t> processedMapPoints(P, 0);
    map_points.clear();
    for (uint32_t kf = 0; kf < n_kf; ++kf) {
      M3 Rwc;
      V3 twc;
      for (int a = 0; a < 3; ++a) {
        for (int b = 0; b < 3; ++b) Rwc.m[a][b] = kfs[kf].Rwc[3 * a + b];
        twc.v[a] = kfs[kf].twc[a];
      }
      const M3 relR = mul(mRcw, Rwc);
      const V3 rt = mul(mRcw, twc);
      const V3 relt = {{rt.v[0] + mtcw.v[0], rt.v[1] + mtcw.v[1], rt.v[2] + mtcw.v[2]}};
      const float yaw = yawOfRotationORBToEulerROS(relR);
      // Eigen stableNorm-free norm(): sqrt of the redux x0*x0 + (x1*x1 + x2*x2)
      const float distBwKfs = sqrtf(dot3(relt.v[0], relt.v[0], relt.v[1], relt.v[1], relt.v[2], relt.v[2]));
      if (!exhaustive_search) {
        if (distBwKfs >= maxDistance * 2.0) continue;
        if (distBwKfs > maxDistance) {
          if (relt.v[2] < 0.0) continue;
          else if (relt.v[2] > 0.0) {
            if ((yaw >= 0 && yaw < M_PI / 2.0) || (yaw > 3.0 * M_PI / 2.0 && yaw < 2 * M_PI)) continue;
          }
        }
        if (distBwKfs <= maxDistance) {
          if (relt.v[2] < 0.0) {
            if ((yaw >= 2.0944 && yaw < 4.18879)) continue;
          }
        }
      }
      numKFsChecked++;
      if (numVisibleMapPoints > maxLandmarks) break;
      for (uint32_t e = kfs[kf].first; e < kfs[kf].first + kfs[kf].count; ++e) {
        const uint32_t pi = point_index[e];
        if (processedMapPoints[pi]) continue;
        processedMapPoints[pi] = 1;
        const float rx = xyz[3 * (uint64_t)pi], ry = xyz[3 * (uint64_t)pi + 1], rz = xyz[3 * (uint64_t)pi + 2];
        V3 Pw = {{-ry, -rz, rx}};
        V3 RP = mul(mRcw, Pw);
        V3 Pc = {{RP.v[0] + mtcw.v[0], RP.v[1] + mtcw.v[1], RP.v[2] + mtcw.v[2]}};
        if (Pc.v[2] < 0.0f) continue;
        float u = fx * Pc.v[0] / Pc.v[2] + cx;
        float v = fy * Pc.v[1] / Pc.v[2] + cy;
        if (u < mnMinX || u > mnMaxX) continue;
        if (v < mnMinY || v > mnMaxY) continue;
        V3 position = mul(tfORBToROS, Pw);
        map_points.push_back(apply(affinePosToMap, position));
        numVisibleMapPoints++;
      }
    }
    numKFsChecked_out = (uint32_t)numKFsChecked;
  }

  // isPoseSafe (cpp:39-115) over the response of the keyframe-structured service
  bool isPoseSafeKeyframes(const fitslam_pose &given_pose, const float *xyz, uint64_t P, const fitslam_keyframe *kfs, uint32_t n_kf,
                           const uint32_t *point_index, int maxLandmarks, float maxDistance, bool exhaustive_search,
                           float &information, uint32_t &n_landmarks) {
    ++latest_version_;
    std::vector<V3> map_points;
    uint32_t checked = 0;
    landmarksInViewKeyframes(given_pose, xyz, P, kfs, n_kf, point_index, maxLandmarks, maxDistance, exhaustive_search, map_points,
                             checked);
    n_landmarks = (uint32_t)map_points.size();
    float total_information = 0.0;
    occupied_voxel_count_ = 0;
    n_used_ = 0;
    total_information_f64_ = 0.0;
    for (const auto &mp : map_points) {
      float info = getInformationFromLookup(mp.v[0], mp.v[1], mp.v[2], latest_version_);
      if (std::isnan(info)) continue;
      ++n_used_;
      total_information_f64_ += (double)info;
      total_information += info;
    }
    information = total_information;
    return total_information > prm.fisher_information_threshold;
  }

  // cpp:39-115
  double total_information_f64_ = 0.0;  // stage dump: the same per-landmark float terms accumulated in double
  uint32_t n_used_ = 0;  // stage dump: landmarks of the last query that hit a lookup-table entry
  bool isPoseSafe(const fitslam_pose &given_pose, const float *xyz, uint64_t P, float &information,
                  uint32_t &n_visible) {
    double fisher_information_threshold = prm.fisher_information_threshold;
    ++latest_version_;
    std::vector<V3> map_points;
    landmarksInView(given_pose, xyz, P, map_points);
    n_visible = (uint32_t)map_points.size();
    float total_information = 0.0;
    occupied_voxel_count_ = 0;
    n_used_ = 0;
    total_information_f64_ = 0.0;
    for (const auto &mp : map_points) {
      // MapPoint.position is float64 on the wire, widened from the float the server computed
      double px = mp.v[0], py = mp.v[1], pz = mp.v[2];
      float info = getInformationFromLookup(px, py, pz, latest_version_);
      if (std::isnan(info)) continue;
      ++n_used_;
      total_information_f64_ += (double)info;
      total_information += info;
    }
    information = total_information;
    if (total_information > fisher_information_threshold) return true;
    return false;
  }
};

}  // namespace

// =================================================================================================
// extern "C" surface used by tests/ and bench.py (ctypes)
// =================================================================================================
extern "C" {

int orc_search_frontiers(const uint8_t *map, uint32_t W, uint32_t H, double res, double ox, double oy,
                         const fitslam_params *p, double rx, double ry, fitslam_frontier *out, uint32_t *n_frontiers,
                         uint32_t *every_cells, uint32_t *n_cells, uint32_t *comp_seed, uint32_t *n_comp,
                         double *centroids_xy) {
  Costmap2D cm = makeMap(map, W, H, res, ox, oy);
  FrontierSearch fs(cm, *p);
  auto list = fs.searchFrom(rx, ry);
  if (fs.robot_out_of_bounds) {
    *n_frontiers = 0;
    if (n_cells) *n_cells = 0;
    if (n_comp) *n_comp = 0;
    return FITSLAM_E_ROBOT_OUT_OF_BOUNDS;
  }
  int status = FITSLAM_OK;
  uint32_t cap = *n_frontiers;
  *n_frontiers = (uint32_t)list.size();
  if (list.size() > cap) status = FITSLAM_E_CAPACITY;
  for (size_t i = 0; i < list.size() && i < cap; ++i) {
    fitslam_frontier f{};
    f.goal_x = list[i].gx;
    f.goal_y = list[i].gy;
    f.goal_my = list[i].gidx / W;
    f.goal_mx = list[i].gidx - f.goal_my * W;
    f.size = list[i].size;
    f.component = list[i].component;
    f.chunk = list[i].chunk;
    f.first_cell = list[i].first_cell;
    out[i] = f;
    if (centroids_xy) {
      centroids_xy[2 * i] = list[i].centroid_x;
      centroids_xy[2 * i + 1] = list[i].centroid_y;
    }
  }
  if (n_cells) {
    uint32_t ccap = *n_cells;
    *n_cells = (uint32_t)fs.every_frontier_cells.size();
    if (every_cells) {
      if (fs.every_frontier_cells.size() > ccap) status = FITSLAM_E_CAPACITY;
      for (size_t i = 0; i < fs.every_frontier_cells.size() && i < ccap; ++i) every_cells[i] = fs.every_frontier_cells[i];
    }
  }
  if (n_comp) {
    uint32_t ccap = *n_comp;
    *n_comp = (uint32_t)fs.component_seed.size();
    if (comp_seed)
      for (size_t i = 0; i < fs.component_seed.size() && i < ccap; ++i) comp_seed[i] = fs.component_seed[i];
  }
  return status;
}

// static frontier mask (isNewFrontierCell with no flags) for stage-wise diffing
int orc_frontier_mask(const uint8_t *map, uint32_t W, uint32_t H, const fitslam_params *p, uint8_t *mask) {
  Costmap2D cm = makeMap(map, W, H, 1.0, 0.0, 0.0);
  FrontierSearch fs(cm, *p);
  fs.bindMap();
  for (uint32_t i = 0; i < W * H; ++i) mask[i] = fs.isFrontierMaskCell(i) ? 1 : 0;
  return FITSLAM_OK;
}

int orc_nearest_free_cell(const uint8_t *map, uint32_t W, uint32_t H, uint32_t start, int val, uint32_t *result) {
  Costmap2D cm = makeMap(map, W, H, 1.0, 0.0, 0.0);
  unsigned int r = 0;
  bool ok = nearestFreeCell(r, start, (unsigned char)val, cm);
  *result = r;
  return ok ? 1 : 0;
}

int orc_ray_shape(const fitslam_params *p, double res, uint32_t *n_rays, uint32_t *k, uint32_t *n_windows,
                  uint32_t *max_length) {
  uint32_t n = 0;
  for (double theta = 0; theta <= (2 * M_PI); theta += p->delta_theta) ++n;
  int kk = static_cast<int>(p->camera_fov / p->delta_theta);
  *n_rays = n;
  *k = (uint32_t)kk;
  *n_windows = n - kk + 1;
  unsigned int ml = p->max_camera_depth / res;
  *max_length = ml;
  return FITSLAM_OK;
}

// Costmap ingest.  (1) TraversabilityThresholdNode::callback, ORB-SLAM3-ROS2-Docker/traversability_mapping/
// traversability_ros_interface/traversability_mapping_ros/src/threshold_occupancy.cpp:25-49.  (2) StaticLayer::interpretValue
// of nav2_costmap_2d (ROS 2 Humble binary package, absent from the reference tree; restated from the published source:
// static_layer.cpp, `unsigned char StaticLayer::interpretValue(unsigned char value)`), applied to every cell of the grid as
// StaticLayer::processMap does.
static unsigned char interpretValue(unsigned char value, const fitslam_occupancy_params &P) {
  const bool track_unknown_space_ = P.track_unknown_space != 0, trinary_costmap_ = P.trinary_costmap != 0;
  const unsigned char unknown_cost_value_ = (unsigned char)P.unknown_cost_value, lethal_threshold_ = (unsigned char)P.lethal_cost_threshold;
  if (track_unknown_space_ && value == unknown_cost_value_) {
    return 255;  // NO_INFORMATION
  } else if (!track_unknown_space_ && value == unknown_cost_value_) {
    return 0;  // FREE_SPACE
  } else if (value >= lethal_threshold_) {
    return 254;  // LETHAL_OBSTACLE
  } else if (trinary_costmap_) {
    return 0;
  }
  double scale = static_cast<double>(value) / lethal_threshold_;
  return scale * 254;
}
int orc_occupancy_to_cost(const int8_t *data, uint64_t n, const fitslam_occupancy_params *p, uint8_t *out) {
  std::vector<int8_t> thresholded;
  const int8_t *grid = data;
  if (p->lethal_obstacle_threshold >= 0) {
    const int lethal_threshold_ = p->lethal_obstacle_threshold;
    thresholded.reserve(n);
    for (uint64_t i = 0; i < n; ++i) {
      const int8_t value = data[i];
      if (value >= lethal_threshold_) thresholded.push_back(100);
      else if (value == -1) thresholded.push_back(-1);
      else thresholded.push_back(0);
    }
    grid = thresholded.data();
  }
  for (uint64_t i = 0; i < n; ++i) out[i] = interpretValue((unsigned char)grid[i], *p);
  return FITSLAM_OK;
}

// LethalMarker (DEP/src/nav2_plugins/lethal_marker.cpp; same layer: FS2/fit_slam2_nav2_plugins/plugins/keepout_layer.cpp).
// An area is cached as the list of cell indices its rays visit (addNewMarkedArea :219-228, addNewMarkedAreaFOV :230-239);
// updateCosts (:303-327) then writes 253 over every cached list (markCells :242-248).
namespace lethal_marker {
// clamp of a ray end point to the grid (cellEnforceBoundaries, :5-11)
void clampToGrid(int64_t &x, int64_t &y, const Costmap2D &costmap) {
  const int64_t last_x = (int64_t)costmap.getSizeInCellsX() - 1, last_y = (int64_t)costmap.getSizeInCellsY() - 1;
  if (x < 0) x = 0;
  if (x > last_x) x = last_x;
  if (y < 0) y = 0;
  if (y > last_y) y = last_y;
}
// the error-term line walk of rayTraceGeneric (:12-41): visits (x0, y0) first and (x1, y1) last; cells whose flat
// index falls outside [0, W * H) are not recorded
void traceLine(const Costmap2D &costmap, int64_t x0, int64_t y0, int64_t x1, int64_t y1, std::vector<unsigned int> &cells) {
  const int64_t map_width = costmap.getSizeInCellsX(), n_cells = (int64_t)(costmap.getSizeInCellsY() * costmap.getSizeInCellsX());
  const int64_t dx = std::abs(x1 - x0), dy = std::abs(y1 - y0);
  const int64_t sx = x0 < x1 ? 1 : -1, sy = y0 < y1 ? 1 : -1;
  int64_t err = dx - dy;
  for (;;) {
    const int64_t index = y0 * map_width + x0;
    if (index >= 0 && index < n_cells) cells.push_back(static_cast<unsigned int>(index));
    if (x0 == x1 && y0 == y1) return;
    const int64_t e2 = 2 * err;
    if (e2 > -dy) { err -= dy; x0 += sx; }
    if (e2 < dx) { err += dx; y0 += sy; }
  }
}
// getPointsInSemiCircle (:52-73)
std::vector<unsigned int> circleCells(unsigned int center_x, unsigned int center_y, unsigned int radius_in_cells,
                                      unsigned int numPoints, double robot_yaw, const Costmap2D &costmap) {
  std::vector<unsigned int> all_cells;
  for (unsigned int i = 0; i < numPoints; ++i) {
    double angle = 2.0 * M_PI * i / numPoints;
    int64_t x = static_cast<int64_t>(center_x + radius_in_cells * std::cos(robot_yaw - M_PI_2 + angle));
    int64_t y = static_cast<int64_t>(center_y + radius_in_cells * std::sin(robot_yaw - M_PI_2 + angle));
    clampToGrid(x, y, costmap);
    traceLine(costmap, center_x, center_y, static_cast<unsigned int>(x), static_cast<unsigned int>(y), all_cells);
  }
  return all_cells;
}
// getPointsInIsoscelesTriangle (:75-127)
std::vector<unsigned int> triangleCells(unsigned int apex_x, unsigned int apex_y, unsigned int triangle_height,
                                        unsigned int numPoints, double direction, double apex_angle, const Costmap2D &costmap) {
  std::vector<unsigned int> all_cells;
  double base_center_x = apex_x + triangle_height * std::cos(direction);
  double base_center_y = apex_y + triangle_height * std::sin(direction);
  double half_base = triangle_height * std::tan(apex_angle / 2.0);
  double left_endpoint_x = base_center_x + half_base * std::cos(direction + M_PI_2);
  double left_endpoint_y = base_center_y + half_base * std::sin(direction + M_PI_2);
  double right_endpoint_x = base_center_x + half_base * std::cos(direction - M_PI_2);
  double right_endpoint_y = base_center_y + half_base * std::sin(direction - M_PI_2);
  for (unsigned int i = 0; i < numPoints; ++i) {
    double t = (numPoints > 1) ? static_cast<double>(i) / (numPoints - 1) : 0.0;
    double sample_x = left_endpoint_x + t * (right_endpoint_x - left_endpoint_x);
    double sample_y = left_endpoint_y + t * (right_endpoint_y - left_endpoint_y);
    int64_t cell_x = static_cast<int64_t>(sample_x);
    int64_t cell_y = static_cast<int64_t>(sample_y);
    clampToGrid(cell_x, cell_y, costmap);
    traceLine(costmap, apex_x, apex_y, static_cast<unsigned int>(cell_x), static_cast<unsigned int>(cell_y), all_cells);
  }
  return all_cells;
}
}  // namespace lethal_marker
int orc_mark_lethal(uint8_t *map, uint32_t W, uint32_t H, double res, double ox, double oy, const fitslam_lethal_area *areas,
                    uint32_t n, uint32_t *n_marked) {
  Costmap2D cm = makeMap(map, W, H, res, ox, oy);
  std::map<int, std::vector<unsigned int>> latest_cells_to_mark_index_;
  int numAdditions_ = 0;
  for (uint32_t a = 0; a < n; ++a) {
    unsigned int robot_xm, robot_ym;
    if (!cm.worldToMap(areas[a].wx, areas[a].wy, robot_xm, robot_ym)) continue;
    auto extent_in_cells = areas[a].extent / cm.getResolution();  // a double, narrowed to unsigned int by the callee
    if (areas[a].kind == 0)
      latest_cells_to_mark_index_[numAdditions_] = lethal_marker::circleCells(robot_xm, robot_ym, extent_in_cells, 360, 0, cm);
    else
      latest_cells_to_mark_index_[numAdditions_] =
          lethal_marker::triangleCells(robot_xm, robot_ym, extent_in_cells, 20, areas[a].yaw, 45 * M_PI / 180, cm);
    numAdditions_++;
  }
  for (int add = 0; add < numAdditions_; add++)
    for (auto index : latest_cells_to_mark_index_.at(add)) map[index] = 253;
  if (n_marked) *n_marked = (uint32_t)numAdditions_;
  return FITSLAM_OK;
}

// SensorSimulator (RE/src/SensorSimulator.cpp): the ray loop of timerCallback (:193-202) and markRay (:211-246) on the
// int8 cells of two nav_msgs/OccupancyGrid messages of the same geometry.
int orc_sensor_mark_rays(const int8_t *latest, int8_t *explored, uint32_t size_x, uint32_t size_y, double resolution, double origin_x,
                         double origin_y, double robot_x, double robot_y, double base_yaw, const fitslam_sensor_params *p,
                         uint32_t *n_rays) {
  // `const double res = latest_map_->info.resolution;` (:217): nav_msgs/MapMetaData.resolution is a float32
  const double res = static_cast<double>(static_cast<float>(resolution));
  const int w = static_cast<int>(size_x);
  const int h = static_cast<int>(size_y);
  uint32_t cast = 0;
  for (double a = p->sensor_min_angle; a <= p->sensor_max_angle; a += p->angular_resolution) {
    const double ray_angle = base_yaw + a;
    const double dx = std::cos(ray_angle);
    const double dy = std::sin(ray_angle);
    ++cast;
    for (double r = 0.0; r <= p->sensor_max_range; r += res) {
      const double wx = robot_x + r * dx;
      const double wy = robot_y + r * dy;
      const int mx = static_cast<int>((wx - origin_x) / res);
      const int my = static_cast<int>((wy - origin_y) / res);
      if (mx < 0 || my < 0 || mx >= w || my >= h) break;
      const std::size_t idx = static_cast<std::size_t>(my) * w + mx;
      explored[idx] = latest[idx];
      if (latest[idx] > 80) break;
    }
  }
  if (n_rays) *n_rays = cast;
  return FITSLAM_OK;
}

// FrontierRoadMap::isConnectable (RE/src/planners/FrontierRoadmap.cpp:671-694) for a batch of node pairs; getTracedCells
// and bresenham2D are the same text in RE/src/Helpers.cpp:7-91 as in DEP (restated above).  max_connection_length_ =
// RADIUS_TO_DECIDE_EDGES * 1.5 (FrontierRoadmap.cpp:32).
int orc_roadmap_connectable(const uint8_t *map, uint32_t W, uint32_t H, double res, double ox, double oy,
                            const double *segments, uint32_t n, double radius_to_decide_edges, uint8_t *connectable,
                            int32_t *num_unknown) {
  Costmap2D cm = makeMap(map, W, H, res, ox, oy);
  const Costmap2D *costmap_ = &cm;
  const double RADIUS_TO_DECIDE_EDGES = radius_to_decide_edges;
  const double max_connection_length_ = RADIUS_TO_DECIDE_EDGES * 1.5;
  for (uint32_t i = 0; i < n; ++i) {
    const double f1x = segments[4 * i], f1y = segments[4 * i + 1], f2x = segments[4 * i + 2], f2y = segments[4 * i + 3];
    std::vector<MapLocation> traced_cells;
    RayTracedCells cell_gatherer(costmap_, traced_cells, 253, 254, 0, 255);
    unsigned int max_length = max_connection_length_ / costmap_->getResolution();
    bool ok = true;
    if (!getTracedCells(f1x, f1y, f2x, f2y, cell_gatherer, max_length, costmap_)) {
      connectable[i] = 0;
      if (num_unknown) num_unknown[i] = -1;
      continue;
    }
    if (cell_gatherer.hasHitObstacle()) ok = false;
    if (cell_gatherer.getNumUnknown() > RADIUS_TO_DECIDE_EDGES / costmap_->getResolution() * 0.3) ok = false;
    connectable[i] = ok ? 1 : 0;
    if (num_unknown) num_unknown[i] = (int32_t)cell_gatherer.getNumUnknown();
  }
  return FITSLAM_OK;
}

int orc_arrival_limits(const uint8_t *map, uint32_t W, uint32_t H, double res, double ox, double oy,
                       const fitslam_params *p, double *max_gt, double *min_gt, int32_t *max_value) {
  Costmap2D cm = makeMap(map, W, H, res, ox, oy);
  FrontierCostCalculator c = makeCalc(&cm, p);
  double mv = c.setMaxArrivalInformation();
  *max_gt = c.max_arrival_info_gt_;
  *min_gt = c.min_arrival_info_gt_;
  if (max_value) *max_value = (int32_t)mv;
  return FITSLAM_OK;
}

int orc_score_arrival(const uint8_t *map, uint32_t W, uint32_t H, double res, double ox, double oy,
                      const fitslam_params *p, const fitslam_frontier *fr, uint32_t n, const double *poly,
                      fitslam_arrival *out, uint16_t *info_along_ray, int32_t *window_sums) {
  Costmap2D cm = makeMap(map, W, H, res, ox, oy);
  FrontierCostCalculator c = makeCalc(&cm, p);
  c.setMaxArrivalInformation();
  uint32_t n_rays, k, n_windows, ml;
  orc_ray_shape(p, res, &n_rays, &k, &n_windows, &ml);
  std::vector<int> rays, result;
  for (uint32_t i = 0; i < n; ++i) {
    c.setArrivalInformationForFrontier(fr[i].goal_x, fr[i].goal_y, fr[i].size, poly, out[i], rays, result);
    if (info_along_ray)
      for (uint32_t r = 0; r < n_rays; ++r) info_along_ray[(size_t)i * n_rays + r] = r < rays.size() ? (uint16_t)rays[r] : 0;
    if (window_sums)
      for (uint32_t w = 0; w < n_windows; ++w) window_sums[(size_t)i * n_windows + w] = w < result.size() ? result[w] : 0;
  }
  return FITSLAM_OK;
}

// DEP/src/CostCalculator.cpp:445-485 setPlanForFrontierEuclidean (robot_yaw = quatToEuler(orientation)[2], an input)
int orc_plan_euclidean(double robot_x, double robot_y, double robot_yaw_in, const fitslam_frontier *fr, uint32_t n,
                       uint8_t *achievable, double *path_length_m, double *path_heading) {
  for (uint32_t i = 0; i < n; ++i) {
    if (!achievable[i]) {
      path_length_m[i] = std::numeric_limits<double>::max();
      path_heading[i] = std::numeric_limits<double>::max();
      continue;
    }
    auto length_to_frontier = sqrt(pow(robot_x - fr[i].goal_x, 2) + pow(robot_y - fr[i].goal_y, 2));
    if (length_to_frontier < 0.5) {
      achievable[i] = 0;
      path_length_m[i] = std::numeric_limits<double>::max();
      path_heading[i] = std::numeric_limits<double>::max();
      continue;
    }
    achievable[i] = 1;
    auto robot_yaw = robot_yaw_in;
    if (robot_yaw < 0) robot_yaw = robot_yaw + (M_PI * 2);
    double goal_yaw = atan2(fr[i].goal_y - robot_y, fr[i].goal_x - robot_x);
    if (goal_yaw < 0) goal_yaw = goal_yaw + (M_PI * 2);
    double heading = std::abs(robot_yaw - goal_yaw);
    if (heading > M_PI) heading = (2 * M_PI) - heading;
    path_length_m[i] = length_to_frontier;
    path_heading[i] = heading;
  }
  return FITSLAM_OK;
}

// DEP/src/FrontierCostsManager.cpp:74-205 : per-frontier bookkeeping (blacklist, recomputeNormalizationFactors
// DEP/src/CostCalculator.cpp:512-520) and the U1 utility block.  arrival[] / achievable[] are the state after
// setArrivalInformationForFrontier + the plan step; path_length == path_length_m as both plan producers set them.
// Blacklisted frontiers: the reference leaves path_heading unset and would throw from getPathHeading() inside the
// U1 loop (DEP/src/Frontier.cpp:200-205); here they keep weighted_cost = DBL_MAX and are skipped.
int orc_normalise(const fitslam_params *p, const fitslam_arrival *arrival, const uint8_t *achievable,
                  const double *path_length, const double *path_heading, const uint8_t *blacklisted, uint32_t n,
                  double max_arrival_info_gt, double *weighted_cost, double *arrival_utility, double *distance_utility,
                  int32_t *error_index) {
  if (n == 0) return FITSLAM_E_NO_FRONTIERS;
  // reset() CostCalculator.hpp:110-116
  double min_traversable_distance = std::numeric_limits<double>::max();
  double max_traversable_distance = -1.0;
  double min_arrival_info_per_frontier = std::numeric_limits<double>::max();
  double max_arrival_info_per_frontier = -1.0;
  for (uint32_t i = 0; i < n; ++i) {
    if (blacklisted && blacklisted[i]) continue;
    if (!achievable[i]) continue;
    min_traversable_distance = std::min(min_traversable_distance, path_length[i]);
    max_traversable_distance = std::max(max_traversable_distance, path_length[i]);
    min_arrival_info_per_frontier = std::min(min_arrival_info_per_frontier, (double)arrival[i].information);
    max_arrival_info_per_frontier = std::max(max_arrival_info_per_frontier, (double)arrival[i].information);
  }
  (void)max_arrival_info_per_frontier;
  const double alpha_ = p->alpha, beta_ = p->beta, max_vx_ = p->vx_max, max_wx_ = p->wz_max;
  for (uint32_t i = 0; i < n; ++i) {
    if (blacklisted && blacklisted[i]) {
      weighted_cost[i] = std::numeric_limits<double>::max();
      arrival_utility[i] = -69.8;
      distance_utility[i] = -1.8;
      continue;
    }
    if (!achievable[i]) {
      weighted_cost[i] = std::numeric_limits<double>::max();
      arrival_utility[i] = -69.8;
      distance_utility[i] = -1.8;
      continue;
    }
    double arrival_info_utility;
    if (static_cast<double>(max_arrival_info_gt - min_arrival_info_per_frontier) == 0.0)
      arrival_info_utility = 0.0;
    else
      arrival_info_utility = static_cast<double>((double)arrival[i].information) / static_cast<double>(max_arrival_info_gt);
    if (arrival_info_utility > 1.0) {
      if (error_index) *error_index = (int32_t)i;
      return FITSLAM_E_ARRIVAL_UTILITY;
    }
    double frontier_plan_utility;
    if ((max_traversable_distance / max_vx_ + M_PI / max_wx_) - (min_traversable_distance / max_vx_ + 0.0 / max_wx_) == 0.0)
      frontier_plan_utility = 1.0;
    else
      frontier_plan_utility = static_cast<double>(path_length[i] / max_vx_ + path_heading[i] / max_wx_) /
                              static_cast<double>(max_traversable_distance / max_vx_ + M_PI / max_wx_);
    frontier_plan_utility = 1.0 - frontier_plan_utility;
    if (arrival_info_utility > 1.0 || arrival_info_utility < 0.0 || frontier_plan_utility > 1.0 ||
        frontier_plan_utility < 0.0) {
      if (error_index) *error_index = (int32_t)i;
      return FITSLAM_E_COST_OUT_OF_BOUNDS;
    }
    double utility = (alpha_ * arrival_info_utility) + ((1.0 - alpha_) * frontier_plan_utility);
    if (utility == 0.0) utility = 1e-16;
    weighted_cost[i] = 1 / (beta_ * utility);
    arrival_utility[i] = arrival_info_utility;
    distance_utility[i] = frontier_plan_utility;
  }
  return FITSLAM_OK;
}

// DEP/src/FullPathOptimizer.cpp:157-227  getFilteredFrontiersN
// eligible[i] = isAchievable() && !isBlacklisted() && !isInBlacklistedRegion()
int orc_select(const fitslam_params *p, const uint8_t *eligible, const double *path_length_m, uint32_t n_frontiers,
               uint32_t *sorted, fitslam_selection *sel) {
  const size_t n = (size_t)p->num_frontiers_in_local_area;
  const double LOCAL_FRONTIER_SEARCH_RADIUS = p->local_frontier_search_radius;
  std::vector<uint32_t> all_frontiers;
  for (uint32_t i = 0; i < n_frontiers; ++i)
    if (eligible[i]) all_frontiers.push_back(i);
  std::sort(all_frontiers.begin(), all_frontiers.end(),
            [path_length_m](uint32_t a, uint32_t b) { return path_length_m[a] < path_length_m[b]; });
  std::memset(sel, 0, sizeof(*sel));
  sel->closest_global = -1;
  sel->n_sorted = (uint32_t)all_frontiers.size();
  for (size_t i = 0; i < all_frontiers.size(); ++i) sorted[i] = all_frontiers[i];
  if (all_frontiers.size() == 0) return FITSLAM_OK;
  bool global_assigned = false;
  size_t counter = 1;
  bool need_to_pop = false;
  std::vector<uint32_t> local_frontiers, global_frontiers;
  int64_t closest = -1;
  size_t pos = 0, global_begin = all_frontiers.size();
  for (const auto &frontier : all_frontiers) {
    if (path_length_m[frontier] <= LOCAL_FRONTIER_SEARCH_RADIUS && counter <= n) {
      local_frontiers.push_back(frontier);
      closest = frontier;
      need_to_pop = true;
    } else if (path_length_m[frontier] <= LOCAL_FRONTIER_SEARCH_RADIUS && counter > n) {
      closest = frontier;
      need_to_pop = false;
    } else if (path_length_m[frontier] > LOCAL_FRONTIER_SEARCH_RADIUS) {
      if (!global_assigned) {
        closest = frontier;
        global_assigned = true;
        global_begin = pos;
      }
      need_to_pop = false;
      global_frontiers.push_back(frontier);
    }
    counter++;
    pos++;
  }
  if (need_to_pop) local_frontiers.pop_back();
  if (global_frontiers.size() == 0 && local_frontiers.size() == 0) {
    global_frontiers.push_back((uint32_t)closest);
    // the only way here is a single eligible frontier: it is sorted[0]
    global_begin = 0;
  }
  sel->n_local = (uint32_t)local_frontiers.size();
  sel->n_global = (uint32_t)global_frontiers.size();
  sel->global_begin = (uint32_t)(global_frontiers.empty() ? all_frontiers.size() : global_begin);
  sel->closest_global = (int32_t)closest;
  return FITSLAM_OK;
}

// std::hash<double> based UID (DEP/include/frontier_exploration/Frontier.hpp:151-161)
uint64_t orc_generate_uid(double x, double y) {
  std::hash<double> hash_fn;
  std::size_t hash1 = hash_fn(x);
  std::size_t hash2 = hash_fn(y);
  return hash1 ^ (hash2 << 1);
}

// ---- Fisher ----
void *orc_fisher_create(const fitslam_params *p) { return new FisherInformationManager(*p); }
void orc_fisher_destroy(void *h) { delete static_cast<FisherInformationManager *>(h); }

int orc_fisher_generate(void *h, const char *save_path) {
  auto *m = static_cast<FisherInformationManager *>(h);
  m->generateLookupTable(m->prm.lut_min_x, m->prm.lut_max_x, m->prm.lut_min_y, m->prm.lut_max_y, m->prm.lut_min_z,
                         m->prm.lut_max_z);
  if (save_path && !m->writeRecords(save_path)) return FITSLAM_E_IO;
  m->loadFromRecords();
  return FITSLAM_OK;
}
int orc_fisher_load(void *h, const char *path) {
  auto *m = static_cast<FisherInformationManager *>(h);
  return m->loadLookupTable(path) ? FITSLAM_OK : FITSLAM_E_NO_LOOKUP_TABLE;
}
uint64_t orc_fisher_num_records(void *h) { return static_cast<FisherInformationManager *>(h)->records_.size(); }
uint64_t orc_fisher_num_entries(void *h) { return static_cast<FisherInformationManager *>(h)->lookup_table_fi_.size(); }
// records in file order: out[4*i .. 4*i+3] = key xyz, value
int orc_fisher_records(void *h, float *out, uint64_t cap) {
  auto *m = static_cast<FisherInformationManager *>(h);
  if (cap < m->records_.size()) return FITSLAM_E_CAPACITY;
  for (size_t i = 0; i < m->records_.size(); ++i) {
    out[4 * i + 0] = m->records_[i].first.components[0];
    out[4 * i + 1] = m->records_[i].first.components[1];
    out[4 * i + 2] = m->records_[i].first.components[2];
    out[4 * i + 3] = m->records_[i].second;
  }
  return FITSLAM_OK;
}
// single lookup without touching the crowding state (value of the voxel holding p, NaN if absent)
float orc_fisher_voxel_information(void *h, float x, float y, float z) {
  auto *m = static_cast<FisherInformationManager *>(h);
  LookupKey k;
  if (!m->getVoxelCoordinate(k.components[0], k.components[1], k.components[2], x, y, z))
    return std::numeric_limits<float>::quiet_NaN();
  auto it = m->lookup_table_fi_.find(k);
  if (it == m->lookup_table_fi_.end()) return std::numeric_limits<float>::quiet_NaN();
  return (float)it->second.information;
}
void orc_fisher_voxel_coordinate(void *h, float x, float y, float z, float *out3) {
  auto *m = static_cast<FisherInformationManager *>(h);
  m->getVoxelCoordinate(out3[0], out3[1], out3[2], x, y, z);
}
float orc_information_of_point_local(float x, float y, float z) {
  V3 p = {{x, y, z}};
  return computeInformationOfPointLocal(p);
}
float orc_crowd_factor(void *h, int num) {
  auto *m = static_cast<FisherInformationManager *>(h);
  return m->getFactorFromNum(num, m->prm.crowd_exponent);
}

int orc_fisher_poses(void *h, const float *xyz, uint64_t P, const fitslam_pose *poses, uint32_t n_poses,
                     float *information, uint8_t *safe, uint32_t *n_visible, uint32_t *n_used, double *information_f64) {
  auto *m = static_cast<FisherInformationManager *>(h);
  if (m->lookup_table_fi_.empty()) return FITSLAM_E_NO_LOOKUP_TABLE;
  for (uint32_t i = 0; i < n_poses; ++i) {
    float info = 0;
    uint32_t nv = 0;
    bool ok = m->isPoseSafe(poses[i], xyz, P, info, nv);
    information[i] = info;
    if (safe) safe[i] = ok ? 1 : 0;
    if (n_visible) n_visible[i] = nv;
    if (n_used) n_used[i] = m->n_used_;
    if (information_f64) information_f64[i] = m->total_information_f64_;
  }
  return FITSLAM_OK;
}

// visible landmarks of one pose, pose frame (the GetLandmarksInView response); returns count
int orc_landmarks_in_view(void *h, const float *xyz, uint64_t P, const fitslam_pose *pose, float *out_xyz, uint32_t cap,
                          uint32_t *count) {
  auto *m = static_cast<FisherInformationManager *>(h);
  std::vector<V3> pts;
  m->landmarksInView(*pose, xyz, P, pts);
  *count = (uint32_t)pts.size();
  for (size_t i = 0; i < pts.size() && i < cap; ++i) {
    out_xyz[3 * i] = pts[i].v[0];
    out_xyz[3 * i + 1] = pts[i].v[1];
    out_xyz[3 * i + 2] = pts[i].v[2];
  }
  return pts.size() > cap ? FITSLAM_E_CAPACITY : FITSLAM_OK;
}

// the keyframe-structured service response (pose frame) and isPoseSafe behind it
int orc_landmarks_in_view_keyframes(void *h, const float *xyz, uint64_t P, const fitslam_keyframe *kfs, uint32_t n_kf,
                                    const uint32_t *point_index, const fitslam_pose *pose, int32_t max_landmarks, float max_distance,
                                    int exhaustive, float *out_xyz, uint32_t cap, uint32_t *count, uint32_t *n_checked) {
  auto *m = static_cast<FisherInformationManager *>(h);
  std::vector<V3> pts;
  m->landmarksInViewKeyframes(*pose, xyz, P, kfs, n_kf, point_index, max_landmarks, max_distance, exhaustive != 0, pts, *n_checked);
  *count = (uint32_t)pts.size();
  for (size_t i = 0; i < pts.size() && i < cap; ++i) {
    out_xyz[3 * i] = pts[i].v[0];
    out_xyz[3 * i + 1] = pts[i].v[1];
    out_xyz[3 * i + 2] = pts[i].v[2];
  }
  return pts.size() > cap ? FITSLAM_E_CAPACITY : FITSLAM_OK;
}
int orc_fisher_pose_keyframes(void *h, const float *xyz, uint64_t P, const fitslam_keyframe *kfs, uint32_t n_kf,
                              const uint32_t *point_index, const fitslam_pose *pose, int32_t max_landmarks, float max_distance,
                              int exhaustive, float *information, double *information_f64, uint8_t *safe, uint32_t *n_landmarks,
                              uint32_t *n_used) {
  auto *m = static_cast<FisherInformationManager *>(h);
  if (m->lookup_table_fi_.empty()) return FITSLAM_E_NO_LOOKUP_TABLE;
  float info = 0;
  const bool ok = m->isPoseSafeKeyframes(*pose, xyz, P, kfs, n_kf, point_index, max_landmarks, max_distance, exhaustive != 0, info,
                                         *n_landmarks);
  *information = info;
  *information_f64 = m->total_information_f64_;
  *safe = ok ? 1 : 0;
  *n_used = m->n_used_;
  return FITSLAM_OK;
}

// ORBW/tests/typeConversionsTests.cpp:8-33 : se3ORBToROS(se3ROSToORB(A)) for an affine given as R (row-major 9) + t
void orc_se3_roundtrip(const float *R9, const float *t3, float *R9_out, float *t3_out) {
  Affine a;
  for (int i = 0; i < 3; ++i) {
    for (int j = 0; j < 3; ++j) a.R.m[i][j] = R9[3 * i + j];
    a.t.v[i] = t3[i];
  }
  Affine orb = se3ROSToORB(a);
  Affine back = se3ORBToROS(orb);
  for (int i = 0; i < 3; ++i) {
    for (int j = 0; j < 3; ++j) R9_out[3 * i + j] = back.R.m[i][j];
    t3_out[i] = back.t.v[i];
  }
}

// tf2 / nav2_util orientationAroundZAxis: setRPY(0, 0, angle) -> (x, y, z, w)
void orc_orientation_around_z(double angle, double *q4) {
  double halfYaw = angle * 0.5;
  double halfPitch = 0.0 * 0.5;
  double halfRoll = 0.0 * 0.5;
  double cosYaw = cos(halfYaw), sinYaw = sin(halfYaw);
  double cosPitch = cos(halfPitch), sinPitch = sin(halfPitch);
  double cosRoll = cos(halfRoll), sinRoll = sin(halfRoll);
  q4[0] = sinRoll * cosPitch * cosYaw - cosRoll * sinPitch * sinYaw;
  q4[1] = cosRoll * sinPitch * cosYaw + sinRoll * cosPitch * sinYaw;
  q4[2] = cosRoll * cosPitch * sinYaw - sinRoll * sinPitch * cosYaw;
  q4[3] = cosRoll * cosPitch * cosYaw + sinRoll * sinPitch * sinYaw;
}

// one full reference-order cycle on the CPU (the timed cpu_baseline): search -> arrival -> euclidean plan ->
// U1 -> select; Fisher over `n_fisher_poses` given poses.  Returns frontier count through *n_frontiers.
int orc_cycle(const uint8_t *map, uint32_t W, uint32_t H, double res, double ox, double oy, const fitslam_params *p,
              double rx, double ry, double ryaw, const double *poly, uint32_t *n_frontiers, uint32_t *n_eligible) {
  Costmap2D cm = makeMap(map, W, H, res, ox, oy);
  FrontierSearch fs(cm, *p);
  auto list = fs.searchFrom(rx, ry);
  *n_frontiers = (uint32_t)list.size();
  if (list.empty()) return FITSLAM_E_NO_FRONTIERS;
  FrontierCostCalculator c = makeCalc(&cm, p);
  c.setMaxArrivalInformation();
  size_t n = list.size();
  std::vector<fitslam_frontier> fr(n);
  std::vector<fitslam_arrival> arr(n);
  std::vector<uint8_t> ach(n);
  std::vector<double> len(n), head(n), wc(n), ua(n), ud(n);
  std::vector<int> rays, result;
  for (size_t i = 0; i < n; ++i) {
    fr[i].goal_x = list[i].gx;
    fr[i].goal_y = list[i].gy;
    fr[i].size = list[i].size;
    c.setArrivalInformationForFrontier(list[i].gx, list[i].gy, list[i].size, poly, arr[i], rays, result);
    ach[i] = arr[i].achievable;
  }
  orc_plan_euclidean(rx, ry, ryaw, fr.data(), (uint32_t)n, ach.data(), len.data(), head.data());
  int32_t err = -1;
  int st = orc_normalise(p, arr.data(), ach.data(), len.data(), head.data(), nullptr, (uint32_t)n, c.max_arrival_info_gt_,
                         wc.data(), ua.data(), ud.data(), &err);
  if (st != FITSLAM_OK) return st;
  std::vector<uint32_t> sorted(n);
  fitslam_selection sel;
  orc_select(p, ach.data(), len.data(), (uint32_t)n, sorted.data(), &sel);
  if (n_eligible) *n_eligible = sel.n_sorted;
  return FITSLAM_OK;
}

void orc_default_params(fitslam_params *p) {
  std::memset(p, 0, sizeof(*p));
  p->min_frontier_cluster_size = 1;
  p->max_frontier_cluster_size = 20;
  p->max_frontier_distance = 50.0;
  p->lethal_threshold = 160;
  p->max_camera_depth = 2.0;
  p->delta_theta = 0.10;
  p->camera_fov = 1.04;
  p->robot_radius = 0.60;
  p->max_arrival_factor = 1.2;
  p->min_arrival_factor = 0.70;
  p->alpha = 0.25;
  p->beta = 1.0;
  p->vx_max = 0.5;
  p->wz_max = 0.5;
  p->num_frontiers_in_local_area = 5;
  p->local_frontier_search_radius = 12.0;
  p->voxel_step = 0.3f;
  p->crowd_exponent = 0.8f;
  p->fisher_information_threshold = 550.0;
  p->lut_min_x = 0.0f;
  p->lut_max_x = 21.0f;
  p->lut_min_y = (float)(-8.5 * 1.732);
  p->lut_max_y = (float)(8.5 * 1.732);
  p->lut_min_z = (float)(-8.5 * 1.732);
  p->lut_max_z = (float)(8.5 * 1.732);
  p->fx = (float)432.496042035043;
  p->fy = (float)432.496042035043;
  p->cx = 320.0f;
  p->cy = 240.0f;
  p->img_max_u = 640.0f;
  p->img_max_v = 480.0f;
  p->candidate_pose_z = 0.3;
}

}  // extern "C"
