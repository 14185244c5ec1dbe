// ref_harness.cpp -- C entry points over the REFERENCE's own frontier search and ray helpers, compiled from
// /root/reference where it lies (oracle/Makefile, target _ref/libfitslam_ref.so): DEP/src/Frontier.cpp,
// DEP/src/FrontierSearch.cpp, the costmap half of DEP/src/Helpers.cpp (lines 1-330: bresenham2D, getTracedCells,
// isRobotFootprintInLethal, nhood4 / nhood8, nearestFreeCell) and four member functions of DEP/src/CostCalculator.cpp
// (setArrivalInformationForFrontier, setMaxArrivalInformation, setPlanForFrontierEuclidean, recomputeNormalizationFactors),
// against the stand-in headers of oracle/ref_shims/.
// Test infrastructure: it pins the restatement in oracle.cpp (tests/test_ref_pin.py) and nothing else links it.
#include <cstdint>
#include <cstring>
#include <vector>

#include <frontier_exploration/CostCalculator.hpp>
#include <frontier_exploration/FrontierCostsManager.hpp>
#include <frontier_exploration/FullPathOptimizer.hpp>
#include <frontier_exploration/fisher_information/FisherInfoManager.hpp>
#include <frontier_exploration/FrontierSearch.hpp>
#include <frontier_exploration/Helpers.hpp>
#include <frontier_exploration/nav2_plugins/lethal_marker.hpp>
#include <nav_msgs/msg/occupancy_grid.hpp>

using frontier_exploration::FrontierSearch;
using nav2_costmap_2d::Costmap2D;

namespace {
Costmap2D make_map(const uint8_t *cells, uint32_t W, uint32_t H, double res, double ox, double oy) {
  Costmap2D m(W, H, res, ox, oy);
  std::memcpy(m.getCharMap(), cells, (size_t)W * H);
  return m;
}
}  // namespace

// where FisherInfoManager.cpp reads / writes its lookup table (ref_shims/frontier_exploration/util/rosVisualizer.hpp)
static std::string g_lookup_path;
std::string fitslam_ref_lookup_path() { return g_lookup_path; }

// TraversabilityThresholdNode::callback (ORB-SLAM3-ROS2-Docker/traversability_mapping/.../threshold_occupancy.cpp:24-49): the
// member function's own text, cut by oracle/Makefile into _ref/threshold_occupancy_callback.inc, inside a class that has the two
// members it names: the threshold and a publisher that keeps the message.
namespace threshold_node {
struct CapturePublisher {
  nav_msgs::msg::OccupancyGrid last;
  void publish(const nav_msgs::msg::OccupancyGrid &m) { last = m; }
};
class TraversabilityThresholdNode {
public:
#include "_ref/threshold_occupancy_callback.inc"
  CapturePublisher capture_, *publisher_ = &capture_;
  int lethal_threshold_ = 85;
};
}  // namespace threshold_node

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

// FrontierCostCalculator::setMaxArrivalInformation then setArrivalInformationForFrontier for n frontiers (DEP/src/CostCalculator.cpp:
// 23-191), as FrontierCostsManager::assignCosts calls them.  goals: n x 3 doubles (goal x, goal y, size); poly: {xmin, ymin, xmax,
// ymax}.  out: n x 6 doubles (arrival information, achievable, goal orientation x y z w); limits: {return value of
// setMaxArrivalInformation, max_arrival_info_gt, min_arrival_info_gt}.
void ref_score_arrival(const uint8_t *cells, uint32_t W, uint32_t H, double res, double ox, double oy, double max_camera_depth,
                       double delta_theta, double camera_fov, double robot_radius, const double *goals, uint32_t n, const double *poly,
                       double *out, double *limits) {
  Costmap2D map = make_map(cells, W, H, res, ox, oy);
  frontier_exploration::FrontierCostCalculator calc(&map, max_camera_depth, delta_theta, camera_fov, robot_radius);
  limits[0] = calc.setMaxArrivalInformation();
  limits[1] = calc.getMaxArrivalInformation();
  limits[2] = calc.getMinArrivalInformationGT();
  std::vector<double> polygon(poly, poly + 4);
  for (uint32_t i = 0; i < n; ++i) {
    FrontierPtr f = std::make_shared<Frontier>();
    f->setGoalPoint(goals[3 * i], goals[3 * i + 1]);
    f->setSize((int)goals[3 * i + 2]);
    f->setAchievability(true);
    calc.setArrivalInformationForFrontier(f, polygon);
    out[6 * i] = f->getArrivalInformation();
    out[6 * i + 1] = f->isAchievable() ? 1.0 : 0.0;
    const auto q = f->getGoalOrientation();
    out[6 * i + 2] = q.x; out[6 * i + 3] = q.y; out[6 * i + 4] = q.z; out[6 * i + 5] = q.w;
  }
}

// setPlanForFrontierEuclidean (DEP/src/CostCalculator.cpp:445-485) + recomputeNormalizationFactors (:512-520) over n frontiers.
// robot: {x, y, qx, qy, qz, qw}; goals: n x 2; achievable_in: n flags; arrival: n doubles (for the normalisation factors).
// out: n x 3 doubles (achievable, path length, path heading); norm: {min, max plan distance, min arrival information};
// *robot_yaw = quatToEuler(orientation)[2] as the function computes it (before its own +2 pi).
void ref_plan_euclidean(const double *robot, const double *goals, const uint8_t *achievable_in, const double *arrival, uint32_t n,
                        double *out, double *norm, double *robot_yaw) {
  Costmap2D map(4, 4, 1.0, 0.0, 0.0);
  frontier_exploration::FrontierCostCalculator calc(&map, 2.0, 0.1, 1.04, 0.6);
  calc.reset();
  geometry_msgs::msg::Pose start;
  start.position.x = robot[0]; start.position.y = robot[1];
  start.orientation.x = robot[2]; start.orientation.y = robot[3]; start.orientation.z = robot[4]; start.orientation.w = robot[5];
  *robot_yaw = frontier_exploration::quatToEuler(start.orientation)[2];
  for (uint32_t i = 0; i < n; ++i) {
    FrontierPtr f = std::make_shared<Frontier>();
    f->setGoalPoint(goals[2 * i], goals[2 * i + 1]);
    f->setAchievability(achievable_in[i] != 0);
    f->setArrivalInformation(arrival[i]);
    calc.setPlanForFrontierEuclidean(start, f, nullptr, false, false);
    calc.recomputeNormalizationFactors(f);
    out[3 * i] = f->isAchievable() ? 1.0 : 0.0;
    out[3 * i + 1] = f->getPathLength();
    out[3 * i + 2] = f->getPathHeading();
  }
  norm[0] = calc.getMinPlanDistance();
  norm[1] = calc.getMaxPlanDistance();
  norm[2] = calc.getMinArrivalInformation();
}

// FrontierCostsManager::assignCosts with cost types {"ArrivalInformation", "EuclideanDistance"} (DEP/src/FrontierCostsManager.cpp:
// 47-213: arrival information, Euclidean plan, normalisation factors, the U1 block) followed by
// FullPathOptimizer::getFilteredFrontiersN (DEP/src/FullPathOptimizer.cpp:157-227) with n = n_local.
// goals: n x 3 (goal x, goal y, size); robot: {x, y, qx, qy, qz, qw}; blacklisted: n flags (setFrontierBlacklist).
// out: n x 7 doubles (weighted cost, arrival_gain_utility, distance_utility, achievable, arrival information, path length, path
// heading); sel: {n_local, n_global, index of closest_global_frontier or -1}; local_idx / global_idx: frontier indices in order.
// Returns 0, 1 if assignCosts returned false, 2 / 3 if it threw "ARRIVAL UTILITY ERROR..." / "Cost out of bounds", 4 for another throw.
int ref_assign_costs(const uint8_t *cells, uint32_t W, uint32_t H, double res, double ox, double oy, double max_camera_depth,
                     double delta_theta, double camera_fov, double robot_radius, double alpha, double beta, double vx_max, double wz_max,
                     const double *goals, uint32_t n, const double *poly, const double *robot, const uint8_t *blacklisted, uint32_t n_local,
                     double *out, int32_t *sel, int32_t *local_idx, int32_t *global_idx) {
  Costmap2D map = make_map(cells, W, H, res, ox, oy);
  frontier_exploration::FrontierCostsManager mgr(&map, max_camera_depth, delta_theta, camera_fov, robot_radius, alpha, beta, vx_max, wz_max);
  std::vector<FrontierPtr> list, black;
  for (uint32_t i = 0; i < n; ++i) {
    FrontierPtr f = std::make_shared<Frontier>();
    f->setGoalPoint(goals[3 * i], goals[3 * i + 1]);
    f->setSize((int)goals[3 * i + 2]);
    f->setUID(generateUID(f));
    f->setAchievability(true);
    f->setBlacklisted(false);
    list.push_back(f);
    if (blacklisted && blacklisted[i]) black.push_back(f);
  }
  if (!black.empty()) mgr.setFrontierBlacklist(black);
  geometry_msgs::msg::Pose start;
  start.position.x = robot[0]; start.position.y = robot[1];
  start.orientation.x = robot[2]; start.orientation.y = robot[3]; start.orientation.z = robot[4]; start.orientation.w = robot[5];
  std::vector<std::vector<std::string>> types(n, {"ArrivalInformation", "EuclideanDistance"});
  std::vector<double> polygon(poly, poly + 4);
  int status = 0;
  try {
    if (!mgr.assignCosts(list, polygon, start, nullptr, types)) status = 1;
  } catch (const std::runtime_error &e) {
    const std::string w = e.what();
    status = w.rfind("ARRIVAL UTILITY ERROR", 0) == 0 ? 2 : (w == "Cost out of bounds" ? 3 : 4);
  }
  if (status != 0) return status;
  for (uint32_t i = 0; i < n; ++i) {
    const FrontierPtr &f = list[i];
    out[7 * i] = f->getWeightedCost();
    const bool blk = blacklisted && blacklisted[i];
    out[7 * i + 1] = blk ? 0.0 : f->getCost("arrival_gain_utility");
    out[7 * i + 2] = blk ? 0.0 : f->getCost("distance_utility");
    out[7 * i + 3] = f->isAchievable() ? 1.0 : 0.0;
    out[7 * i + 4] = f->getArrivalInformation();
    out[7 * i + 5] = f->getPathLength();
    out[7 * i + 6] = blk ? 0.0 : f->getPathHeading();
    if (blk) f->setBlacklisted(true);  // what the behaviour tree does with a blacklisted frontier before the selection
  }
  frontier_exploration::FullPathOptimizer opt;
  frontier_exploration::SortedFrontiers sorted;
  geometry_msgs::msg::PoseStamped robotP;
  robotP.pose = start;
  opt.getFilteredFrontiersN(list, n_local, sorted, robotP);
  auto index_of = [&](const FrontierPtr &f) { for (uint32_t i = 0; i < n; ++i) if (list[i] == f) return (int32_t)i; return (int32_t)-1; };
  sel[0] = (int32_t)sorted.local_frontiers.size();
  sel[1] = (int32_t)sorted.global_frontiers.size();
  sel[2] = sorted.closest_global_frontier ? index_of(sorted.closest_global_frontier) : -1;
  for (size_t i = 0; i < sorted.local_frontiers.size(); ++i) local_idx[i] = index_of(sorted.local_frontiers[i]);
  for (size_t i = 0; i < sorted.global_frontiers.size(); ++i) global_idx[i] = index_of(sorted.global_frontiers[i]);
  return 0;
}

// ---- stage F: the reference's FisherInformationManager (DEP/src/fisher_information/FisherInfoManager.cpp, whole) --------------
// generateLookupTable into the file `path` (FisherInfoManager.cpp:117-229 + the Fisher helpers of DEP/src/Helpers.cpp)
void ref_fisher_generate(const char *path, float min_x, float max_x, float min_y, float max_y, float min_z, float max_z) {
  g_lookup_path = path;
  frontier_exploration::FisherInformationManager m;
  m.generateLookupTable(min_x, max_x, min_y, max_y, min_z, max_z);
}
// a manager that loads the table at `path` (the constructor's loadLookupTable, :231-262); 0 if the file cannot be read
void *ref_fisher_create(const char *path) {
  g_lookup_path = path;
  try {
    return new frontier_exploration::FisherInformationManager(rclcpp::Node::make_shared("ref"));
  } catch (const std::runtime_error &) {
    return nullptr;
  }
}
void ref_fisher_destroy(void *h) { delete static_cast<frontier_exploration::FisherInformationManager *>(h); }
// isPoseSafe (:39-115) with the service answering `n` landmarks (pose frame, float64 as on the wire).  Returns safe.
int ref_fisher_pose(void *h, const double *landmarks, uint32_t n, double threshold, float *information) {
  using Srv = slam_msgs::srv::GetLandmarksInView;
  parameterInstance.set("fisherInformation/fisher_information_threshold", threshold);
  rclcpp::Client<Srv>::handler() = [landmarks, n](const Srv::Request &, Srv::Response &resp) {
    resp.map_points.resize(n);
    for (uint32_t i = 0; i < n; ++i) {
      resp.map_points[i].position.x = landmarks[3 * i];
      resp.map_points[i].position.y = landmarks[3 * i + 1];
      resp.map_points[i].position.z = landmarks[3 * i + 2];
    }
  };
  geometry_msgs::msg::Pose pose;
  pose.orientation.w = 1.0;
  float info = 0.0f;
  const bool safe = static_cast<frontier_exploration::FisherInformationManager *>(h)->isPoseSafe(pose, true, info);
  *information = info;
  return safe ? 1 : 0;
}
// getVoxelCoordinate and getFactorFromNum (FisherInfoManager.hpp:102-123)
int ref_fisher_voxel(float x, float y, float z, float *out3) {
  frontier_exploration::FisherInformationManager m;
  return m.getVoxelCoordinate(out3[0], out3[1], out3[2], x, y, z) ? 1 : 0;
}
float ref_crowd_factor(int num, float s) {
  frontier_exploration::FisherInformationManager m;
  return m.getFactorFromNum(num, s);
}
// computeInformationOfPointLocal(p, I) (DEP/src/Helpers.cpp:440-449)
float ref_information_of_point(float x, float y, float z) {
  Eigen::Vector3f p(x, y, z);
  return frontier_exploration::computeInformationOfPointLocal(p, Eigen::Matrix3f::Identity());
}

// LethalMarker::addNewMarkedArea / addNewMarkedAreaFOV per area, then one updateCosts over the whole map (DEP/src/nav2_plugins/
// lethal_marker.cpp:218-236, 305-327).  areas: n * 5 doubles (wx, wy, extent, yaw, kind); cells is stamped in place.
// Returns the number of areas that were on the map (numAdditions_).
int ref_mark_lethal(uint8_t *cells, uint32_t W, uint32_t H, double res, double ox, double oy, const double *areas, uint32_t n) {
  Costmap2D map = make_map(cells, W, H, res, ox, oy);
  nav2_costmap_2d::LayeredCostmap layered(&map);
  frontier_exploration::LethalMarker layer(&layered);
  for (uint32_t a = 0; a < n; ++a) {
    const double *r = areas + 5 * a;
    if (r[4] == 0.0)
      layer.addNewMarkedArea(r[0], r[1], r[2]);
    else
      layer.addNewMarkedAreaFOV(r[0], r[1], r[3], r[2]);
  }
  layer.updateCosts(map, 0, 0, (int)W, (int)H);
  std::memcpy(cells, map.getCharMap(), (size_t)W * H);
  return layer.additions();
}

// The threshold node on n int8 cells: out = the published message's data.
void ref_threshold_occupancy(const int8_t *data, uint64_t n, int lethal_threshold, int8_t *out) {
  threshold_node::TraversabilityThresholdNode node;
  node.lethal_threshold_ = lethal_threshold;
  auto msg = std::make_shared<nav_msgs::msg::OccupancyGrid>();
  msg->data.assign(data, data + n);
  node.callback(msg);
  std::memcpy(out, node.capture_.last.data.data(), node.capture_.last.data.size());
}

}  // extern "C"
