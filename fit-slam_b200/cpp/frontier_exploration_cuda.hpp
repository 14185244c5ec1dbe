// frontier_exploration_cuda.hpp -- the reference's `frontier_exploration` class interfaces over the fitslam C ABI.
//
// Header-only C++17 host layer: same class names, method names, argument meaning and error behaviour as the classes it
// replaces (file:line below, paths relative to dev_ws/src/DEPRECATED/frontier_exploration/frontier_exploration/), every
// method forwarding to one call of include/fitslam_cuda.h.  No ROS dependency: with FITSLAM_HAVE_ROS defined the real
// nav2_costmap_2d / geometry_msgs headers are used; otherwise the minimal stand-ins below (the members this path reads,
// with the semantics of ROS 2 Humble's costmap_2d.cpp) are.  There is no CPU fallback: construction throws if the CUDA
// library cannot create a context.
//
//   FrontierSearch              include/frontier_exploration/FrontierSearch.hpp:28-65,  src/FrontierSearch.cpp:21-254
//   Frontier / FrontierPtr      include/frontier_exploration/Frontier.hpp:39-136,        src/Frontier.cpp:5-246
//   FrontierCostCalculator      include/frontier_exploration/CostCalculator.hpp:56-116,  src/CostCalculator.cpp:23-191,512-520
//   FisherInformationManager    include/.../fisher_information/FisherInfoManager.hpp:69-127, src/.../FisherInfoManager.cpp:39-324
#pragma once
#include <cmath>
#include <cstdint>
#include <functional>
#include <limits>
#include <map>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

#include "fitslam_cuda.h"

#ifdef FITSLAM_HAVE_ROS
#include <geometry_msgs/msg/point.hpp>
#include <geometry_msgs/msg/pose.hpp>
#include <nav2_costmap_2d/costmap_2d.hpp>
#else
// ---- stand-ins (only what the path reads) -------------------------------------------------------------------------
namespace geometry_msgs { namespace msg {
struct Point { double x = 0, y = 0, z = 0; };
struct Quaternion { double x = 0, y = 0, z = 0, w = 1; };
struct Pose { Point position; Quaternion orientation; };
}}  // namespace geometry_msgs::msg
namespace nav2_costmap_2d {
static constexpr unsigned char NO_INFORMATION = 255, LETHAL_OBSTACLE = 254, INSCRIBED_INFLATED_OBSTACLE = 253, FREE_SPACE = 0;
class Costmap2D {
public:
  Costmap2D(unsigned int size_x, unsigned int size_y, double resolution, double origin_x, double origin_y, unsigned char def = 0)
      : size_x_(size_x), size_y_(size_y), resolution_(resolution), origin_x_(origin_x), origin_y_(origin_y),
        costmap_((size_t)size_x * size_y, def) {}
  unsigned char *getCharMap() { return costmap_.data(); }
  const unsigned char *getCharMap() const { return costmap_.data(); }
  unsigned int getSizeInCellsX() const { return size_x_; }
  unsigned int getSizeInCellsY() const { return size_y_; }
  double getResolution() const { return resolution_; }
  double getOriginX() const { return origin_x_; }
  double getOriginY() const { return origin_y_; }
  unsigned int getIndex(unsigned int mx, unsigned int my) const { return my * size_x_ + mx; }
  void mapToWorld(unsigned int mx, unsigned int my, double &wx, double &wy) const {
    wx = origin_x_ + (mx + 0.5) * resolution_;
    wy = origin_y_ + (my + 0.5) * resolution_;
  }
  bool worldToMap(double wx, double wy, unsigned int &mx, unsigned int &my) const {
    if (wx < origin_x_ || wy < origin_y_) return false;
    mx = static_cast<unsigned int>((wx - origin_x_) / resolution_);
    my = static_cast<unsigned int>((wy - origin_y_) / resolution_);
    return mx < size_x_ && my < size_y_;
  }
private:
  unsigned int size_x_, size_y_;
  double resolution_, origin_x_, origin_y_;
  std::vector<unsigned char> costmap_;
};
}  // namespace nav2_costmap_2d
#endif

namespace frontier_exploration {

// One CUDA context shared by the classes of one exploration server (the reference shares one Costmap2D the same way,
// src/ExplorationBT.cpp:1204-1213).  Status codes become the reference's own exceptions / return values.
class CudaContext {
public:
  explicit CudaContext(const fitslam_params *params = nullptr, int device = 0) {
    fitslam_params p;
    fitslam_default_params(&p);
    if (params) p = *params;
    const fitslam_status st = fitslam_create(&p, device, &ctx_);
    if (st != FITSLAM_OK) throw std::runtime_error(std::string("fitslam_create: ") + fitslam_status_string(st));
  }
  ~CudaContext() { fitslam_destroy(ctx_); }
  CudaContext(const CudaContext &) = delete;
  CudaContext &operator=(const CudaContext &) = delete;
  fitslam_ctx *get() const { return ctx_; }
  // snapshot of the costmap (the caller holds costmap->getMutex() as in ExplorationBT.cpp:170-195)
  void updateCostmap(const nav2_costmap_2d::Costmap2D &cm) {
    check(fitslam_set_costmap(ctx_, cm.getCharMap(), cm.getSizeInCellsX(), cm.getSizeInCellsY(), cm.getResolution(),
                              cm.getOriginX(), cm.getOriginY(), nullptr), "fitslam_set_costmap");
  }
  // the same from a nav_msgs/OccupancyGrid (data: int8 cells, row-major): threshold node + static layer on the device
  // (threshold_occupancy.cpp:25-49, nav2 StaticLayer::interpretValue); xywh = a dirty rectangle of a resident map, or null
  void updateOccupancyGrid(const int8_t *data, uint32_t size_x, uint32_t size_y, double resolution, double origin_x, double origin_y,
                           const fitslam_occupancy_params *params = nullptr, const uint32_t *xywh = nullptr) {
    check(fitslam_set_occupancy_grid(ctx_, data, size_x, size_y, resolution, origin_x, origin_y, params, xywh), "fitslam_set_occupancy_grid");
  }
  void check(fitslam_status st, const char *what) const {
    if (st != FITSLAM_OK) throw std::runtime_error(std::string(what) + ": " + fitslam_status_string(st) + " (" + fitslam_last_error(ctx_) + ")");
  }
private:
  fitslam_ctx *ctx_ = nullptr;
};
using CudaContextPtr = std::shared_ptr<CudaContext>;

// ---- Frontier (Frontier.hpp:39-136): getters throw std::runtime_error on unset fields (Frontier.cpp:142-239) --------
inline size_t generateUID(double x, double y) {  // Frontier.hpp:151-161
  return std::hash<double>{}(x) ^ (std::hash<double>{}(y) << 1);
}
class Frontier {
public:
  Frontier() { is_achievable_ = std::make_shared<bool>(true); is_blacklisted_ = std::make_shared<bool>(false); }
  void setUID(size_t uid) { set(unique_id_, uid); }
  void setSize(int sz) { set(size_, sz); }
  void setGoalPoint(geometry_msgs::msg::Point gp) { set(goal_point_, gp); }
  void setGoalPoint(double x, double y) { geometry_msgs::msg::Point p; p.x = x; p.y = y; set(goal_point_, p); }
  void setGoalOrientation(double theta) {  // nav2_util::geometry_utils::orientationAroundZAxis = tf2 setRPY(0, 0, theta)
    set(theta_s_star_, theta);
    geometry_msgs::msg::Quaternion q;
    q.x = 0.0; q.y = 0.0; q.z = std::sin(theta * 0.5); q.w = std::cos(theta * 0.5);
    set(best_orientation_, q);
  }
  void setArrivalInformation(double info) { set(information_, info); }
  void setPathLength(double pl) { set(path_length_, pl); }
  void setPathLengthInM(double pl) { set(path_length_m_, pl); }
  void setPathHeading(double h) { set(path_heading_, h); }
  void setFisherInformation(double fi) { set(fisher_information_in_path_, fi); }
  void setCost(std::string name, double value) { if (!costs_) costs_ = std::make_shared<std::map<std::string, double>>(); (*costs_)[name] = value; }
  void setWeightedCost(double c) { set(weighted_cost_, c); }
  void setAchievability(bool v) { set(is_achievable_, v); }
  void setBlacklisted(bool v) { set(is_blacklisted_, v); }
  bool operator==(const Frontier &o) const { return getGoalPoint().x == o.getGoalPoint().x && getGoalPoint().y == o.getGoalPoint().y; }
  bool operator<(const Frontier &o) const { return getUID() < o.getUID(); }
  size_t getUID() const { return get(unique_id_, "Goal uid frontier property is null"); }
  int getSize() const { return get(size_, "Size frontier property is null"); }
  geometry_msgs::msg::Point &getGoalPoint() const { need(goal_point_, "Goal point frontier property is null"); return *goal_point_; }
  geometry_msgs::msg::Quaternion &getGoalOrientation() const { need(best_orientation_, "Goal orientation frontier property is null"); return *best_orientation_; }
  double getArrivalInformation() const { return get(information_, "Arrival information frontier property is null"); }
  double getPathLength() const { return get(path_length_, "Path length frontier property is null"); }
  double getPathLengthInM() const { return get(path_length_m_, "Path length in m frontier property is null"); }
  double getPathHeading() const { return get(path_heading_, "Path heading frontier property is null"); }
  double getFisherInformation() const { return get(fisher_information_in_path_, "Fisher information frontier property is null"); }
  double getCost(const std::string &name) const {
    if (!costs_ || !costs_->count(name)) throw std::runtime_error("Cost frontier property is null");
    return costs_->at(name);
  }
  double getWeightedCost() const { return get(weighted_cost_, "Weighted cost frontier property is null"); }
  bool isAchievable() const { return get(is_achievable_, "Achievability frontier property is null"); }
  bool isBlacklisted() const { return get(is_blacklisted_, "Blacklist frontier property is null"); }
  bool isFrontierNull() const { return !unique_id_ && !size_ && !goal_point_; }
  // not in the reference: the cell the goal point is the centre of, kept so that a batch call needs no worldToMap
  uint32_t goal_mx = 0, goal_my = 0;
private:
  template <typename T> static void set(std::shared_ptr<T> &p, const T &v) { if (p) *p = v; else p = std::make_shared<T>(v); }
  template <typename T> static void need(const std::shared_ptr<T> &p, const char *msg) { if (!p) throw std::runtime_error(msg); }
  template <typename T> static T get(const std::shared_ptr<T> &p, const char *msg) { need(p, msg); return *p; }
  std::shared_ptr<size_t> unique_id_;
  std::shared_ptr<int> size_;
  std::shared_ptr<geometry_msgs::msg::Point> goal_point_;
  std::shared_ptr<geometry_msgs::msg::Quaternion> best_orientation_;
  std::shared_ptr<double> theta_s_star_, information_, path_length_, path_length_m_, path_heading_, fisher_information_in_path_;
  std::shared_ptr<bool> is_achievable_, is_blacklisted_;
  std::shared_ptr<std::map<std::string, double>> costs_;
  std::shared_ptr<double> weighted_cost_;
};
using FrontierPtr = std::shared_ptr<Frontier>;

// ---- FrontierSearch (FrontierSearch.hpp:28-65) -------------------------------------------------------------------------
class FrontierSearch {
public:
  // the reference constructor reads frontierSearch/* from ParameterHandler (FrontierSearch.cpp:10-19); here the values come
  // from the context's fitslam_params
  explicit FrontierSearch(nav2_costmap_2d::Costmap2D &costmap, CudaContextPtr ctx = nullptr, double max_frontier_distance = 50.0)
      : costmap_(costmap), ctx_(ctx ? ctx : std::make_shared<CudaContext>()), max_frontier_distance_(max_frontier_distance),
        original_search_distance_(max_frontier_distance) {}
  void reset() { every_frontier_list.clear(); }
  void incrementSearchDistance(double value) { max_frontier_distance_ += value; }
  void resetSearchDistance() { max_frontier_distance_ = original_search_distance_; }
  // searchFrom (FrontierSearch.cpp:21-96): robot off the map -> empty list (the reference logs CRITICAL and returns {})
  std::vector<FrontierPtr> searchFrom(geometry_msgs::msg::Point position) {
    ctx_->updateCostmap(costmap_);
    ctx_->check(fitslam_set_search_distance(ctx_->get(), max_frontier_distance_), "fitslam_set_search_distance");
    std::vector<fitslam_frontier> raw(1024);
    std::vector<uint32_t> cells(1 << 16);
    for (;;) {
      uint32_t n = (uint32_t)raw.size(), nc = (uint32_t)cells.size();
      const fitslam_status st = fitslam_search_frontiers(ctx_->get(), position.x, position.y, raw.data(), &n, cells.data(), &nc);
      if (st == FITSLAM_E_ROBOT_OUT_OF_BOUNDS) return {};
      if (st == FITSLAM_E_CAPACITY) { raw.resize(std::max<size_t>(n, raw.size())); cells.resize(std::max<size_t>(nc, cells.size())); continue; }
      ctx_->check(st, "fitslam_search_frontiers");
      raw.resize(n);
      cells.resize(nc);
      break;
    }
    std::vector<FrontierPtr> out;
    out.reserve(raw.size());
    for (const fitslam_frontier &r : raw) {  // FrontierSearch.cpp:157-186
      auto f = std::make_shared<Frontier>();
      f->setSize(r.size);
      f->setGoalPoint(r.goal_x, r.goal_y);
      f->setUID(generateUID(r.goal_x, r.goal_y));
      f->goal_mx = r.goal_mx;
      f->goal_my = r.goal_my;
      out.push_back(f);
    }
    const unsigned int W = costmap_.getSizeInCellsX();
    for (uint32_t c : cells) {  // getAllFrontiers: cell centres in acceptance order (FrontierSearch.cpp:107-110,139-142)
      double wx, wy;
      costmap_.mapToWorld(c % W, c / W, wx, wy);
      every_frontier_list.push_back({wx, wy});
    }
    return out;
  }
  std::vector<std::vector<double>> getAllFrontiers() { return every_frontier_list; }
  CudaContextPtr context() const { return ctx_; }
private:
  nav2_costmap_2d::Costmap2D &costmap_;
  CudaContextPtr ctx_;
  double max_frontier_distance_, original_search_distance_;
  std::vector<std::vector<double>> every_frontier_list;
};

// ---- FrontierCostCalculator (CostCalculator.hpp:56-116): the arrival-information half --------------------------------
class FrontierCostCalculator {
public:
  FrontierCostCalculator(nav2_costmap_2d::Costmap2D *costmap, CudaContextPtr ctx) : exploration_costmap_(costmap), ctx_(ctx) { reset(); }
  // setArrivalInformationForFrontier (CostCalculator.cpp:23-121) for a whole list in one launch; a failed ray leaves
  // information 0 / orientation 0 for that frontier and is not an error (:50-55)
  void setArrivalInformationForFrontiers(std::vector<FrontierPtr> &frontiers, std::vector<double> &polygon_xy_min_max) {
    if (frontiers.empty()) return;
    if (polygon_xy_min_max.size() < 4) throw std::runtime_error("polygon_xy_min_max needs 4 values");
    if (max_arrival_info_gt_ == 0.0) setMaxArrivalInformation();
    std::vector<fitslam_frontier> in(frontiers.size());
    for (size_t i = 0; i < frontiers.size(); ++i) {
      fitslam_frontier &r = in[i];
      r = fitslam_frontier{};
      r.goal_x = frontiers[i]->getGoalPoint().x;
      r.goal_y = frontiers[i]->getGoalPoint().y;
      r.size = frontiers[i]->getSize();
      unsigned int mx = 0, my = 0;
      exploration_costmap_->worldToMap(r.goal_x, r.goal_y, mx, my);
      r.goal_mx = mx;
      r.goal_my = my;
    }
    std::vector<fitslam_arrival> out(frontiers.size());
    ctx_->check(fitslam_score_arrival(ctx_->get(), in.data(), (uint32_t)in.size(), polygon_xy_min_max.data(), out.data(), nullptr, nullptr),
                "fitslam_score_arrival");
    for (size_t i = 0; i < frontiers.size(); ++i) {
      frontiers[i]->setArrivalInformation((double)out[i].information);
      frontiers[i]->setGoalOrientation(out[i].theta_s_star);
      if (!out[i].achievable) frontiers[i]->setAchievability(false);  // CostCalculator.cpp:77-82, 112-117
    }
  }
  void setArrivalInformationForFrontier(FrontierPtr &goal_point_w, std::vector<double> &polygon_xy_min_max) {
    std::vector<FrontierPtr> one{goal_point_w};
    setArrivalInformationForFrontiers(one, polygon_xy_min_max);
  }
  double setMaxArrivalInformation() {  // CostCalculator.cpp:123-191
    int32_t max_value = 0;
    ctx_->check(fitslam_arrival_limits(ctx_->get(), &max_arrival_info_gt_, &min_arrival_info_gt_, &max_value), "fitslam_arrival_limits");
    return max_arrival_info_gt_;
  }
  void recomputeNormalizationFactors(FrontierPtr &frontier) {  // CostCalculator.cpp:512-520
    if (!frontier->isAchievable()) return;
    min_traversable_distance = std::min(min_traversable_distance, frontier->getPathLength());
    max_traversable_distance = std::max(max_traversable_distance, frontier->getPathLength());
    min_arrival_info_per_frontier = std::min(min_arrival_info_per_frontier, frontier->getArrivalInformation());
    max_arrival_info_per_frontier = std::max(max_arrival_info_per_frontier, frontier->getArrivalInformation());
  }
  double getMinPlanDistance() { return min_traversable_distance; }
  double getMaxPlanDistance() { return max_traversable_distance; }
  double getMinArrivalInformation() { return min_arrival_info_per_frontier; }
  double getMaxArrivalInformation() { return max_arrival_info_gt_; }
  void reset() {
    min_traversable_distance = std::numeric_limits<double>::max();
    max_traversable_distance = -1.0;
    min_arrival_info_per_frontier = std::numeric_limits<double>::max();
    max_arrival_info_per_frontier = -1.0;
  }
private:
  nav2_costmap_2d::Costmap2D *exploration_costmap_;
  CudaContextPtr ctx_;
  double min_traversable_distance, max_traversable_distance, min_arrival_info_per_frontier, max_arrival_info_per_frontier;
  double max_arrival_info_gt_ = 0.0, min_arrival_info_gt_ = 0.0;
};

// ---- FisherInformationManager (FisherInfoManager.hpp:69-127) ----------------------------------------------------------
// The landmark service round trip of the reference (FisherInfoManager.cpp:48-75) is replaced by the map points resident
// on the device (setMapPoints); everything after it is the same arithmetic.
class FisherInformationManager {
public:
  explicit FisherInformationManager(CudaContextPtr ctx, std::string lookup_path = "") : ctx_(ctx), lookup_path_(std::move(lookup_path)) {}
  void setMapPoints(const float *xyz, uint64_t n) { ctx_->check(fitslam_set_map_points(ctx_->get(), xyz, n), "fitslam_set_map_points"); }
  void generateLookupTable(float minX, float maxX, float minY, float maxY, float minZ, float maxZ) {  // FisherInfoManager.cpp:117-229
    (void)minX; (void)maxX; (void)minY; (void)maxY; (void)minZ; (void)maxZ;  // the bounds are part of fitslam_params (lut_min/max_*)
    ctx_->check(fitslam_fisher_generate_lookup(ctx_->get(), lookup_path_.empty() ? nullptr : lookup_path_.c_str()), "fitslam_fisher_generate_lookup");
  }
  void loadLookupTable() {  // FisherInfoManager.cpp:231-262: throws when the file is missing (:238-241)
    const fitslam_status st = fitslam_fisher_load_lookup(ctx_->get(), lookup_path_.c_str());
    if (st == FITSLAM_E_NO_LOOKUP_TABLE) throw std::runtime_error("Cannot load lookup table. Please generate it first.");
    ctx_->check(st, "fitslam_fisher_load_lookup");
  }
  inline float getFactorFromNum(int num, float s) { return std::exp(1 - std::pow(num, s)); }  // FisherInfoManager.hpp:102-106
  inline bool getVoxelCoordinate(float &vx, float &vy, float &vz, float lx, float ly, float lz) {  // FisherInfoManager.hpp:108-123
    const double corrected_step = step_max;  // subSampleVoxelUntil_m = -1: the fine branch is dead
    if (corrected_step == 0) return false;
    vx = std::round(lx * (1 / corrected_step)) * corrected_step;
    vy = std::round(ly * (1 / corrected_step)) * corrected_step;
    vz = std::round(lz * (1 / corrected_step)) * corrected_step;
    return true;
  }
  bool isPoseSafe(geometry_msgs::msg::Pose &given_pose, bool exhaustiveSearch, float &information) {  // FisherInfoManager.cpp:39-115
    (void)exhaustiveSearch;  // the resident cloud is always searched exhaustively
    fitslam_pose p{given_pose.position.x, given_pose.position.y, given_pose.position.z, given_pose.orientation.x,
                   given_pose.orientation.y, given_pose.orientation.z, given_pose.orientation.w};
    uint8_t safe = 0;
    const fitslam_status st = fitslam_score_fisher(ctx_->get(), &p, 1, &information, &safe, nullptr);
    if (st == FITSLAM_E_NO_MAP_POINTS) return false;  // "service unavailable" (:52-57)
    ctx_->check(st, "fitslam_score_fisher");
    return safe != 0;
  }
  // batch form: one launch for all candidate poses
  void arePosesSafe(const std::vector<geometry_msgs::msg::Pose> &poses, std::vector<float> &information, std::vector<uint8_t> &safe) {
    std::vector<fitslam_pose> p(poses.size());
    for (size_t i = 0; i < poses.size(); ++i)
      p[i] = fitslam_pose{poses[i].position.x, poses[i].position.y, poses[i].position.z, poses[i].orientation.x,
                          poses[i].orientation.y, poses[i].orientation.z, poses[i].orientation.w};
    information.assign(poses.size(), 0.0f);
    safe.assign(poses.size(), 0);
    if (!poses.empty())
      ctx_->check(fitslam_score_fisher(ctx_->get(), p.data(), (uint32_t)p.size(), information.data(), safe.data(), nullptr), "fitslam_score_fisher");
  }
  // the landmark service itself (slam_msgs/srv/GetLandmarksInView, exhaustive search): positions of the visible map
  // points in the frame of `pose`, in map-point order, widened to double as geometry_msgs/Point carries them
  // (ORBW/src/rgbd/rgbd-slam-node.cpp:406-429 over orb_slam3_interface.cpp:188-317)
  std::vector<geometry_msgs::msg::Point> getLandmarksInView(const geometry_msgs::msg::Pose &pose) {
    const fitslam_pose p{pose.position.x, pose.position.y, pose.position.z, pose.orientation.x, pose.orientation.y, pose.orientation.z,
                         pose.orientation.w};
    uint32_t n = 0;
    fitslam_status st = fitslam_landmarks_in_view(ctx_->get(), &p, nullptr, 0, &n);
    if (st != FITSLAM_OK && st != FITSLAM_E_CAPACITY) ctx_->check(st, "fitslam_landmarks_in_view");
    std::vector<float> xyz(3 * (size_t)n);
    if (n) ctx_->check(fitslam_landmarks_in_view(ctx_->get(), &p, xyz.data(), n, &n), "fitslam_landmarks_in_view");
    std::vector<geometry_msgs::msg::Point> out(n);
    for (uint32_t i = 0; i < n; ++i) { out[i].x = xyz[3 * i]; out[i].y = xyz[3 * i + 1]; out[i].z = xyz[3 * i + 2]; }
    return out;
  }
private:
  CudaContextPtr ctx_;
  std::string lookup_path_;
  float step_max = 0.3f;
};

// ---- FrontierRoadMap, the edge test only (RE/include/roadmap_explorer/planners/FrontierRoadmap.hpp:173-188) -------------
// isConnectable (RE/src/planners/FrontierRoadmap.cpp:671-694) decides the roadmap's edges: one traced ray per node pair.
// The node store, A* and publishing stay with the caller; the batch form scores every candidate pair of a node in one launch.
class FrontierRoadMap {
public:
  FrontierRoadMap(CudaContextPtr ctx, double radius_to_decide_edges) : ctx_(ctx), RADIUS_TO_DECIDE_EDGES(radius_to_decide_edges) {
    max_connection_length_ = RADIUS_TO_DECIDE_EDGES * 1.5;  // FrontierRoadmap.cpp:32
  }
  bool isConnectable(const FrontierPtr &f1, const FrontierPtr &f2) {
    const double seg[4] = {f1->getGoalPoint().x, f1->getGoalPoint().y, f2->getGoalPoint().x, f2->getGoalPoint().y};
    uint8_t ok = 0;
    ctx_->check(fitslam_roadmap_connectable(ctx_->get(), seg, 1, RADIUS_TO_DECIDE_EDGES, &ok, nullptr), "fitslam_roadmap_connectable");
    return ok != 0;
  }
  // f against every node of `others` (populateNodes' inner loop, FrontierRoadmap.cpp:316-345): one launch
  std::vector<uint8_t> areConnectable(const FrontierPtr &f, const std::vector<FrontierPtr> &others) {
    std::vector<double> seg(4 * others.size());
    for (size_t i = 0; i < others.size(); ++i) {
      seg[4 * i] = f->getGoalPoint().x; seg[4 * i + 1] = f->getGoalPoint().y;
      seg[4 * i + 2] = others[i]->getGoalPoint().x; seg[4 * i + 3] = others[i]->getGoalPoint().y;
    }
    std::vector<uint8_t> ok(others.size(), 0);
    if (!others.empty())
      ctx_->check(fitslam_roadmap_connectable(ctx_->get(), seg.data(), (uint32_t)others.size(), RADIUS_TO_DECIDE_EDGES, ok.data(), nullptr),
                  "fitslam_roadmap_connectable");
    return ok;
  }
private:
  CudaContextPtr ctx_;
  double max_connection_length_;
  double RADIUS_TO_DECIDE_EDGES;
};

// ---- LethalMarker (include/frontier_exploration/nav2_plugins/lethal_marker.hpp:33-104; FS2 keepout_layer.hpp) ------------
// The costmap layer keeps every marked area and writes 253 over all of them on each updateCosts (lethal_marker.cpp:303-327).
// Here the areas are kept as their specifications (as FS2's zone_specs_, keepout_layer.cpp:201-203) and one launch stamps
// them into the context's resident costmap; call updateCosts() after every updateCostmap / updateOccupancyGrid, which is
// where the layered costmap would run this layer.
class LethalMarker {
public:
  explicit LethalMarker(CudaContextPtr ctx) : ctx_(ctx) {}
  void addNewMarkedArea(double center_wx, double center_wy, double radius) {  // lethal_marker.cpp:219-228
    areas_.push_back(fitslam_lethal_area{center_wx, center_wy, radius, 0.0, 0, 0});
  }
  void addNewMarkedAreaFOV(double center_wx, double center_wy, double angleYaw, double height) {  // :230-239
    areas_.push_back(fitslam_lethal_area{center_wx, center_wy, height, angleYaw, 1, 0});
  }
  void setEnabled(bool enabled) { enabled_ = enabled; }  // the `enabled` parameter (:250-275)
  void clear() { areas_.clear(); }
  size_t numAdditions() const { return areas_.size(); }
  // returns the number of areas whose centre lies on the map (the others are skipped, :222-223)
  uint32_t updateCosts() {
    if (!enabled_ || areas_.empty()) return 0;  // "Costs returned because not enabled." (:307-311)
    uint32_t marked = 0;
    ctx_->check(fitslam_mark_lethal(ctx_->get(), areas_.data(), (uint32_t)areas_.size(), &marked), "fitslam_mark_lethal");
    return marked;
  }
private:
  CudaContextPtr ctx_;
  std::vector<fitslam_lethal_area> areas_;
  bool enabled_ = true;
};

// ---- SensorSimulator, the marking step (RE/include/roadmap_explorer/SensorSimulator.hpp; SensorSimulator.cpp:193-246) ----
// timerCallback's ray loop over the int8 cells of latest_map_ / explored_map_ (nav_msgs/OccupancyGrid::data); subscriptions,
// publishing, map saving and the geometry-change reprojection stay with the caller.
class SensorSimulator {
public:
  explicit SensorSimulator(CudaContextPtr ctx) : ctx_(ctx) { fitslam_default_sensor_params(&params_); }
  fitslam_sensor_params &parameters() { return params_; }  // sensorSimulator.* of RE/src/Parameters.cpp:227-240
  // explored (size_x * size_y cells) is updated in place; returns the number of rays cast
  uint32_t markRays(const std::vector<int8_t> &latest, std::vector<int8_t> &explored, uint32_t size_x, uint32_t size_y, double resolution,
                    double origin_x, double origin_y, const geometry_msgs::msg::Pose &base_pose, double base_yaw) {
    if (latest.size() != (size_t)size_x * size_y || explored.size() != latest.size())
      throw std::runtime_error("SensorSimulator::markRays: grids do not match the geometry");
    uint32_t n = 0;
    ctx_->check(fitslam_sensor_mark_rays(ctx_->get(), latest.data(), explored.data(), size_x, size_y, resolution, origin_x, origin_y,
                                         base_pose.position.x, base_pose.position.y, base_yaw, &params_, &n), "fitslam_sensor_mark_rays");
    return n;
  }
private:
  CudaContextPtr ctx_;
  fitslam_sensor_params params_;
};

}  // namespace frontier_exploration

namespace roadmap_explorer {  // the maintained package uses the same classes under this namespace (RE/include/roadmap_explorer/)
using frontier_exploration::CudaContext;
using frontier_exploration::CudaContextPtr;
using frontier_exploration::FisherInformationManager;
using frontier_exploration::Frontier;
using frontier_exploration::FrontierCostCalculator;
using frontier_exploration::FrontierPtr;
using frontier_exploration::FrontierRoadMap;
using frontier_exploration::FrontierSearch;
using frontier_exploration::LethalMarker;
using frontier_exploration::SensorSimulator;
}  // namespace roadmap_explorer
