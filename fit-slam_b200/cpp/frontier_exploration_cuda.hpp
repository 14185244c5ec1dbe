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
//   FrontierCostsManager        include/frontier_exploration/FrontierCostsManager.hpp:40-95, src/FrontierCostsManager.cpp:47-222
//   CostAssigner (+ request / response)  include/frontier_exploration/CostAssigner.hpp:43-100, src/CostAssigner.cpp:29-168
//   FullPathOptimizer (selection part)   include/frontier_exploration/FullPathOptimizer.hpp:24-130, src/FullPathOptimizer.cpp:157-227,342-352,548-639
//   Fisher helper free functions         include/frontier_exploration/Helpers.hpp:144-165, src/Helpers.cpp:333-449
#pragma once
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <functional>
#include <mutex>
#include <unordered_map>
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
struct PoseStamped { struct { std::string frame_id; } header; Pose pose; };
struct Point32 { float x = 0, y = 0, z = 0; };
struct Polygon { std::vector<Point32> points; };
struct PolygonStamped { struct { std::string frame_id; } header; Polygon polygon; };
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
  double getSizeInMetersX() const { return (size_x_ - 1 + 0.5) * resolution_; }
  double getSizeInMetersY() const { return (size_y_ - 1 + 0.5) * resolution_; }
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
  bool hasPathLength() const { return (bool)path_length_ || !isAchievable(); }  // not in the reference: a planner has filled it in
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
  // ---- the service as it really runs on a multi-keyframe atlas (orb_slam3_interface.cpp:188-317): keyframes in service order
  // with GetPoseInverse() and their map-point lists, the per-keyframe prefilter unless exhaustive, and the cutoff after more
  // than maxLandmarks landmarks (1000 in rgbd-slam-node.cpp:413)
  void setKeyframes(const std::vector<fitslam_keyframe> &keyframes, const std::vector<uint32_t> &point_index) {
    ctx_->check(fitslam_set_keyframes(ctx_->get(), keyframes.data(), (uint32_t)keyframes.size(), point_index.data(), point_index.size()),
                "fitslam_set_keyframes");
  }
  std::vector<geometry_msgs::msg::Point> getLandmarksInView(const geometry_msgs::msg::Pose &pose, float max_dist_pose_observation,
                                                            float max_angle_pose_observation, bool exhaustive_search,
                                                            int maxLandmarks = 1000) {
    const fitslam_pose p{pose.position.x, pose.position.y, pose.position.z, pose.orientation.x, pose.orientation.y, pose.orientation.z,
                         pose.orientation.w};
    uint32_t n = 0, checked = 0;
    fitslam_status st = fitslam_landmarks_in_view_keyframes(ctx_->get(), &p, maxLandmarks, max_dist_pose_observation,
                                                            max_angle_pose_observation, exhaustive_search ? 1 : 0, nullptr, 0, &n, &checked);
    if (st != FITSLAM_OK && st != FITSLAM_E_CAPACITY) ctx_->check(st, "fitslam_landmarks_in_view_keyframes");
    std::vector<float> xyz(3 * (size_t)n);
    if (n)
      ctx_->check(fitslam_landmarks_in_view_keyframes(ctx_->get(), &p, maxLandmarks, max_dist_pose_observation, max_angle_pose_observation,
                                                      exhaustive_search ? 1 : 0, xyz.data(), n, &n, &checked),
                  "fitslam_landmarks_in_view_keyframes");
    std::vector<geometry_msgs::msg::Point> out(n);
    for (uint32_t i = 0; i < n; ++i) { out[i].x = xyz[3 * i]; out[i].y = xyz[3 * i + 1]; out[i].z = xyz[3 * i + 2]; }
    return out;
  }
  // isPoseSafe with the request fields the reference fills (FisherInfoManager.cpp:48-57) behind that service
  bool isPoseSafe(geometry_msgs::msg::Pose &given_pose, bool exhaustiveSearch, float max_dist_pose_observation, float &information,
                  int maxLandmarks = 1000) {
    const fitslam_pose p{given_pose.position.x, given_pose.position.y, given_pose.position.z, given_pose.orientation.x,
                         given_pose.orientation.y, given_pose.orientation.z, given_pose.orientation.w};
    uint8_t safe = 0;
    ctx_->check(fitslam_score_fisher_keyframes(ctx_->get(), &p, maxLandmarks, max_dist_pose_observation, exhaustiveSearch ? 1 : 0,
                                               &information, &safe, nullptr, nullptr),
                "fitslam_score_fisher_keyframes");
    return safe != 0;
  }
private:
  CudaContextPtr ctx_;
  std::string lookup_path_;
  float step_max = 0.3f;
};

// ---- Fisher helper free functions (Helpers.hpp:144-165, Helpers.cpp:333-449) -------------------------------------------
// With FITSLAM_HAVE_EIGEN the reference's own definitions can be compiled against real Eigen; otherwise these tiny
// fixed-size stand-ins carry the same signatures.  The arithmetic follows Eigen 3.4's evaluation order for fixed-size float
// expressions (3-term dot products as x0*y0 + (x1*y1 + x2*y2), Matrix3f::inverse through cofactors), which is what makes
// computeInformationOfPointLocal(voxel centre, I) equal the lookup table the device generates, bit for bit
// (tests/cpp/shim_harness.cpp checks that against fitslam_fisher_lookup_copy).
#ifndef FITSLAM_HAVE_EIGEN
}  // namespace frontier_exploration
namespace Eigen {
template <typename T, int R, int C>
struct Matrix {
  T m[R][C];
  Matrix() { for (int i = 0; i < R; ++i) for (int j = 0; j < C; ++j) m[i][j] = T(0); }
  Matrix(T x, T y, T z) { static_assert(R == 3 && C == 1, "3-vector"); m[0][0] = x; m[1][0] = y; m[2][0] = z; }
  T &operator()(int i, int j) { return m[i][j]; }
  const T &operator()(int i, int j) const { return m[i][j]; }
  T &operator()(int i) { return m[i][0]; }
  const T &operator()(int i) const { return m[i][0]; }
  static Matrix Identity() { Matrix r; for (int i = 0; i < (R < C ? R : C); ++i) r.m[i][i] = T(1); return r; }
  T trace() const {  // Eigen's unrolled redux: halves, then x0 + (x1 + x2) inside a half of three
    static_assert(R == C, "square");
    if (R == 6) return (m[0][0] + (m[1][1] + m[2][2])) + (m[3][3] + (m[4][4] + m[5][5]));
    T t = m[0][0];
    for (int i = 1; i < R; ++i) t += m[i][i];
    return t;
  }
};
using Matrix3f = Matrix<float, 3, 3>;
using Vector3f = Matrix<float, 3, 1>;
}  // namespace Eigen
namespace frontier_exploration {
#endif
namespace detail {
inline float dot3(float a0, float b0, float a1, float b1, float a2, float b2) { return a0 * b0 + (a1 * b1 + a2 * b2); }
inline Eigen::Matrix3f inverse3(const Eigen::Matrix3f &a) {  // Eigen compute_inverse<Matrix3f>: cofactors, determinant from column 0
  auto cof = [&a](int i, int j) {
    const int i1 = (i + 1) % 3, i2 = (i + 2) % 3, j1 = (j + 1) % 3, j2 = (j + 2) % 3;
    return a(i1, j1) * a(i2, j2) - a(i1, j2) * a(i2, j1);
  };
  const float c0 = cof(0, 0), c1 = cof(1, 0), c2 = cof(2, 0);
  const float invdet = 1.0f / dot3(c0, a(0, 0), c1, a(1, 0), c2, a(2, 0));
  Eigen::Matrix3f r;
  r(0, 0) = c0 * invdet; r(0, 1) = c1 * invdet; r(0, 2) = c2 * invdet;
  for (int i = 1; i < 3; ++i)
    for (int j = 0; j < 3; ++j) r(i, j) = cof(j, i) * invdet;
  return r;
}
}  // namespace detail
inline Eigen::Matrix3f getSkewMatrix(const Eigen::Vector3f &v) {  // Helpers.cpp:333-340
  Eigen::Matrix3f s;
  s(0, 0) = 0; s(0, 1) = -v(2); s(0, 2) = v(1);
  s(1, 0) = v(2); s(1, 1) = 0; s(1, 2) = -v(0);
  s(2, 0) = -v(1); s(2, 1) = v(0); s(2, 2) = 0;
  return s;
}
inline Eigen::Matrix<float, 3, 6> computeJacobianForPointLocal(Eigen::Vector3f &p) {  // Helpers.cpp:397-417
  const float n = std::sqrt(detail::dot3(p(0), p(0), p(1), p(1), p(2), p(2)));
  const float a = 1 / n, s = 1 / (n * n * n);
  Eigen::Matrix3f df;
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) df(i, j) = a * (i == j ? 1.0f : 0.0f) - (s * p(i)) * p(j);
  Eigen::Matrix<float, 3, 6> right;
  const Eigen::Matrix3f skew = getSkewMatrix(p);
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) {
      right(i, j) = (float)((-1.0) * (i == j ? 1.0f : 0.0f));
      right(i, 3 + j) = skew(i, j);
    }
  Eigen::Matrix<float, 3, 6> J;
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 6; ++j) J(i, j) = detail::dot3(df(i, 0), right(0, j), df(i, 1), right(1, j), df(i, 2), right(2, j));
  return J;
}
inline Eigen::Matrix<float, 6, 6> computeFIM(Eigen::Matrix<float, 3, 6> &J, Eigen::Matrix3f &Q) {  // Helpers.cpp:419-422
  const Eigen::Matrix3f Qi = detail::inverse3(Q);
  Eigen::Matrix<float, 6, 6> F;
  for (int i = 0; i < 6; ++i) {
    float q[3];
    for (int k = 0; k < 3; ++k) q[k] = detail::dot3(J(0, i), Qi(0, k), J(1, i), Qi(1, k), J(2, i), Qi(2, k));
    for (int j = 0; j < 6; ++j) F(i, j) = detail::dot3(q[0], J(0, j), q[1], J(1, j), q[2], J(2, j));
  }
  return F;
}
inline float computeInformationOfPointLocal(Eigen::Vector3f &p, Eigen::Matrix3f Q) {  // Helpers.cpp:440-449
  auto jac = computeJacobianForPointLocal(p);
  auto fim = computeFIM(jac, Q);
  return fim.trace();
}

// ---- FrontierCostsManager (FrontierCostsManager.hpp:40-95) ------------------------------------------------------------
struct FrontierHash {  // Frontier.hpp:163-180
  size_t operator()(const FrontierPtr &key) const { return generateUID(key->getGoalPoint().x, key->getGoalPoint().y); }
};
struct FrontierGoalPointEquality {  // Frontier.hpp:139-143
  bool operator()(const FrontierPtr &a, const FrontierPtr &b) const { return a->getGoalPoint().x == b->getGoalPoint().x && a->getGoalPoint().y == b->getGoalPoint().y && a->getGoalPoint().z == b->getGoalPoint().z; }
};
class FrontierCostsManager {
public:
  FrontierCostsManager(nav2_costmap_2d::Costmap2D *costmap, CudaContextPtr ctx)
      : costmap_(costmap), ctx_(ctx), costCalculator_(std::make_shared<FrontierCostCalculator>(costmap, ctx)) {}
  // assignCosts (FrontierCostsManager.cpp:47-213).  costTypes per frontier: "ArrivalInformation" and a distance method.
  // "EuclideanDistance" is setPlanForFrontierEuclidean (CostCalculator.cpp:445-485) on the host (the reference's own libm).
  // "RoadmapPlannerDistance" / "A*PlannerDistance" are the reference's planners (out of scope here): a frontier that asks
  // for one must already carry the planner's result (setPathLength / setPathLengthInM / setPathHeading, or
  // setAchievability(false)); one that carries nothing gets the Euclidean values.  The arrival information of the whole
  // list is one raycast launch, the U1 block one launch (bit-exact weighted costs).  map_data is unused, as in the reference.
  bool assignCosts(std::vector<FrontierPtr> &frontier_list, std::vector<double> polygon_xy_min_max, geometry_msgs::msg::Pose start_pose_w,
                   std::shared_ptr<void> map_data, std::vector<std::vector<std::string>> &costTypes) {
    (void)map_data;
    costCalculator_->reset();
    costCalculator_->setMaxArrivalInformation();
    if (frontier_list.size() == 0) return false;       // "No frontiers found from frontier search."
    if (polygon_xy_min_max.size() <= 0) return false;  // "Frontier cannot be selected, no polygon."
    for (size_t i = 0; i < frontier_list.size(); ++i)  // findDuplicates (:25-45): the same pointer twice
      for (size_t j = i + 1; j < frontier_list.size(); ++j)
        if (frontier_list[i] == frontier_list[j]) throw std::runtime_error("Duplicate frontiers found.");
    const size_t n = frontier_list.size();
    std::vector<uint8_t> blacklisted(n, 0);
    {
      std::lock_guard<std::mutex> lock(blacklist_mutex_);
      for (size_t i = 0; i < n; ++i) blacklisted[i] = frontier_blacklist_.count(frontier_list[i]) > 0;
    }
    auto has = [&costTypes](size_t i, const char *what) {
      return i < costTypes.size() && std::find(costTypes[i].begin(), costTypes[i].end(), std::string(what)) != costTypes[i].end();
    };
    // arrival information: one launch over the list (IMPORTANT in the reference too: before the planning); the results go to
    // the frontiers that ask for it, and the scored list stays resident for the U1 launch below
    std::vector<fitslam_frontier> in(n);
    for (size_t i = 0; i < n; ++i) {
      in[i] = fitslam_frontier{};
      in[i].goal_x = frontier_list[i]->getGoalPoint().x;
      in[i].goal_y = frontier_list[i]->getGoalPoint().y;
      in[i].size = frontier_list[i]->getSize();
      unsigned int mx = 0, my = 0;
      costmap_->worldToMap(in[i].goal_x, in[i].goal_y, mx, my);
      in[i].goal_mx = mx;
      in[i].goal_my = my;
    }
    std::vector<fitslam_arrival> arr(n);
    ctx_->check(fitslam_score_arrival(ctx_->get(), in.data(), (uint32_t)n, polygon_xy_min_max.data(), arr.data(), nullptr, nullptr), "fitslam_score_arrival");
    for (size_t i = 0; i < n; ++i) {
      if (blacklisted[i] || !has(i, "ArrivalInformation")) continue;
      frontier_list[i]->setArrivalInformation((double)arr[i].information);
      frontier_list[i]->setGoalOrientation(arr[i].theta_s_star);
      if (!arr[i].achievable) frontier_list[i]->setAchievability(false);  // CostCalculator.cpp:77-82, 112-117
    }
    std::vector<double> len(n, std::numeric_limits<double>::max()), head(n, std::numeric_limits<double>::max());
    for (size_t i = 0; i < n; ++i) {
      FrontierPtr &f = frontier_list[i];
      if (blacklisted[i]) {  // :77-86
        f->setArrivalInformation(0.0);
        f->setGoalOrientation(0.0);
        f->setFisherInformation(0.0);
        f->setPathLength(std::numeric_limits<double>::max());
        f->setPathLengthInM(std::numeric_limits<double>::max());
        f->setWeightedCost(std::numeric_limits<double>::max());
        continue;
      }
      const bool planner = has(i, "A*PlannerDistance") || has(i, "RoadmapPlannerDistance");
      if (!(planner && f->hasPathLength())) setPlanForFrontierEuclidean(start_pose_w, f);
      if (f->isAchievable()) { len[i] = f->getPathLength(); head[i] = f->getPathHeading(); }
      costCalculator_->recomputeNormalizationFactors(f);
    }
    // U1 (:121-205) on the device (a frontier without a plan, length DBL_MAX, counts as not achievable); its two throws come
    // back as statuses
    std::vector<double> wc(n), ua(n), ud(n);
    const fitslam_status st = fitslam_normalise_select(ctx_->get(), len.data(), head.data(), blacklisted.data(), (uint32_t)n, wc.data(), ua.data(),
                                                       ud.data(), nullptr, nullptr, nullptr);
    if (st == FITSLAM_E_ARRIVAL_UTILITY || st == FITSLAM_E_COST_OUT_OF_BOUNDS) {
      const std::string msg = fitslam_last_error(ctx_->get());  // "... (frontier K)"
      const size_t at = msg.rfind("(frontier ");
      const size_t k = at == std::string::npos ? 0 : (size_t)std::strtoul(msg.c_str() + at + 10, nullptr, 10);
      if (st == FITSLAM_E_COST_OUT_OF_BOUNDS) throw std::runtime_error("Cost out of bounds");
      throw std::runtime_error("ARRIVAL UTILITY ERROR! MORE THAN 1. It is: " + std::to_string(frontier_list[std::min(k, n - 1)]->getArrivalInformation()));
    }
    ctx_->check(st, "fitslam_normalise_select");
    for (size_t i = 0; i < n; ++i) {
      if (blacklisted[i]) continue;
      frontier_list[i]->setWeightedCost(wc[i]);
      frontier_list[i]->setCost("arrival_gain_utility", ua[i]);
      frontier_list[i]->setCost("distance_utility", ud[i]);
    }
    return true;
  }
  void setFrontierBlacklist(std::vector<FrontierPtr> &blacklist) {  // :215-222
    std::lock_guard<std::mutex> lock(blacklist_mutex_);
    for (auto frontier : blacklist) frontier_blacklist_[frontier] = true;
  }
  std::shared_ptr<FrontierCostCalculator> getCostCalcPtr() { return costCalculator_; }
  // setPlanForFrontierEuclidean (CostCalculator.cpp:445-485); robot yaw = tf2 getRPY of the pose's quaternion (GeometryUtils.hpp:22-32)
  static void setPlanForFrontierEuclidean(geometry_msgs::msg::Pose start_pose_w, FrontierPtr &goal_point_w) {
    const auto unachievable = [&goal_point_w]() {
      goal_point_w->setAchievability(false);
      goal_point_w->setPathLength(std::numeric_limits<double>::max());
      goal_point_w->setPathLengthInM(std::numeric_limits<double>::max());
      goal_point_w->setPathHeading(std::numeric_limits<double>::max());
      goal_point_w->setFisherInformation(0);
    };
    if (goal_point_w->isAchievable() == false) return unachievable();
    const auto start_point_w = start_pose_w.position;
    const double length_to_frontier = sqrt(pow(start_point_w.x - goal_point_w->getGoalPoint().x, 2) + pow(start_point_w.y - goal_point_w->getGoalPoint().y, 2));
    if (length_to_frontier < 0.5) return unachievable();
    goal_point_w->setAchievability(true);
    double robot_yaw = yawOf(start_pose_w.orientation);
    if (robot_yaw < 0) robot_yaw = robot_yaw + (M_PI * 2);
    double goal_yaw = atan2(goal_point_w->getGoalPoint().y - start_pose_w.position.y, goal_point_w->getGoalPoint().x - start_pose_w.position.x);
    if (goal_yaw < 0) goal_yaw = goal_yaw + (M_PI * 2);
    double path_heading = std::abs(robot_yaw - goal_yaw);
    if (path_heading > M_PI) path_heading = (2 * M_PI) - path_heading;
    goal_point_w->setPathLength(length_to_frontier);
    goal_point_w->setPathLengthInM(length_to_frontier);
    goal_point_w->setPathHeading(path_heading);
    goal_point_w->setFisherInformation(0.0);
  }
  static double yawOf(const geometry_msgs::msg::Quaternion &q) {  // tf2::Matrix3x3(q).getRPY yaw (the first Euler solution)
    const double d = q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w, s = 2.0 / d;
    const double ys = q.y * s, zs = q.z * s, wy = q.w * ys, wz = q.w * zs, xy = q.x * ys, xz = q.x * zs, yy = q.y * ys, zz = q.z * zs;
    const double m00 = 1.0 - (yy + zz), m10 = xy + wz, m20 = xz - wy;
    if (std::fabs(m20) >= 1) return 0.0;
    const double pitch = -std::asin(m20);
    return std::atan2(m10 / std::cos(pitch), m00 / std::cos(pitch));
  }
private:
  nav2_costmap_2d::Costmap2D *costmap_;
  CudaContextPtr ctx_;
  std::shared_ptr<FrontierCostCalculator> costCalculator_;
  std::unordered_map<FrontierPtr, bool, FrontierHash, FrontierGoalPointEquality> frontier_blacklist_;
  std::mutex blacklist_mutex_;
};

// ---- CostAssigner (CostAssigner.hpp:43-100) ------------------------------------------------------------------------------
struct GetFrontierCostsRequest {
  geometry_msgs::msg::PoseStamped start_pose;
  std::vector<FrontierPtr> frontier_list;
  std::vector<std::vector<double>> every_frontier;
  std::vector<FrontierPtr> prohibited_frontiers;
};
struct GetFrontierCostsResponse {
  bool success = false;
  std::vector<FrontierPtr> frontier_list;
  std::vector<double> frontier_costs;
  std::vector<double> frontier_distances;
  std::vector<double> frontier_arrival_information;
  std::vector<double> frontier_path_information;
};
class CostAssigner {
public:
  // (the reference takes the Costmap2DROS and reads its layered costmap: CostAssigner.cpp:8-27)
  CostAssigner(nav2_costmap_2d::Costmap2D *costmap, CudaContextPtr ctx, std::vector<std::string> methods = {"EuclideanDistance", "ArrivalInformation"})
      : costmap_(costmap), frontierCostsManager_(std::make_shared<FrontierCostsManager>(costmap, ctx)), chosenMethods_(std::move(methods)) {}
  // updateBoundaryPolygon (CostAssigner.cpp:121-167): an empty boundary means the whole map, with the reference's own corner
  // values (origin and size in metres, as it mixes them)
  bool updateBoundaryPolygon(geometry_msgs::msg::PolygonStamped &explore_boundary) {
    for (const auto &point32 : explore_boundary.polygon.points) polygon_.points.push_back(point32);
    if (polygon_.points.empty()) {
      geometry_msgs::msg::Point32 temp;
      temp.x = (float)costmap_->getOriginX();
      temp.y = (float)costmap_->getOriginY();
      polygon_.points.push_back(temp);
      temp.y = (float)costmap_->getSizeInMetersY();
      polygon_.points.push_back(temp);
      temp.x = (float)costmap_->getSizeInMetersX();
      polygon_.points.push_back(temp);
      temp.y = (float)costmap_->getOriginY();
      polygon_.points.push_back(temp);
    }
    double min_x = std::numeric_limits<double>::infinity(), min_y = std::numeric_limits<double>::infinity();
    double max_x = -std::numeric_limits<double>::infinity(), max_y = -std::numeric_limits<double>::infinity();
    for (const auto &point : polygon_.points) {
      min_x = std::min(min_x, (double)point.x);
      min_y = std::min(min_y, (double)point.y);
      max_x = std::max(max_x, (double)point.x);
      max_y = std::max(max_y, (double)point.y);
    }
    polygon_xy_min_max_.push_back(min_x);
    polygon_xy_min_max_.push_back(min_y);
    polygon_xy_min_max_.push_back(max_x);
    polygon_xy_min_max_.push_back(max_y);
    return true;
  }
  // getFrontierCosts (CostAssigner.cpp:73-119): the response carries the SAME pointers in the SAME order
  bool getFrontierCosts(std::shared_ptr<GetFrontierCostsRequest> requestData, std::shared_ptr<GetFrontierCostsResponse> resultData) {
    frontierCostsManager_->setFrontierBlacklist(requestData->prohibited_frontiers);
    std::vector<std::vector<std::string>> costTypes(requestData->frontier_list.size(), chosenMethods_);
    const bool costsResult = frontierCostsManager_->assignCosts(requestData->frontier_list, polygon_xy_min_max_, requestData->start_pose.pose,
                                                                nullptr, costTypes);
    if (costsResult == false) {
      resultData->success = false;
      return resultData->success;
    }
    resultData->success = true;
    std::vector<FrontierPtr> frontiers_list;
    std::vector<double> frontier_costs, frontier_distances, frontier_arrival_information;
    for (auto &frontier : requestData->frontier_list) {
      frontiers_list.push_back(frontier);
      frontier_costs.push_back(frontier->getWeightedCost());
      frontier_distances.push_back(frontier->getPathLengthInM());
      frontier_arrival_information.push_back(frontier->getArrivalInformation());
    }
    resultData->frontier_list = frontiers_list;
    resultData->frontier_costs = frontier_costs;
    resultData->frontier_distances = frontier_distances;
    resultData->frontier_arrival_information = frontier_arrival_information;
    if (resultData->frontier_list != requestData->frontier_list) throw std::runtime_error("Lists are not SAME!");
    return resultData->success;
  }
  std::shared_ptr<FrontierCostsManager> getCostManagerPtr() { return frontierCostsManager_; }
  const std::vector<double> &polygonXYMinMax() const { return polygon_xy_min_max_; }
private:
  nav2_costmap_2d::Costmap2D *costmap_;
  std::shared_ptr<FrontierCostsManager> frontierCostsManager_;
  std::vector<std::string> chosenMethods_;
  geometry_msgs::msg::Polygon polygon_;
  std::vector<double> polygon_xy_min_max_;
};

// ---- FullPathOptimizer, the selection part (FullPathOptimizer.hpp:24-130) -----------------------------------------------
enum PathSafetyStatus { SAFE, UNSAFE, UNDETERMINED };  // FullPathOptimizer.hpp:24-28
struct SortedFrontiers {                              // :70-75
  std::vector<FrontierPtr> local_frontiers;
  std::vector<FrontierPtr> global_frontiers;
  FrontierPtr closest_global_frontier;
};
class FullPathOptimizer {
public:
  static constexpr double NUM_FRONTIERS_IN_LOCAL_AREA = 5.0, LOCAL_FRONTIER_SEARCH_RADIUS = 12.0, BLACKLISTING_CIRCLE_RADIUS = 1.7;  // :13-19
  explicit FullPathOptimizer(std::shared_ptr<FisherInformationManager> fisher = nullptr, bool exhaustiveLandmarkSearch = true)
      : fisher_information_manager_(std::move(fisher)), exhaustiveLandmarkSearch_(exhaustiveLandmarkSearch) {}
  bool isInBlacklistedRegion(const FrontierPtr &frontier) {  // :118-130
    for (auto &zone : circularBlacklistCenters_) {
      const double dx = frontier->getGoalPoint().x - zone->getGoalPoint().x, dy = frontier->getGoalPoint().y - zone->getGoalPoint().y;
      if (sqrt(pow(dx, 2) + pow(dy, 2)) < BLACKLISTING_CIRCLE_RADIUS) return true;
    }
    return false;
  }
  void blacklistFrontier(FrontierPtr &frontier) { circularBlacklistCenters_.push_back(frontier); }
  // getFilteredFrontiersN (FullPathOptimizer.cpp:157-227): host logic over the F pointers, with the library's own std::sort
  // (unstable on equal path lengths, as the reference's)
  void getFilteredFrontiersN(std::vector<FrontierPtr> &frontier_list, size_t n, SortedFrontiers &sortedFrontiers, geometry_msgs::msg::PoseStamped &robotP) {
    (void)robotP;
    std::vector<FrontierPtr> all_frontiers;
    for (const auto &frontier : frontier_list)
      if (frontier->isAchievable() && !frontier->isBlacklisted() && !isInBlacklistedRegion(frontier)) all_frontiers.push_back(frontier);
    std::sort(all_frontiers.begin(), all_frontiers.end(), [](const FrontierPtr &a, const FrontierPtr &b) { return a->getPathLengthInM() < b->getPathLengthInM(); });
    bool global_assigned = false;
    if (all_frontiers.size() == 0) return;  // "No achievable frontiers exist during sorting."
    size_t counter = 1;
    bool need_to_pop = false;
    for (const auto &frontier : all_frontiers) {
      if (frontier->getPathLengthInM() <= LOCAL_FRONTIER_SEARCH_RADIUS && counter <= n) {
        sortedFrontiers.local_frontiers.push_back(frontier);
        sortedFrontiers.closest_global_frontier = frontier;
        need_to_pop = true;
      } else if (frontier->getPathLengthInM() <= LOCAL_FRONTIER_SEARCH_RADIUS && counter > n) {
        sortedFrontiers.closest_global_frontier = frontier;
        need_to_pop = false;
      } else if (frontier->getPathLengthInM() > LOCAL_FRONTIER_SEARCH_RADIUS) {
        if (!global_assigned) {
          sortedFrontiers.closest_global_frontier = frontier;
          global_assigned = true;
        }
        need_to_pop = false;
        sortedFrontiers.global_frontiers.push_back(frontier);
      }
      counter++;
    }
    if (need_to_pop) sortedFrontiers.local_frontiers.pop_back();
    if (sortedFrontiers.global_frontiers.size() == 0 && sortedFrontiers.local_frontiers.size() == 0)
      sortedFrontiers.global_frontiers.push_back(sortedFrontiers.closest_global_frontier);
  }
  PathSafetyStatus isRobotPoseSafe(geometry_msgs::msg::Pose &robotPose) {  // :342-352
    float information;
    const bool safety_value = fisher_information_manager_->isPoseSafe(robotPose, exhaustiveLandmarkSearch_, information);
    return safety_value ? SAFE : UNSAFE;
  }
  // The order in which the local frontiers are visited is the reference's travelling-salesman step (getBestFullPath, out of
  // scope: SURVEY.md 2.2); a caller that has it plugs it in here.  bestPath[0] is the robot, bestPath[1] the next goal.
  // Default: the local frontiers in order of path length.
  std::function<bool(SortedFrontiers &, std::vector<FrontierPtr> &, geometry_msgs::msg::PoseStamped &)> getBestFullPath;
  // getNextGoal (FullPathOptimizer.cpp:548-639)
  PathSafetyStatus getNextGoal(std::vector<FrontierPtr> &frontier_list, FrontierPtr &nextFrontier, size_t n, geometry_msgs::msg::PoseStamped &robotP,
                               geometry_msgs::msg::PoseStamped &robotPFI, bool use_fi) {
    (void)n;  // the reference ignores it too and uses NUM_FRONTIERS_IN_LOCAL_AREA (:552)
    SortedFrontiers sortedFrontiers;
    getFilteredFrontiersN(frontier_list, (size_t)NUM_FRONTIERS_IN_LOCAL_AREA, sortedFrontiers, robotP);
    FrontierPtr zeroFrontier = std::make_shared<Frontier>();
    std::vector<FrontierPtr> bestFrontierWaypoint;
    if (sortedFrontiers.local_frontiers.size() == 0) {
      if (sortedFrontiers.global_frontiers.size() >= 1) {
        nextFrontier = sortedFrontiers.closest_global_frontier;
        if (use_fi && isRobotPoseSafe(robotPFI.pose) == UNSAFE) return PathSafetyStatus::UNSAFE;  // "Dead reckoning!!"
        return PathSafetyStatus::SAFE;
      }
      nextFrontier = zeroFrontier;
      return PathSafetyStatus::UNDETERMINED;
    }
    bool have_path;
    if (getBestFullPath) {
      have_path = getBestFullPath(sortedFrontiers, bestFrontierWaypoint, robotP);
    } else {
      bestFrontierWaypoint.push_back(zeroFrontier);  // the robot's place in the path
      for (auto &f : sortedFrontiers.local_frontiers) bestFrontierWaypoint.push_back(f);
      have_path = true;
    }
    if (!have_path || bestFrontierWaypoint.size() < 2) {
      nextFrontier = zeroFrontier;
      return PathSafetyStatus::UNDETERMINED;
    }
    nextFrontier = bestFrontierWaypoint[1];
    if (use_fi && isRobotPoseSafe(robotPFI.pose) == UNSAFE) return PathSafetyStatus::UNSAFE;
    return PathSafetyStatus::SAFE;
  }
private:
  std::shared_ptr<FisherInformationManager> fisher_information_manager_;
  bool exhaustiveLandmarkSearch_;
  std::vector<FrontierPtr> circularBlacklistCenters_;
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
using frontier_exploration::CostAssigner;
using frontier_exploration::FrontierCostCalculator;
using frontier_exploration::FrontierCostsManager;
using frontier_exploration::FrontierPtr;
using frontier_exploration::FullPathOptimizer;
using frontier_exploration::GetFrontierCostsRequest;
using frontier_exploration::GetFrontierCostsResponse;
using frontier_exploration::PathSafetyStatus;
using frontier_exploration::SortedFrontiers;
using frontier_exploration::FrontierRoadMap;
using frontier_exploration::FrontierSearch;
using frontier_exploration::LethalMarker;
using frontier_exploration::SensorSimulator;
}  // namespace roadmap_explorer
