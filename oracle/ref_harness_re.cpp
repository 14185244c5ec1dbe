// ref_harness_re.cpp -- C entry points over the roadmap-explorer package's own code (the REFERENCE's second copy of the helpers,
// namespace roadmap_explorer), compiled from /root/reference where it lies into oracle/_ref/libfitslam_ref_re.so (oracle/Makefile):
// RE/src/Frontier.cpp, the costmap half of RE/src/Helpers.cpp (bresenham2D, getTracedCells, nhood*, nearestFreeCell),
// FrontierRoadMap::isConnectable (RE/src/planners/FrontierRoadmap.cpp:671-694) and SensorSimulator::markRay
// (RE/src/SensorSimulator.cpp:212-246), against stand-in headers (oracle/ref_shims, oracle/ref_shims_re).  A library of its own:
// the package's global `Frontier` class has the same name as the deprecated package's.  Test infrastructure only
// (tests/test_ref_pin.py pins the oracle's rows 8f-2 and 8f-4 with it).
#include <cstdint>
#include <cstring>
#include <memory>
#include <vector>

#include <roadmap_explorer/SensorSimulator.hpp>
#include <roadmap_explorer/planners/FrontierRoadmap.hpp>

using nav2_costmap_2d::Costmap2D;

extern "C" {

// FrontierRoadMap::isConnectable for n node pairs (x1, y1, x2, y2) on one costmap; out[i] = 1 if connectable
void ref_re_connectable(const uint8_t *cells, uint32_t W, uint32_t H, double res, double ox, double oy, double radius_to_decide_edges,
                        const double *pairs, uint32_t n, uint8_t *out) {
  Costmap2D map(W, H, res, ox, oy);
  std::memcpy(map.getCharMap(), cells, (size_t)W * H);
  roadmap_explorer::FrontierRoadMap roadmap(&map, radius_to_decide_edges);
  for (uint32_t i = 0; i < n; ++i) {
    FrontierPtr a = std::make_shared<Frontier>(), b = std::make_shared<Frontier>();
    a->setGoalPoint(pairs[4 * i], pairs[4 * i + 1]);
    b->setGoalPoint(pairs[4 * i + 2], pairs[4 * i + 3]);
    out[i] = roadmap.isConnectable(a, b) ? 1 : 0;
  }
}

// The ray fan of SensorSimulator::timerCallback (RE/src/SensorSimulator.cpp:193-202: for (a = min; a <= max; a += delta)
// markRay(pose, yaw + a)) around SensorSimulator::markRay itself.  explored is updated in place; returns the rays cast.
uint32_t ref_re_mark_rays(const int8_t *latest, int8_t *explored, uint32_t W, uint32_t H, double res, double ox, double oy, double px,
                          double py, double base_yaw, double min_angle, double max_angle, double delta_theta, double max_range) {
  auto lat = std::make_shared<nav_msgs::msg::OccupancyGrid>();
  auto exp = std::make_shared<nav_msgs::msg::OccupancyGrid>();
  for (auto &g : {lat, exp}) {
    g->info.resolution = (float)res;
    g->info.width = W;
    g->info.height = H;
    g->info.origin.position.x = ox;
    g->info.origin.position.y = oy;
  }
  lat->data.assign(latest, latest + (size_t)W * H);
  exp->data.assign(explored, explored + (size_t)W * H);
  parameterInstance.set("sensorSimulator.sensor_max_range", max_range);
  roadmap_explorer::SensorSimulator sim(lat, exp);
  geometry_msgs::msg::Pose pose;
  pose.position.x = px;
  pose.position.y = py;
  uint32_t rays = 0;
  for (double a = min_angle; a <= max_angle; a += delta_theta) {
    sim.markRay(pose, base_yaw + a);
    ++rays;
  }
  std::memcpy(explored, exp->data.data(), (size_t)W * H);
  return rays;
}

}  // extern "C"
