// shim_harness.cpp -- drives the reference-shaped C++ classes (fit-slam_b200/cpp/frontier_exploration_cuda.hpp) the
// way the behaviour tree drives the reference (DEP/src/ExplorationBT.cpp:166-197, 410, FullPathOptimizer.cpp:346):
// FrontierSearch::reset / searchFrom / getAllFrontiers, FrontierCostCalculator::setArrivalInformationForFrontier,
// FisherInformationManager::isPoseSafe.  tests/test_cpp_shim.py builds it with g++ and compares its output with the oracle.
//
//   shim_harness --no-gpu              host-only checks (no compute call): Frontier getters throw on unset fields, the
//                                      Costmap2D stand-in, generateUID, the library links and answers
//   shim_harness <input.bin>           input: u32 W, H, n_points; f64 res, ox, oy, rx, ry; W*H bytes; n_points*3 f32
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "frontier_exploration_cuda.hpp"

using namespace frontier_exploration;

static int host_only() {
  int bad = 0;
  Frontier f;
  try { (void)f.getUID(); ++bad; } catch (const std::runtime_error &) {}
  try { (void)f.getGoalPoint(); ++bad; } catch (const std::runtime_error &) {}
  try { (void)f.getWeightedCost(); ++bad; } catch (const std::runtime_error &) {}
  if (!f.isAchievable() || f.isBlacklisted()) ++bad;  // defaults (Frontier.cpp:5-20)
  f.setGoalPoint(1.5, -2.25);
  f.setUID(generateUID(1.5, -2.25));
  f.setGoalOrientation(1.0);
  if (f.getGoalOrientation().z != std::sin(0.5) || f.getGoalOrientation().w != std::cos(0.5)) ++bad;
  if (generateUID(0.0, 0.0) != 0) ++bad;  // std::hash<double>(0.0) == 0
  nav2_costmap_2d::Costmap2D cm(10, 8, 0.05, -1.0, 2.0, 255);
  unsigned int mx, my;
  double wx, wy;
  if (cm.worldToMap(-1.01, 2.0, mx, my)) ++bad;
  if (!cm.worldToMap(-1.0 + 9.5 * 0.05, 2.0 + 7.5 * 0.05, mx, my) || mx != 9 || my != 7) ++bad;
  cm.mapToWorld(0, 0, wx, wy);
  if (wx != -1.0 + 0.5 * 0.05 || wy != 2.0 + 0.5 * 0.05) ++bad;
  fitslam_params p;
  fitslam_default_params(&p);
  if (p.max_frontier_cluster_size != 20 || fitslam_abi_version() != FITSLAM_ABI_VERSION) ++bad;
  if (std::strlen(fitslam_status_string(FITSLAM_E_NO_LOOKUP_TABLE)) == 0) ++bad;
  std::printf("host-only checks: %d failed\n", bad);
  return bad;
}

int main(int argc, char **argv) {
  if (argc < 2) return 2;
  if (!std::strcmp(argv[1], "--no-gpu")) return host_only();
  FILE *fp = std::fopen(argv[1], "rb");
  if (!fp) return 2;
  uint32_t hdr[3];
  double g[5];
  if (std::fread(hdr, 4, 3, fp) != 3 || std::fread(g, 8, 5, fp) != 5) return 2;
  const uint32_t W = hdr[0], H = hdr[1], P = hdr[2];
  nav2_costmap_2d::Costmap2D costmap(W, H, g[0], g[1], g[2]);
  if (std::fread(costmap.getCharMap(), 1, (size_t)W * H, fp) != (size_t)W * H) return 2;
  std::vector<float> pts(3 * (size_t)P);
  if (P && std::fread(pts.data(), 4, pts.size(), fp) != pts.size()) return 2;
  std::fclose(fp);

  auto ctx = std::make_shared<CudaContext>();
  FrontierSearch search(costmap, ctx);
  FrontierCostCalculator calc(&costmap, ctx);
  FisherInformationManager fisher(ctx);
  fisher.generateLookupTable(0, 21, -14.722f, 14.722f, -14.722f, 14.722f);
  fisher.setMapPoints(pts.data(), P);

  geometry_msgs::msg::Point robot;
  robot.x = g[3];
  robot.y = g[4];
  search.reset();
  std::vector<FrontierPtr> frontiers = search.searchFrom(robot);
  auto every = search.getAllFrontiers();
  std::vector<double> poly{g[1], g[2], g[1] + W * g[0], g[2] + H * g[0]};
  std::printf("limits %a\n", calc.setMaxArrivalInformation());
  for (auto &f : frontiers) calc.setArrivalInformationForFrontier(f, poly);  // one at a time, as the reference's loop does
  std::printf("frontiers %zu cells %zu\n", frontiers.size(), every.size());
  for (auto &f : frontiers) {
    geometry_msgs::msg::Pose pose;
    pose.position = f->getGoalPoint();
    pose.position.z = 0.3;
    pose.orientation = f->getGoalOrientation();
    float information = 0;
    const bool safe = fisher.isPoseSafe(pose, true, information);
    std::printf("F %a %a %d %zu %a %a %d %a %d\n", f->getGoalPoint().x, f->getGoalPoint().y, f->getSize(), f->getUID(),
                f->getArrivalInformation(), f->getGoalOrientation().z, f->isAchievable() ? 1 : 0, (double)information, safe ? 1 : 0);
  }
  // roadmap edges: the first frontier against all the others (one launch), and one pair through the single-pair form
  if (frontiers.size() > 1) {
    roadmap_explorer::FrontierRoadMap roadmap(ctx, 2.6);
    std::vector<FrontierPtr> others(frontiers.begin() + 1, frontiers.end());
    const std::vector<uint8_t> ok = roadmap.areConnectable(frontiers[0], others);
    std::printf("edges");
    for (uint8_t v : ok) std::printf(" %d", (int)v);
    std::printf("\nedge1 %d\n", roadmap.isConnectable(frontiers[0], frontiers[1]) ? 1 : 0);
  }
  double sx = 0, sy = 0;
  for (auto &c : every) { sx += c[0]; sy += c[1]; }
  std::printf("cellsum %a %a\n", sx, sy);
  // the landmark service for the first frontier's pose (GetLandmarksInView, exhaustive)
  if (!frontiers.empty()) {
    geometry_msgs::msg::Pose pose;
    pose.position = frontiers[0]->getGoalPoint();
    pose.position.z = 0.3;
    pose.orientation = frontiers[0]->getGoalOrientation();
    const auto lm = fisher.getLandmarksInView(pose);
    double sx = 0, sy = 0, sz = 0;
    for (const auto &p : lm) { sx += p.x; sy += p.y; sz += p.z; }
    std::printf("landmarks %zu %a %a %a\n", lm.size(), sx, sy, sz);
  }
  // sensor simulator: a fan around the robot on an occupancy grid derived from the costmap (254 -> 100, 255 -> -1, else 0)
  {
    std::vector<int8_t> latest((size_t)W * H), explored((size_t)W * H, -1);
    for (size_t i = 0; i < latest.size(); ++i) {
      const unsigned char v = costmap.getCharMap()[i];
      latest[i] = v == 254 ? 100 : (v == 255 ? -1 : 0);
    }
    roadmap_explorer::SensorSimulator sim(ctx);
    geometry_msgs::msg::Pose base;
    base.position.x = g[3];
    base.position.y = g[4];
    const uint32_t n_rays = sim.markRays(latest, explored, W, H, g[0], g[1], g[2], base, 0.4);
    uint64_t seen = 0, wsum = 0;
    for (size_t i = 0; i < explored.size(); ++i)
      if (explored[i] != -1) { ++seen; wsum += (uint64_t)i * (uint64_t)(explored[i] + 2); }
    std::printf("sensor %u %llu %llu\n", n_rays, (unsigned long long)seen, (unsigned long long)wsum);
  }
  // lethal marker: the two calls of ExplorationBT.cpp:621-622 stamped into the resident costmap
  {
    frontier_exploration::LethalMarker marker(ctx);
    marker.addNewMarkedAreaFOV(g[3] + 1.0, g[4], 0.7, 3.5);
    marker.addNewMarkedArea(g[3] - 2.0, g[4] + 1.0, 1.7);
    marker.addNewMarkedArea(g[1] - 5.0, g[2], 1.7);  // off the map: skipped
    const uint32_t marked = marker.updateCosts();
    std::vector<uint8_t> stamped((size_t)W * H);
    ctx->check(fitslam_get_costmap(ctx->get(), stamped.data(), stamped.size()), "fitslam_get_costmap");
    uint64_t n253 = 0, isum = 0, changed = 0;
    for (size_t i = 0; i < stamped.size(); ++i) {
      if (stamped[i] == 253) { ++n253; isum += i; }
      if (stamped[i] != costmap.getCharMap()[i]) ++changed;
    }
    std::printf("lethal %u %llu %llu %llu\n", marked, (unsigned long long)n253, (unsigned long long)isum, (unsigned long long)changed);
  }
  // robot off the map: empty list, no exception (FrontierSearch.cpp:28-32)
  robot.x = g[1] - 1.0;
  std::printf("offmap %zu\n", search.searchFrom(robot).size());
  return 0;
}
