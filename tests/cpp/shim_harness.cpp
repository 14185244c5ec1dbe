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
  // Fisher helpers (Helpers.cpp:397-449): trace(FIM)(p) = 2 + 2 / |p|^2, skew matrix layout
  {
    Eigen::Vector3f p(1.0f, 0.0f, 0.0f);
    if (computeInformationOfPointLocal(p, Eigen::Matrix3f::Identity()) != 4.0f) ++bad;
    Eigen::Vector3f q(3.0f, 0.0f, 0.0f);
    if (std::fabs(computeInformationOfPointLocal(q, Eigen::Matrix3f::Identity()) - (2.0f + 2.0f / 9.0f)) > 1e-6f) ++bad;
    const Eigen::Matrix3f sk = getSkewMatrix(Eigen::Vector3f(1.0f, 2.0f, 3.0f));
    if (sk(0, 1) != -3.0f || sk(0, 2) != 2.0f || sk(1, 0) != 3.0f || sk(1, 2) != -1.0f || sk(2, 0) != -2.0f || sk(2, 1) != 1.0f || sk(0, 0) != 0.0f) ++bad;
    Eigen::Matrix<float, 3, 6> J = computeJacobianForPointLocal(q);
    Eigen::Matrix3f I = Eigen::Matrix3f::Identity();
    if (computeFIM(J, I).trace() != computeInformationOfPointLocal(q, I)) ++bad;
  }
  // getFilteredFrontiersN known answers (FullPathOptimizer.cpp:157-227)
  {
    auto make = [](int k) {
      std::vector<FrontierPtr> v;
      for (int i = 0; i < k; ++i) {
        auto f = std::make_shared<Frontier>();
        f->setGoalPoint(10.0 + i, 0.0);
        f->setUID(generateUID(10.0 + i, 0.0));
        f->setSize(5);
        f->setPathLengthInM(1.0 + ((i * 3) % k));  // lengths 1..k in a scrambled order
        v.push_back(f);
      }
      return v;
    };
    geometry_msgs::msg::PoseStamped robotP;
    FullPathOptimizer opt;
    SortedFrontiers s7;
    auto v7 = make(7);
    opt.getFilteredFrontiersN(v7, 5, s7, robotP);
    if (s7.local_frontiers.size() != 5 || !s7.global_frontiers.empty() || s7.closest_global_frontier->getPathLengthInM() != 7.0) ++bad;
    for (size_t i = 0; i < s7.local_frontiers.size(); ++i)
      if (s7.local_frontiers[i]->getPathLengthInM() != 1.0 + i) ++bad;
    SortedFrontiers s5;
    auto v5 = make(5);
    opt.getFilteredFrontiersN(v5, 5, s5, robotP);
    if (s5.local_frontiers.size() != 4 || s5.closest_global_frontier->getPathLengthInM() != 5.0) ++bad;  // the last one is popped (:221-222)
    // a blacklisted region (1.7 m around a frontier) and an unachievable frontier are filtered out
    auto v3 = make(3);
    v3[0]->setAchievability(false);
    opt.blacklistFrontier(v3[1]);
    SortedFrontiers s3;
    opt.getFilteredFrontiersN(v3, 5, s3, robotP);  // v3[2] is 1 m from the blacklisted centre: inside the circle too
    if (!s3.local_frontiers.empty() || !s3.global_frontiers.empty()) ++bad;
    // getNextGoal without local or global frontiers: a zero frontier and UNDETERMINED (:583-588)
    FrontierPtr next;
    if (opt.getNextGoal(v3, next, 5, robotP, robotP, false) != UNDETERMINED || !next->isFrontierNull()) ++bad;
  }
  // setPlanForFrontierEuclidean (CostCalculator.cpp:445-485)
  {
    geometry_msgs::msg::Pose start;
    start.orientation.z = std::sin(0.35);  // yaw 0.7
    start.orientation.w = std::cos(0.35);
    auto f = std::make_shared<Frontier>();
    f->setGoalPoint(3.0, 4.0);
    FrontierCostsManager::setPlanForFrontierEuclidean(start, f);
    if (f->getPathLength() != 5.0 || std::fabs(f->getPathHeading() - (std::atan2(4.0, 3.0) - 0.7)) > 1e-12 || !f->isAchievable()) ++bad;
    auto near = std::make_shared<Frontier>();
    near->setGoalPoint(0.3, 0.0);
    FrontierCostsManager::setPlanForFrontierEuclidean(start, near);
    if (near->isAchievable() || near->getPathLength() != std::numeric_limits<double>::max()) ++bad;
  }
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
    // the same service over a keyframe atlas: two keyframes at the identity pose, even points then odd points then the first
    // 10 again; a cutoff of 20 landmarks stops after the first keyframe (orb_slam3_interface.cpp:270-271)
    std::vector<fitslam_keyframe> kfs(3);
    std::vector<uint32_t> index;
    for (int k = 0; k < 3; ++k) {
      for (int i = 0; i < 9; ++i) kfs[k].Rwc[i] = (i % 4 == 0) ? 1.0f : 0.0f;
      kfs[k].twc[0] = kfs[k].twc[1] = kfs[k].twc[2] = 0.0f;
      kfs[k].first = (uint32_t)index.size();
      if (k < 2) { for (uint32_t i = (uint32_t)k; i < P; i += 2) index.push_back(i); }
      else { for (uint32_t i = 0; i < 10 && i < P; ++i) index.push_back(i); }
      kfs[k].count = (uint32_t)index.size() - kfs[k].first;
    }
    fisher.setKeyframes(kfs, index);
    const auto all = fisher.getLandmarksInView(pose, 5.0f, 0.0f, true, 1 << 30);
    const auto cut = fisher.getLandmarksInView(pose, 5.0f, 0.0f, true, 20);
    double ax = 0, cx = 0;
    for (const auto &p : all) ax += p.x;
    for (const auto &p : cut) cx += p.x;
    float info_kf = 0;
    const bool safe_kf = fisher.isPoseSafe(pose, true, 5.0f, info_kf, 20);
    std::printf("landmarks_kf %zu %a %zu %a %a %d\n", all.size(), ax, cut.size(), cx, (double)info_kf, safe_kf ? 1 : 0);
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
  // ---- cost assignment and goal selection in the behaviour tree's order (ExplorationBT.cpp:371-423, 517-637):
  // CostAssigner::updateBoundaryPolygon / getFrontierCosts, then FullPathOptimizer::getNextGoal
  {
    ctx->updateCostmap(costmap);  // (the lethal marker above stamped the resident copy)
    std::vector<FrontierPtr> list = search.searchFrom(robot);
    CostAssigner assigner(&costmap, ctx);
    geometry_msgs::msg::PolygonStamped boundary;  // the map's corners, as float32 (geometry_msgs/Point32)
    const float bx[4] = {(float)g[1], (float)g[1], (float)(g[1] + W * g[0]), (float)(g[1] + W * g[0])};
    const float by[4] = {(float)g[2], (float)(g[2] + H * g[0]), (float)(g[2] + H * g[0]), (float)g[2]};
    for (int k = 0; k < 4; ++k) {
      geometry_msgs::msg::Point32 p32;
      p32.x = bx[k];
      p32.y = by[k];
      boundary.polygon.points.push_back(p32);
    }
    assigner.updateBoundaryPolygon(boundary);
    auto req = std::make_shared<GetFrontierCostsRequest>();
    auto res = std::make_shared<GetFrontierCostsResponse>();
    req->start_pose.pose.position = robot;
    req->frontier_list = list;
    req->every_frontier = search.getAllFrontiers();
    if (list.size() > 4) req->prohibited_frontiers.push_back(list[3]);  // one blacklisted frontier (:77-86)
    const bool ok = assigner.getFrontierCosts(req, res);
    std::printf("costs %d %zu\n", ok ? 1 : 0, res->frontier_list.size());
    for (size_t i = 0; i < res->frontier_list.size(); ++i) {
      const FrontierPtr &f = res->frontier_list[i];
      if (f != list[i]) std::printf("NOT THE SAME POINTER %zu\n", i);
      std::printf("C %a %a %a %d\n", res->frontier_costs[i], res->frontier_distances[i], res->frontier_arrival_information[i], f->isAchievable() ? 1 : 0);
    }
    auto fisher_mgr = std::make_shared<FisherInformationManager>(ctx);
    FullPathOptimizer optimizer(fisher_mgr);
    geometry_msgs::msg::PoseStamped robotP;
    robotP.pose.position = robot;
    robotP.pose.position.z = 0.3;
    FrontierPtr next;
    const PathSafetyStatus status = optimizer.getNextGoal(list, next, 5, robotP, robotP, true);
    float robot_info = 0;
    const bool robot_safe = fisher_mgr->isPoseSafe(robotP.pose, true, robot_info);
    std::printf("next %d %a %a %d %a\n", (int)status, next->isFrontierNull() ? 0.0 : next->getGoalPoint().x, next->isFrontierNull() ? 0.0 : next->getGoalPoint().y,
                robot_safe ? 1 : 0, (double)robot_info);
    // the reference's exceptions: the same frontier twice, and a response list that is not the request's
    try {
      std::vector<FrontierPtr> twice{list[0], list[0]};
      std::vector<std::vector<std::string>> types(2, {"EuclideanDistance", "ArrivalInformation"});
      assigner.getCostManagerPtr()->assignCosts(twice, assigner.polygonXYMinMax(), robotP.pose, nullptr, types);
      std::printf("duplicates NOT DETECTED\n");
    } catch (const std::runtime_error &e) {
      std::printf("duplicates %s\n", e.what());
    }
    // the Fisher helper free functions against the device's lookup table: every entry along two lines of voxels
    int32_t dims[3], org[3];
    ctx->check(fitslam_fisher_lookup_dims(ctx->get(), dims, org), "fitslam_fisher_lookup_dims");
    std::vector<float> lut((size_t)dims[0] * dims[1] * dims[2]);
    ctx->check(fitslam_fisher_lookup_copy(ctx->get(), lut.data(), lut.size()), "fitslam_fisher_lookup_copy");
    int lut_bad = 0, lut_n = 0;
    for (int ix = 1; ix < dims[0]; ++ix)
      for (int iy = 0; iy < dims[1]; iy += 7) {
        const int iz = (ix * 5 + iy) % dims[2];
        float vx = 0, vy = 0, vz = 0;
        fisher_mgr->getVoxelCoordinate(vx, vy, vz, (float)((ix + org[0]) * 0.3), (float)((iy + org[1]) * 0.3), (float)((iz + org[2]) * 0.3));
        Eigen::Vector3f p(vx, vy, vz);
        const float want = lut[((size_t)ix * dims[1] + iy) * dims[2] + iz];
        const float got = computeInformationOfPointLocal(p, Eigen::Matrix3f::Identity());
        ++lut_n;
        if (std::memcmp(&want, &got, 4) != 0) ++lut_bad;
      }
    std::printf("lut %d %d\n", lut_n, lut_bad);
  }
  // robot off the map: empty list, no exception (FrontierSearch.cpp:28-32)
  robot.x = g[1] - 1.0;
  std::printf("offmap %zu\n", search.searchFrom(robot).size());
  return 0;
}
