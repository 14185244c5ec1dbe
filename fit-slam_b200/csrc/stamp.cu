// stamp.cu -- the two writers next to the path (SURVEY 8f-3 stamping half, 8f-4):
//   * LethalMarker: fans of Bresenham rays stamped with 253 into the resident costmap
//     (DEP/src/nav2_plugins/lethal_marker.cpp; FS2/fit_slam2_nav2_plugins/plugins/keepout_layer.cpp);
//   * SensorSimulator::markRay: rays that copy the latest occupancy grid into the explored grid up to the first
//     occupied cell (RE/src/SensorSimulator.cpp:193-246).
// Every transcendental (ray end points, ray directions) and every accumulated double (the angle and range loops) is
// evaluated on the host with libm exactly as the reference does and uploaded as a small table; the kernels only do
// integer stepping and IEEE double + - * / (compiled with -fmad=false), so cells are bit-exact.
#include <cmath>
#include <vector>

#include "ctx.h"

namespace fitslam {

// ---------------------------------------------------------------------------------------------- lethal marker
// rayTraceGeneric (lethal_marker.cpp:12-41) with markCells (:242-248) as the visitor: one thread per ray.  All rays
// write the same value, so the order in which overlapping rays reach a cell does not matter.
__global__ void __launch_bounds__(128) lethal_rays_kernel(uint8_t *__restrict__ map, uint32_t W, uint32_t H, uint32_t pitch,
                                                           const int4 *__restrict__ rays, uint32_t n) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int4 r = rays[i];
  int x0 = r.x, y0 = r.y;
  const int x1 = r.z, y1 = r.w;
  const int dx = abs(x1 - x0), dy = abs(y1 - y0);
  const int sx = (x0 < x1) ? 1 : -1, sy = (y0 < y1) ? 1 : -1;
  int err = dx - dy;
  while (true) {
    // `index >= 0 && index < height * width` of the reference: both ends are inside the map, so every step is
    if ((uint32_t)x0 < W && (uint32_t)y0 < H) map[(size_t)y0 * pitch + x0] = 253;
    if (x0 == x1 && y0 == y1) break;
    const int e2 = 2 * err;
    if (e2 > -dy) {
      err -= dy;
      x0 += sx;
    }
    if (e2 < dx) {
      err += dx;
      y0 += sy;
    }
  }
}

// cellEnforceBoundaries (lethal_marker.cpp:5-11)
static inline void cell_enforce_boundaries(int64_t &x, int64_t &y, uint32_t W, uint32_t H) {
  x = (x < 0) ? 0 : x;
  x = (x > (int64_t)W - 1) ? (int64_t)W - 1 : x;
  y = (y < 0) ? 0 : y;
  y = (y > (int64_t)H - 1) ? (int64_t)H - 1 : y;
}

// Costmap2D::worldToMap (nav2_costmap_2d, ROS 2 Humble)
static inline bool world_to_map_host(const fitslam_ctx *c, double wx, double wy, unsigned int &mx, unsigned int &my) {
  if (wx < c->ox || wy < c->oy) return false;
  mx = static_cast<unsigned int>((wx - c->ox) / c->res);
  my = static_cast<unsigned int>((wy - c->oy) / c->res);
  return mx < c->W && my < c->H;
}

// getPointsInSemiCircle with numPoints 360, robot_yaw 0 (lethal_marker.cpp:52-73, called from :226)
static void circle_rays(const fitslam_ctx *c, unsigned int center_x, unsigned int center_y, unsigned int radius_in_cells,
                        std::vector<int4> &rays) {
  const unsigned int numPoints = 360;
  const double robot_yaw = 0;
  for (unsigned int i = 0; i < numPoints; ++i) {
    const double angle = 2.0 * M_PI * i / numPoints;
    int64_t x = static_cast<int64_t>(center_x + radius_in_cells * std::cos(robot_yaw - M_PI_2 + angle));
    int64_t y = static_cast<int64_t>(center_y + radius_in_cells * std::sin(robot_yaw - M_PI_2 + angle));
    cell_enforce_boundaries(x, y, c->W, c->H);
    rays.push_back(make_int4((int)center_x, (int)center_y, (int)x, (int)y));
  }
}

// getPointsInIsoscelesTriangle with numPoints 20, apex angle 45 degrees (lethal_marker.cpp:75-127, called from :237)
static void triangle_rays(const fitslam_ctx *c, unsigned int apex_x, unsigned int apex_y, unsigned int triangle_height,
                          double direction, std::vector<int4> &rays) {
  const unsigned int numPoints = 20;
  const double apex_angle = 45 * M_PI / 180;
  const double base_center_x = apex_x + triangle_height * std::cos(direction);
  const double base_center_y = apex_y + triangle_height * std::sin(direction);
  const double half_base = triangle_height * std::tan(apex_angle / 2.0);
  const double left_x = base_center_x + half_base * std::cos(direction + M_PI_2);
  const double left_y = base_center_y + half_base * std::sin(direction + M_PI_2);
  const double right_x = base_center_x + half_base * std::cos(direction - M_PI_2);
  const double right_y = base_center_y + half_base * std::sin(direction - M_PI_2);
  for (unsigned int i = 0; i < numPoints; ++i) {
    const double t = static_cast<double>(i) / (numPoints - 1);
    const double sample_x = left_x + t * (right_x - left_x);
    const double sample_y = left_y + t * (right_y - left_y);
    int64_t cell_x = static_cast<int64_t>(sample_x);
    int64_t cell_y = static_cast<int64_t>(sample_y);
    cell_enforce_boundaries(cell_x, cell_y, c->W, c->H);
    rays.push_back(make_int4((int)apex_x, (int)apex_y, (int)cell_x, (int)cell_y));
  }
}

// ---------------------------------------------------------------------------------------------- sensor simulator
// markRay (SensorSimulator.cpp:211-246), one warp per ray.  The range loop `for (r = 0; r <= range; r += res)` only
// depends on the parameters: its values come as a table, so the 32 lanes evaluate 32 consecutive steps at once, a
// ballot finds the first step that leaves the map or meets a cell > 80, and the lanes up to it copy their cells.
// Rays that cross copy the same value into the same cell.
struct SensorRays {
  const int8_t *latest;  // rows [row0, row0 + rows) of the grid
  int8_t *explored;
  int w, h, row0, rows;
  double res, ox, oy, px, py;
  const double *dir;    // n_rays x {cos, sin}(base_yaw + a)
  const double *range;  // n_steps values of r
  uint32_t n_rays, n_steps;
};

__global__ void __launch_bounds__(128) sensor_rays_kernel(const SensorRays S) {
  const uint32_t ray = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31u;
  if (ray >= S.n_rays) return;
  const double dx = S.dir[2 * ray], dy = S.dir[2 * ray + 1];
  for (uint32_t base = 0; base < S.n_steps; base += 32) {
    const uint32_t k = base + lane;
    int state = 1;  // 0: copy and go on, 1: stop without a copy (outside the map / past the range), 2: copy and stop
    size_t idx = 0;
    int8_t v = 0;
    if (k < S.n_steps) {
      const double r = S.range[k];
      const double wx = S.px + r * dx;
      const double wy = S.py + r * dy;
      const int mx = static_cast<int>((wx - S.ox) / S.res);
      const int my = static_cast<int>((wy - S.oy) / S.res);
      if (!(mx < 0 || my < 0 || mx >= S.w || my >= S.h) && my >= S.row0 && my < S.row0 + S.rows) {
        idx = (size_t)(my - S.row0) * S.w + mx;
        v = S.latest[idx];
        state = (v > 80) ? 2 : 0;
      }
    }
    const uint32_t stop = __ballot_sync(0xffffffffu, state != 0);
    const uint32_t first = stop ? (uint32_t)__ffs(stop) - 1u : 32u;
    if (lane < first || (lane == first && state == 2)) S.explored[idx] = v;
    if (stop) break;
  }
}

}  // namespace fitslam

using namespace fitslam;

extern "C" {

fitslam_status fitslam_mark_lethal(fitslam_ctx *c, const fitslam_lethal_area *areas, uint32_t n, uint32_t *n_marked) {
  if (!c || (!areas && n)) return fail(c, FITSLAM_E_INVALID, "mark_lethal: null argument");
  if (!c->have_map) return fail(c, FITSLAM_E_NO_COSTMAP, "no costmap set");
  if (n_marked) *n_marked = 0;
  std::vector<int4> rays;
  uint32_t marked = 0;
  for (uint32_t a = 0; a < n; ++a) {
    const fitslam_lethal_area &A = areas[a];
    if (A.kind != 0 && A.kind != 1) return fail(c, FITSLAM_E_INVALID, "mark_lethal: kind must be 0 (circle) or 1 (FOV triangle)");
    if (!std::isfinite(A.wx) || !std::isfinite(A.wy) || !std::isfinite(A.yaw) || !(A.extent >= 0.0) ||
        !(A.extent / c->res < 2147483648.0))
      return fail(c, FITSLAM_E_INVALID, "mark_lethal: bad area");
    unsigned int mx, my;
    if (!world_to_map_host(c, A.wx, A.wy, mx, my)) continue;  // lethal_marker.cpp:222-223, :233-234
    // `auto radius_in_cells = radius / resolution` is a double that the callee takes as unsigned int (:225-226, :236-237)
    const unsigned int cells = static_cast<unsigned int>(A.extent / c->res);
    if (A.kind == 0) circle_rays(c, mx, my, cells, rays);
    else triangle_rays(c, mx, my, cells, A.yaw, rays);
    ++marked;
  }
  if (n_marked) *n_marked = marked;
  if (rays.empty()) return FITSLAM_OK;
  cudaSetDevice(c->device);
  const uint32_t nr = (uint32_t)rays.size();
  FS_CUDA(c, c->stamp_tab.reserve((size_t)nr * 2));  // int4 = 2 doubles
  FS_CUDA(c, cudaMemcpyAsync(c->stamp_tab.p, rays.data(), (size_t)nr * sizeof(int4), cudaMemcpyHostToDevice, c->stream));
  lethal_rays_kernel<<<(nr + 127) / 128, 128, 0, c->stream>>>(c->map.p, c->W, c->H, c->pitch,
                                                              reinterpret_cast<const int4 *>(c->stamp_tab.p), nr);
  ++c->launches;
  FS_CUDA(c, cudaGetLastError());
  FS_CUDA(c, cudaStreamSynchronize(c->stream));  // `rays` is pageable host memory that dies with this call
  return FITSLAM_OK;
}

void fitslam_default_sensor_params(fitslam_sensor_params *p) {
  if (!p) return;
  p->sensor_min_angle = -M_PI / 4;  // RE/src/Parameters.cpp:233-234
  p->sensor_max_angle = M_PI / 4;   // :236-237
  p->angular_resolution = 0.013;    // :227-228
  p->sensor_max_range = 3.0;        // :239-240
}

fitslam_status fitslam_sensor_mark_rays(fitslam_ctx *c, const int8_t *latest, int8_t *explored, uint32_t size_x, uint32_t size_y,
                                        double resolution_in, double origin_x, double origin_y, double robot_x, double robot_y,
                                        double base_yaw, const fitslam_sensor_params *p, uint32_t *n_rays) {
  // `const double res = latest_map_->info.resolution;` (SensorSimulator.cpp:217): nav_msgs/MapMetaData.resolution is a float32,
  // so the reference steps with 0.05f widened, not with 0.05 (a cell's difference here and there; found against the reference's
  // own markRay, tests/test_ref_pin.py)
  const double resolution = static_cast<double>(static_cast<float>(resolution_in));
  if (!c || !latest || !explored || size_x == 0 || size_y == 0 || !(resolution > 0.0))
    return fail(c, FITSLAM_E_INVALID, "sensor_mark_rays: bad argument");
  if ((uint64_t)size_x * size_y >= (1ull << 31)) return fail(c, FITSLAM_E_INVALID, "sensor_mark_rays: more than 2^31 cells");
  fitslam_sensor_params P;
  if (p) P = *p; else fitslam_default_sensor_params(&P);
  if (!std::isfinite(robot_x) || !std::isfinite(robot_y) || !std::isfinite(base_yaw) || !std::isfinite(origin_x) || !std::isfinite(origin_y) ||
      !std::isfinite(P.sensor_min_angle) || !std::isfinite(P.sensor_max_angle) || !std::isfinite(P.sensor_max_range))
    return fail(c, FITSLAM_E_INVALID, "sensor_mark_rays: non-finite argument");
  if (!(P.angular_resolution > 0.0)) return fail(c, FITSLAM_E_INVALID, "sensor_mark_rays: angular_resolution must be positive");
  if ((P.sensor_max_angle - P.sensor_min_angle) / P.angular_resolution > 1e7 || P.sensor_max_range / resolution > 1e7)
    return fail(c, FITSLAM_E_INVALID, "sensor_mark_rays: more than 1e7 rays or steps");
  // the two loops of the reference, with their accumulated doubles (SensorSimulator.cpp:198-201, :226)
  std::vector<double> tab;
  for (double a = P.sensor_min_angle; a <= P.sensor_max_angle; a += P.angular_resolution) {
    const double ray_angle = base_yaw + a;
    tab.push_back(std::cos(ray_angle));
    tab.push_back(std::sin(ray_angle));
  }
  const uint32_t nr = (uint32_t)(tab.size() / 2);
  for (double r = 0.0; r <= P.sensor_max_range; r += resolution) tab.push_back(r);
  const uint32_t ns = (uint32_t)(tab.size() - 2 * (size_t)nr);
  if (n_rays) *n_rays = nr;
  if (nr == 0 || ns == 0) return FITSLAM_OK;
  const bool dev = [&] {
    cudaPointerAttributes a;
    const bool d = cudaPointerGetAttributes(&a, latest) == cudaSuccess && (a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged);
    cudaGetLastError();
    return d;
  }();
  cudaSetDevice(c->device);
  // rows a ray can reach: |wy - robot_y| <= r <= max_range (one row of slack on either side for the roundings)
  int64_t row0 = 0, row1 = (int64_t)size_y - 1;
  if (!dev) {
    const double lo = std::floor((robot_y - P.sensor_max_range - origin_y) / resolution) - 1.0;
    const double hi = std::floor((robot_y + P.sensor_max_range - origin_y) / resolution) + 1.0;
    if (hi < 0.0 || lo > (double)size_y - 1.0) return FITSLAM_OK;  // the robot's own cell is off the map: every ray stops at r = 0
    row0 = lo < 0.0 ? 0 : (int64_t)lo;
    row1 = hi > (double)size_y - 1.0 ? (int64_t)size_y - 1 : (int64_t)hi;
  }
  const size_t rows = (size_t)(row1 - row0 + 1), bytes = rows * size_x;
  SensorRays S;
  if (dev) {
    S.latest = latest;
    S.explored = explored;
  } else {
    FS_CUDA(c, c->occ_stage.reserve(2 * bytes));
    FS_CUDA(c, cudaMemcpyAsync(c->occ_stage.p, latest + (size_t)row0 * size_x, bytes, cudaMemcpyHostToDevice, c->stream));
    FS_CUDA(c, cudaMemcpyAsync(c->occ_stage.p + bytes, explored + (size_t)row0 * size_x, bytes, cudaMemcpyHostToDevice, c->stream));
    S.latest = c->occ_stage.p;
    S.explored = c->occ_stage.p + bytes;
  }
  FS_CUDA(c, c->stamp_tab.reserve(tab.size()));
  FS_CUDA(c, cudaMemcpyAsync(c->stamp_tab.p, tab.data(), tab.size() * sizeof(double), cudaMemcpyHostToDevice, c->stream));
  S.w = (int)size_x;
  S.h = (int)size_y;
  S.row0 = (int)row0;
  S.rows = (int)rows;
  S.res = resolution;
  S.ox = origin_x;
  S.oy = origin_y;
  S.px = robot_x;
  S.py = robot_y;
  S.dir = c->stamp_tab.p;
  S.range = c->stamp_tab.p + 2 * (size_t)nr;
  S.n_rays = nr;
  S.n_steps = ns;
  sensor_rays_kernel<<<(nr * 32 + 127) / 128, 128, 0, c->stream>>>(S);
  ++c->launches;
  FS_CUDA(c, cudaGetLastError());
  if (!dev) FS_CUDA(c, cudaMemcpyAsync(explored + (size_t)row0 * size_x, S.explored, bytes, cudaMemcpyDeviceToHost, c->stream));
  FS_CUDA(c, cudaStreamSynchronize(c->stream));
  return FITSLAM_OK;
}

}  // extern "C"
