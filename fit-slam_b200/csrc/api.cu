// api.cu -- C ABI entry points: lifecycle, inputs, ray-fan tables, arrival-information calls.
// The stage kernels live in raycast.cu / frontier.cu / fisher.cu / select.cu.
#include "ctx.h"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>

using namespace fitslam;

static_assert(sizeof(fitslam_params) == 192, "ABI layout");
static_assert(sizeof(fitslam_frontier) == 40 && sizeof(fitslam_arrival) == 24 && sizeof(fitslam_pose) == 56, "ABI layout");
static_assert(sizeof(fitslam_selection) == 24 && sizeof(fitslam_cycle_result) == 96 && sizeof(ScoreRecord) == 16, "ABI layout");

namespace fitslam {

fitslam_status ensure_pinned(fitslam_ctx *c, size_t bytes) {
  if (bytes <= c->pinned_cap) return FITSLAM_OK;
  if (c->pinned) cudaFreeHost(c->pinned);
  c->pinned = nullptr;
  c->pinned_cap = 0;
  size_t want = bytes + bytes / 2 + 4096;
  FS_CUDA(c, cudaMallocHost(&c->pinned, want));
  c->pinned_cap = want;
  return FITSLAM_OK;
}

// Ray fan geometry and the per-ray end-point offsets.  The theta loop is the reference's own
// (`for (double theta = 0; theta <= 2*M_PI; theta += DELTA_THETA)`, DEP/src/CostCalculator.cpp:36) and
// cos/sin are evaluated once here on the host, because they are functions of the parameters only.
fitslam_status update_ray_shape(fitslam_ctx *c) {
  if (!c->have_map) return fail(c, FITSLAM_E_NO_COSTMAP, "no costmap set");
  if (c->shape_res == c->res) return FITSLAM_OK;
  const fitslam_params &p = c->prm;
  if (!(p.delta_theta > 0.0) || !(p.camera_fov > 0.0) || !(p.max_camera_depth > 0.0))
    return fail(c, FITSLAM_E_INVALID, "delta_theta, camera_fov and max_camera_depth must be positive");
  std::vector<double> tab;
  for (double theta = 0; theta <= (2 * M_PI); theta += p.delta_theta) {
    tab.push_back(p.max_camera_depth * cos(theta));
    tab.push_back(p.max_camera_depth * sin(theta));
    if (tab.size() > 2 * 1024) return fail(c, FITSLAM_E_INVALID, "delta_theta gives more than 1024 rays");
  }
  RayShape s;
  s.n_rays = (uint32_t)(tab.size() / 2);
  const int k = static_cast<int>(p.camera_fov / p.delta_theta);  // :87
  if (k < 1 || (uint32_t)k > s.n_rays) return fail(c, FITSLAM_E_INVALID, "camera_fov / delta_theta out of range");
  s.k = (uint32_t)k;
  s.n_windows = s.n_rays - s.k + 1;
  s.max_length = (unsigned int)(p.max_camera_depth / c->res);  // :28
  s.radius_cells = (uint32_t)std::ceil(p.robot_radius / c->res);  // :77
  s.half = std::max(s.max_length, s.radius_cells);  // Bresenham takes at most max_length steps
  if (2 * s.half + 1 > 256) return fail(c, FITSLAM_E_INVALID, "ray length exceeds the 127-cell staged window");
  FS_CUDA(c, c->ray_dxdy.reserve(tab.size()));
  FS_CUDA(c, cudaMemcpyAsync(c->ray_dxdy.p, tab.data(), tab.size() * sizeof(double), cudaMemcpyHostToDevice, c->stream));
  FS_CUDA(c, cudaStreamSynchronize(c->stream));
  c->shape = s;
  c->shape_res = c->res;
  return FITSLAM_OK;
}

}  // namespace fitslam

extern "C" {

int fitslam_abi_version(void) { return FITSLAM_ABI_VERSION; }

void fitslam_default_params(fitslam_params *p) {
  if (!p) return;
  memset(p, 0, sizeof(*p));
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
  p->lut_min_y = (float)(-8.5 * 1.732);  // GenerateLookupMain.cpp:8, double product narrowed at the call
  p->lut_max_y = (float)(8.5 * 1.732);
  p->lut_min_z = (float)(-8.5 * 1.732);  // GenerateLookupMain.cpp:8, double product narrowed at the call
  p->lut_max_z = (float)(8.5 * 1.732);
  p->fx = (float)432.496042035043;
  p->fy = (float)432.496042035043;
  p->cx = 320.0f;
  p->cy = 240.0f;
  p->img_max_u = 640.0f;
  p->img_max_v = 480.0f;
  p->candidate_pose_z = 0.3;
}

const char *fitslam_status_string(fitslam_status s) {
  switch (s) {
    case FITSLAM_OK: return "ok";
    case FITSLAM_E_INVALID: return "invalid argument";
    case FITSLAM_E_CUDA: return "CUDA error";
    case FITSLAM_E_CAPACITY: return "output buffer too small";
    case FITSLAM_E_NO_COSTMAP: return "no costmap set";
    case FITSLAM_E_ROBOT_OUT_OF_BOUNDS: return "Robot out of costmap bounds, cannot search for frontiers";
    case FITSLAM_E_NO_FRONTIERS: return "No frontiers found from frontier search.";
    case FITSLAM_E_NO_POLYGON: return "Frontier cannot be selected, no polygon.";
    case FITSLAM_E_ARRIVAL_UTILITY: return "ARRIVAL UTILITY ERROR! MORE THAN 1";
    case FITSLAM_E_COST_OUT_OF_BOUNDS: return "Cost out of bounds";
    case FITSLAM_E_NO_LOOKUP_TABLE: return "Cannot load lookup table. Please generate it first.";
    case FITSLAM_E_NO_MAP_POINTS: return "no map points set";
    case FITSLAM_E_IO: return "I/O error";
  }
  return "unknown status";
}

// The context stream carries the critical path of a cycle; the launches that score finished batches of frontiers beside
// it run on lowest-priority streams (frontier.cu), so it takes the highest priority the device offers.
static cudaError_t create_context_stream(cudaStream_t *st) {
  int lo = 0, hi = 0;
  cudaError_t e = cudaDeviceGetStreamPriorityRange(&lo, &hi);
  if (e != cudaSuccess) return e;
  return cudaStreamCreateWithPriority(st, cudaStreamNonBlocking, hi);
}

fitslam_status fitslam_create(const fitslam_params *p, int device, fitslam_ctx **out) {
  if (!out) return FITSLAM_E_INVALID;
  *out = nullptr;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0 || device < 0 || device >= ndev) {
    cudaGetLastError();
    return FITSLAM_E_CUDA;  // no CPU fallback: the path needs a device
  }
  fitslam_ctx *c = new (std::nothrow) fitslam_ctx();
  if (!c) return FITSLAM_E_INVALID;
  if (p) c->prm = *p; else fitslam_default_params(&c->prm);
  const fitslam_params &q = c->prm;
  if (q.max_frontier_cluster_size < 1 || q.max_frontier_cluster_size > 63 || q.min_frontier_cluster_size < 0 ||
      q.lethal_threshold < 1 || q.lethal_threshold > 254) {
    delete c;
    return FITSLAM_E_INVALID;
  }
  c->device = device;
  if (cudaSetDevice(device) != cudaSuccess || create_context_stream(&c->stream) != cudaSuccess) {
    cudaGetLastError();
    delete c;
    return FITSLAM_E_CUDA;
  }
  const char *e = getenv("FITSLAM_RAYCAST_TMA");
  if (e && e[0] == '0') c->use_tma = 0;
  *out = c;
  return FITSLAM_OK;
}

void fitslam_destroy(fitslam_ctx *c) {
  if (!c) return;
  cudaSetDevice(c->device);
  if (c->stream) cudaStreamSynchronize(c->stream);
  destroy_frontier_state(c);
  destroy_fisher_state(c);
  destroy_select_state(c);
  destroy_cycle_state(c);
  destroy_comm_state(c);
  c->map.release();
  c->ray_dxdy.release();
  c->frontiers.release();
  c->arrival.release();
  c->info_ray.release();
  c->win_sums.release();
  c->records.release();
  c->scratch.release();
  if (c->pinned) cudaFreeHost(c->pinned);
  c->edge_seg.release(); c->edge_out.release(); c->occ_stage.release(); c->stamp_tab.release();
  if (c->stream) cudaStreamDestroy(c->stream);
  delete c;
}

const char *fitslam_last_error(const fitslam_ctx *c) { return c ? c->err.c_str() : "null context"; }

fitslam_status fitslam_set_shard(fitslam_ctx *c, int rank, int nranks) {
  if (!c || nranks < 1 || rank < 0 || rank >= nranks) return fail(c, FITSLAM_E_INVALID, "bad shard");
  c->rank = rank;
  c->nranks = nranks;
  return FITSLAM_OK;
}

void *fitslam_stream(fitslam_ctx *c) { return c ? (void *)c->stream : nullptr; }
uint64_t fitslam_launch_count(const fitslam_ctx *c) { return c ? c->launches : 0; }


/* ------------------------------------------------------------------ inputs */
static bool is_device_ptr(const void *p) {
  cudaPointerAttributes a;
  if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
    cudaGetLastError();
    return false;
  }
  return a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged;
}

// ---- costmap ingest
}  // extern "C"

namespace fitslam {

// threshold_occupancy.cpp:33-45, then StaticLayer::interpretValue (see include/fitslam_cuda.h)
__host__ __device__ __forceinline__ uint8_t occupancy_to_cost(int8_t raw, const fitslam_occupancy_params P) {
  int8_t v = raw;
  if (P.lethal_obstacle_threshold >= 0) v = raw >= P.lethal_obstacle_threshold ? (int8_t)100 : (raw == -1 ? (int8_t)-1 : (int8_t)0);
  const unsigned char value = (unsigned char)v;  // the static layer reads the grid as unsigned char
  const unsigned char unknown = (unsigned char)P.unknown_cost_value;
  if (value == unknown) return P.track_unknown_space ? 255 : 0;
  if (value >= (unsigned char)P.lethal_cost_threshold) return 254;
  if (P.trinary_costmap) return 0;
  const double scale = (double)value / (unsigned char)P.lethal_cost_threshold;
  return (uint8_t)(scale * 254);
}

struct OccSimd {  // the trinary cases as per-byte selects (mode 0: none, use occupancy_to_cost per cell)
  int mode;       // 1: threshold node present (signed >= on the raw cell), 2: static layer only (unsigned >= lethal)
  uint32_t unknown4, thr4, if_unknown4, if_lethal4, if_free4;
};

// full map, rows of the device costmap are `pitch` bytes (a multiple of 16): one thread per 16 cells of a row
__global__ void __launch_bounds__(256) occupancy_rows_kernel(const int8_t *__restrict__ in, uint32_t W, uint32_t H, uint32_t in_pitch,
                                                              uint8_t *__restrict__ out, uint32_t out_pitch, uint32_t x_off, uint32_t y_off,
                                                              const fitslam_occupancy_params P, int vec, const OccSimd simd) {
  const uint32_t chunks = (W + 15u) / 16u;
  // grid.x covers the chunks of a row, grid.y strides over the rows: no division per chunk
  for (uint32_t y = blockIdx.y; y < H; y += gridDim.y)
  for (uint32_t cx = blockIdx.x * blockDim.x + threadIdx.x; cx < chunks; cx += gridDim.x * blockDim.x) {
    const uint32_t x = cx * 16u;
    const int8_t *src = in + (size_t)y * in_pitch + x;
    uint8_t *dst = out + (size_t)(y + y_off) * out_pitch + x + x_off;
    if (vec && x + 16u <= W) {  // both rows 16-byte aligned: one 16-byte load, one 16-byte store
      const uint4 q = __ldcs(reinterpret_cast<const uint4 *>(src));  // read once: streaming
      const uint32_t w[4] = {q.x, q.y, q.z, q.w};
      uint32_t o[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        uint32_t r = 0;
        if (simd.mode) {
          // trinary translation, four cells per instruction: the function takes three values, selected by two per-byte
          // compares (with the threshold node: == -1, >= threshold signed; without: == unknown, >= lethal unsigned)
          const uint32_t u = __vcmpeq4(w[k], simd.unknown4);
          const uint32_t t = (simd.mode == 1 ? __vcmpges4(w[k], simd.thr4) : __vcmpgeu4(w[k], simd.thr4)) & ~u;
          r = (u & simd.if_unknown4) | (t & simd.if_lethal4) | (~(u | t) & simd.if_free4);
        } else {
#pragma unroll
          for (int b = 0; b < 4; ++b) r |= (uint32_t)occupancy_to_cost((int8_t)((w[k] >> (8 * b)) & 0xFFu), P) << (8 * b);
        }
        o[k] = r;
      }
      *reinterpret_cast<uint4 *>(dst) = make_uint4(o[0], o[1], o[2], o[3]);
    } else {
      for (uint32_t k = 0; k < 16u && x + k < W; ++k) dst[k] = occupancy_to_cost(src[k], P);
    }
  }
}

}  // namespace fitslam

extern "C" {

void fitslam_default_occupancy_params(fitslam_occupancy_params *p) {
  if (!p) return;
  p->lethal_obstacle_threshold = 85;  // threshold_occupancy.cpp:17
  p->lethal_cost_threshold = 100;     // nav2 StaticLayer defaults; the reference's yaml leaves them alone
  p->unknown_cost_value = -1;
  p->trinary_costmap = 1;
  p->track_unknown_space = 1;         // DEP/params/exploration_params.yaml:25
}

fitslam_status fitslam_set_occupancy_grid(fitslam_ctx *c, const int8_t *data, uint32_t size_x, uint32_t size_y, double resolution,
                                          double origin_x, double origin_y, const fitslam_occupancy_params *p, const uint32_t *dirty) {
  if (!c || !data || size_x == 0 || size_y == 0 || !(resolution > 0.0))
    return fail(c, FITSLAM_E_INVALID, "set_occupancy_grid: bad argument");
  if ((uint64_t)size_x * size_y >= (1ull << 31)) return fail(c, FITSLAM_E_INVALID, "set_occupancy_grid: more than 2^31 cells");
  fitslam_occupancy_params P;
  if (p) P = *p; else fitslam_default_occupancy_params(&P);
  if (P.lethal_cost_threshold < 0 || P.lethal_cost_threshold > 100) return fail(c, FITSLAM_E_INVALID, "set_occupancy_grid: lethal_cost_threshold outside 0..100");
  cudaSetDevice(c->device);
  const bool same_geom = c->have_map && c->W == size_x && c->H == size_y && c->res == resolution &&
                         c->ox == origin_x && c->oy == origin_y;
  const bool dev = is_device_ptr(data);
  uint32_t x0 = 0, y0 = 0, w = size_x, h = size_y;
  if (dirty) {
    if (!same_geom) return fail(c, FITSLAM_E_INVALID, "set_occupancy_grid: dirty rectangle needs a resident map of the same geometry");
    x0 = dirty[0]; y0 = dirty[1]; w = dirty[2]; h = dirty[3];
    if (w == 0 || h == 0) return FITSLAM_OK;
    if ((uint64_t)x0 + w > size_x || (uint64_t)y0 + h > size_y) return fail(c, FITSLAM_E_INVALID, "set_occupancy_grid: dirty rectangle outside the map");
  } else {
    const uint32_t pitch = (size_x + 15u) & ~15u;  // TMA needs 16-byte row strides
    FS_CUDA(c, c->map.reserve((size_t)pitch * size_y + 256));
    if (pitch != size_x) FS_CUDA(c, cudaMemsetAsync(c->map.p, 0, (size_t)pitch * size_y, c->stream));
    c->pitch = pitch;
  }
  // the int8 cells of the rectangle: read in place from a device grid, staged (rows of 16-byte-aligned pitch) from the host
  const int8_t *src = data + (size_t)y0 * size_x + x0;
  uint32_t src_pitch = size_x;
  if (!dev) {
    const uint32_t sp = (w + 15u) & ~15u;
    FS_CUDA(c, c->occ_stage.reserve((size_t)sp * h + 16));
    FS_CUDA(c, cudaMemcpy2DAsync(c->occ_stage.p, sp, src, size_x, w, h, cudaMemcpyHostToDevice, c->stream));
    src = c->occ_stage.p;
    src_pitch = sp;
  }
  const int vec = ((uintptr_t)src % 16 == 0) && (src_pitch % 16 == 0) && (x0 % 16 == 0) ? 1 : 0;
  const uint32_t row_chunks = (w + 15u) / 16u;
  const dim3 grid(std::min<uint32_t>((row_chunks + 255u) / 256u, 64u), std::min<uint32_t>(h, 148u * 32u));
  OccSimd simd{};
  auto rep4 = [](uint8_t b) { return (uint32_t)b * 0x01010101u; };
  if (P.trinary_costmap && P.lethal_obstacle_threshold >= 0 && P.lethal_obstacle_threshold <= 127) {
    // after the threshold node a cell is -1, 100 or 0: three values of the static layer's function
    simd.mode = 1;
    simd.unknown4 = rep4(0xFF);
    simd.thr4 = rep4((uint8_t)P.lethal_obstacle_threshold);
    simd.if_unknown4 = rep4(occupancy_to_cost((int8_t)-1, P));
    simd.if_lethal4 = rep4(occupancy_to_cost((int8_t)P.lethal_obstacle_threshold, P));
    simd.if_free4 = rep4(occupancy_to_cost((int8_t)(P.lethal_obstacle_threshold > 0 ? 0 : -2), P));
    if (P.lethal_obstacle_threshold == 0) simd.mode = 0;  // (every non-negative cell is lethal and 0 is not free: keep the scalar form)
  } else if (P.trinary_costmap && P.lethal_obstacle_threshold < 0 && (uint8_t)P.unknown_cost_value >= (uint8_t)P.lethal_cost_threshold) {
    simd.mode = 2;  // unknown is tested first and lies above the lethal threshold, as with the defaults (255 vs 100)
    simd.unknown4 = rep4((uint8_t)P.unknown_cost_value);
    simd.thr4 = rep4((uint8_t)P.lethal_cost_threshold);
    simd.if_unknown4 = rep4(P.track_unknown_space ? 255 : 0);
    simd.if_lethal4 = rep4(254);
    simd.if_free4 = 0;
  }
  occupancy_rows_kernel<<<grid, 256, 0, c->stream>>>(src, w, h, src_pitch, c->map.p, c->pitch, x0, y0, P, vec, simd);
  ++c->launches;
  FS_CUDA(c, cudaGetLastError());
  FS_CUDA(c, cudaStreamSynchronize(c->stream));  // the caller may reuse `data` on return
  if (dirty) return FITSLAM_OK;
  c->W = size_x;
  c->H = size_y;
  c->res = resolution;
  c->ox = origin_x;
  c->oy = origin_y;
  c->have_map = true;
  if (!same_geom) {
    c->shape_res = -1.0;
    c->limits_set = false;
  }
  return update_ray_shape(c);
}

fitslam_status fitslam_get_costmap(fitslam_ctx *c, uint8_t *out, uint64_t capacity) {
  if (!c || !out) return fail(c, FITSLAM_E_INVALID, "get_costmap: null argument");
  if (!c->have_map) return fail(c, FITSLAM_E_NO_COSTMAP, "no costmap set");
  if (capacity < (uint64_t)c->W * c->H) return fail(c, FITSLAM_E_CAPACITY, "get_costmap: buffer too small");
  cudaSetDevice(c->device);
  FS_CUDA(c, cudaMemcpy2DAsync(out, c->W, c->map.p, c->pitch, c->W, c->H, cudaMemcpyDefault, c->stream));
  FS_CUDA(c, cudaStreamSynchronize(c->stream));
  return FITSLAM_OK;
}

fitslam_status fitslam_set_costmap(fitslam_ctx *c, const uint8_t *data, uint32_t size_x, uint32_t size_y,
                                   double resolution, double origin_x, double origin_y, const uint32_t *dirty) {
  // with a communicator (fitslam_comm_init) the call is collective: rank 0 uploads, the others pass data = NULL and receive
  const bool receive = c && comm_active(c) && c->rank != 0;
  if (!c || (!data && !receive) || size_x == 0 || size_y == 0 || !(resolution > 0.0))
    return fail(c, FITSLAM_E_INVALID, "set_costmap: bad argument");
  if ((uint64_t)size_x * size_y >= (1ull << 31)) return fail(c, FITSLAM_E_INVALID, "set_costmap: more than 2^31 cells");
  cudaSetDevice(c->device);
  const bool same_geom = c->have_map && c->W == size_x && c->H == size_y && c->res == resolution &&
                         c->ox == origin_x && c->oy == origin_y;
  const cudaMemcpyKind kind = (data && is_device_ptr(data)) ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice;
  if (dirty) {
    if (!same_geom) return fail(c, FITSLAM_E_INVALID, "set_costmap: dirty rectangle needs a resident map of the same geometry");
    const uint32_t x0 = dirty[0], y0 = dirty[1], w = dirty[2], h = dirty[3];
    if (w == 0 || h == 0) return FITSLAM_OK;
    if ((uint64_t)x0 + w > size_x || (uint64_t)y0 + h > size_y) return fail(c, FITSLAM_E_INVALID, "set_costmap: dirty rectangle outside the map");
    if (!receive)
      FS_CUDA(c, cudaMemcpy2DAsync(c->map.p + (size_t)y0 * c->pitch + x0, c->pitch, data + (size_t)y0 * size_x + x0, size_x,
                                   w, h, kind, c->stream));
    // (the rows of the rectangle, whole: the ranks' maps agree outside it)
    fitslam_status bst = comm_broadcast(c, c->map.p + (size_t)y0 * c->pitch, (size_t)h * c->pitch);
    if (bst != FITSLAM_OK) return bst;
    FS_CUDA(c, cudaStreamSynchronize(c->stream));  // the caller may reuse `data` on return
    return FITSLAM_OK;
  }
  const uint32_t pitch = (size_x + 15u) & ~15u;  // TMA needs 16-byte row strides
  FS_CUDA(c, c->map.reserve((size_t)pitch * size_y + 256));
  if (pitch != size_x) FS_CUDA(c, cudaMemsetAsync(c->map.p, 0, (size_t)pitch * size_y, c->stream));
  if (!receive) FS_CUDA(c, cudaMemcpy2DAsync(c->map.p, pitch, data, size_x, size_x, size_y, kind, c->stream));
  fitslam_status bst = comm_broadcast(c, c->map.p, (size_t)pitch * size_y);
  if (bst != FITSLAM_OK) return bst;
  FS_CUDA(c, cudaStreamSynchronize(c->stream));
  c->W = size_x;
  c->H = size_y;
  c->pitch = pitch;
  c->res = resolution;
  c->ox = origin_x;
  c->oy = origin_y;
  c->have_map = true;
  if (!same_geom) {
    c->shape_res = -1.0;
    // arrival_info_limits_set_ lives as long as the FrontierCostCalculator; a new map geometry is a new calculator
    c->limits_set = false;
  }
  fitslam_status st = update_ray_shape(c);
  if (st != FITSLAM_OK) return st;
  return FITSLAM_OK;
}

/* ------------------------------------------------------------------ arrival information */
fitslam_status fitslam_ray_shape(const fitslam_ctx *c, uint32_t *n_rays, uint32_t *k, uint32_t *n_windows,
                                 uint32_t *max_length) {
  if (!c || !c->have_map) return FITSLAM_E_NO_COSTMAP;
  if (n_rays) *n_rays = c->shape.n_rays;
  if (k) *k = c->shape.k;
  if (n_windows) *n_windows = c->shape.n_windows;
  if (max_length) *max_length = c->shape.max_length;
  return FITSLAM_OK;
}

fitslam_status fitslam_arrival_limits(fitslam_ctx *c, double *max_gt, double *min_gt, int32_t *max_value) {
  if (!c) return FITSLAM_E_INVALID;
  if (!c->have_map) return fail(c, FITSLAM_E_NO_COSTMAP, "no costmap set");
  cudaSetDevice(c->device);
  fitslam_status st = raycast_limits(c);
  if (st != FITSLAM_OK) return st;
  if (max_gt) *max_gt = c->max_arrival_info_gt;
  if (min_gt) *min_gt = c->min_arrival_info_gt;
  if (max_value) *max_value = c->limits_max_value;
  return FITSLAM_OK;
}

fitslam_status fitslam_score_arrival(fitslam_ctx *c, const fitslam_frontier *frontiers, uint32_t n,
                                     const double poly[4], fitslam_arrival *out, uint16_t *info_along_ray,
                                     int32_t *window_sums) {
  if (!c) return FITSLAM_E_INVALID;
  if (!c->have_map) return fail(c, FITSLAM_E_NO_COSTMAP, "no costmap set");
  if (!poly) return fail(c, FITSLAM_E_NO_POLYGON, "Frontier cannot be selected, no polygon.");
  cudaSetDevice(c->device);
  if (frontiers) {
    FS_CUDA(c, c->frontiers.reserve(n));
    if (n) FS_CUDA(c, cudaMemcpyAsync(c->frontiers.p, frontiers, (size_t)n * sizeof(fitslam_frontier),
                                      is_device_ptr(frontiers) ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, c->stream));
    c->n_frontiers = n;
  } else {
    n = c->n_frontiers;
  }
  if (n == 0) return fail(c, FITSLAM_E_NO_FRONTIERS, "No frontiers found from frontier search.");
  fitslam_status st = raycast_limits(c);  // assignCosts: setMaxArrivalInformation() first (FrontierCostsManager.cpp:67)
  if (st != FITSLAM_OK) return st;
  const RayShape &s = c->shape;
  FS_CUDA(c, c->arrival.reserve(n));
  const size_t n_rec = (size_t)shard_block(c, n) * c->nranks;  // nranks equal blocks (in-place all-gather layout)
  FS_CUDA(c, c->records.reserve(n_rec));
  c->records_rows = n_rec;
  if (info_along_ray) FS_CUDA(c, c->info_ray.reserve((size_t)n * s.n_rays));
  if (window_sums) FS_CUDA(c, c->win_sums.reserve((size_t)n * s.n_windows));
  if (c->nranks > 1) {  // records of the other shards stay zero until the caller merges them
    FS_CUDA(c, cudaMemsetAsync(c->arrival.p, 0, (size_t)n * sizeof(fitslam_arrival), c->stream));
    FS_CUDA(c, cudaMemsetAsync(c->records.p, 0, n_rec * sizeof(ScoreRecord), c->stream));
    if (info_along_ray) FS_CUDA(c, cudaMemsetAsync(c->info_ray.p, 0, (size_t)n * s.n_rays * 2, c->stream));
    if (window_sums) FS_CUDA(c, cudaMemsetAsync(c->win_sums.p, 0, (size_t)n * s.n_windows * 4, c->stream));
  }
  st = raycast_score(c, c->frontiers.p, n, poly, c->arrival.p, info_along_ray ? c->info_ray.p : nullptr,
                     window_sums ? c->win_sums.p : nullptr, c->records.p);
  if (st != FITSLAM_OK) return st;
  if (out) FS_CUDA(c, cudaMemcpyAsync(out, c->arrival.p, (size_t)n * sizeof(fitslam_arrival),
                                      is_device_ptr(out) ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost, c->stream));
  if (info_along_ray) FS_CUDA(c, cudaMemcpyAsync(info_along_ray, c->info_ray.p, (size_t)n * s.n_rays * 2, cudaMemcpyDefault, c->stream));
  if (window_sums) FS_CUDA(c, cudaMemcpyAsync(window_sums, c->win_sums.p, (size_t)n * s.n_windows * 4, cudaMemcpyDefault, c->stream));
  FS_CUDA(c, cudaStreamSynchronize(c->stream));
  return FITSLAM_OK;
}

fitslam_status fitslam_score_records(fitslam_ctx *c, void **device_ptr, uint64_t *bytes) {
  if (!c || !device_ptr || !bytes) return fail(c, FITSLAM_E_INVALID, "score_records: null argument");
  *device_ptr = c->records.p;
  *bytes = c->records_rows * sizeof(ScoreRecord);  // cycle: one row per frontier; score_arrival: nranks equal blocks
  return FITSLAM_OK;
}

}  // extern "C"
