// ctx.h -- internal state of one fitslam context (one CUDA stream, device-resident map + points).
// Not part of the ABI; include/fitslam_cuda.h is.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <string>
#include <vector>

#include "../../include/fitslam_cuda.h"

namespace fitslam {

// Device view of a nav2_costmap_2d::Costmap2D snapshot (T1).  Rows are stored with a pitch that is a
// multiple of 16 bytes so every window row can be fetched with 16-byte aligned TMA bulk copies; map indices at the ABI are my*W+mx.
struct MapView {
  const uint8_t *data;
  uint32_t W, H, pitch;
  double res, ox, oy;
};

// Ray fan geometry (function of the parameters only; DEP/src/CostCalculator.cpp:28,36,42,87).
struct RayShape {
  uint32_t n_rays;     // iterations of `for (theta = 0; theta <= 2*pi; theta += delta_theta)`
  uint32_t k;          // static_cast<int>(camera_fov / delta_theta)
  uint32_t n_windows;  // n_rays - k + 1
  uint32_t max_length; // (unsigned)(max_camera_depth / resolution)
  uint32_t radius_cells; // ceil(robot_radius / resolution)
  uint32_t half;       // half size of the staged window: max(max_length, radius_cells)
};

template <typename T>
struct DevBuf {
  T *p = nullptr;
  size_t cap = 0;  // elements
  cudaError_t reserve(size_t n) {
    if (n <= cap) return cudaSuccess;
    if (p) cudaFree(p);
    p = nullptr;
    cap = 0;
    size_t want = n + n / 4 + 64;
    cudaError_t e = cudaMalloc((void **)&p, want * sizeof(T));
    if (e == cudaSuccess) cap = want;
    return e;
  }
  void release() {
    if (p) cudaFree(p);
    p = nullptr;
    cap = 0;
  }
};

struct FrontierState;
struct FisherState;
struct SelectState;
struct CycleState;
struct CommState;

struct ScoreRecord {  // 16 bytes, the all-gather unit
  int32_t information;
  int32_t max_index;
  float fisher;
  uint32_t flags;  // bit0 achievable, bit1 footprint_in_lethal, bit2 ray_failed, bit3 fisher valid, bit4 fisher safe
};

}  // namespace fitslam

struct fitslam_ctx {
  fitslam_params prm;
  int device = 0;
  cudaStream_t stream = nullptr;
  std::string err;
  uint64_t launches = 0;
  float last_bfs_ms = 0;  // device time of the outer wavefront kernel in the last search
  int rank = 0, nranks = 1;

  // ---- costmap
  fitslam::DevBuf<uint8_t> map;
  uint32_t W = 0, H = 0, pitch = 0;
  double res = 0, ox = 0, oy = 0;
  bool have_map = false;
  int use_tma = 1;        // FITSLAM_RAYCAST_TMA=0 switches to the direct-gather variant (ablation)
  size_t ray_smem_attr = 0;  // dynamic shared memory the raycast kernels are opted in for

  // ---- ray fan tables (host-built from parameters, uploaded)
  fitslam::RayShape shape{};
  fitslam::DevBuf<double> ray_dxdy;  // 2*n_rays doubles: depth*cos(theta_i), depth*sin(theta_i)
  double shape_res = -1.0;
  bool limits_set = false;
  double max_arrival_info_gt = 0.0, min_arrival_info_gt = 0.0;
  int32_t limits_max_value = 0;

  // ---- frontier list resident on the device (last search / last upload)
  fitslam::DevBuf<fitslam_frontier> frontiers;
  uint32_t n_frontiers = 0;
  fitslam::DevBuf<fitslam_arrival> arrival;
  fitslam::DevBuf<uint16_t> info_ray;
  fitslam::DevBuf<int32_t> win_sums;
  fitslam::DevBuf<fitslam::ScoreRecord> records;
  uint64_t records_rows = 0;  // rows of `records` the last scoring call laid out (fitslam_score_records)

  fitslam::DevBuf<int8_t> occ_stage;  // fitslam_set_occupancy_grid: the int8 grid (or its dirty rectangle) before translation
  fitslam::DevBuf<double> stamp_tab;  // fitslam_mark_lethal / fitslam_sensor_mark_rays: host-built ray tables
  // ---- roadmap edge batches (fitslam_roadmap_connectable)
  fitslam::DevBuf<double> edge_seg;
  fitslam::DevBuf<int32_t> edge_out;

  // ---- scratch
  fitslam::DevBuf<uint8_t> scratch;
  void *pinned = nullptr;
  size_t pinned_cap = 0;

  // opaque per-stage state (defined in the stage's .cu file)
  fitslam::FrontierState *fs = nullptr;
  fitslam::FisherState *fi = nullptr;
  fitslam::SelectState *sel = nullptr;
  fitslam::CycleState *cy = nullptr;
  fitslam::CommState *comm = nullptr;  // NCCL communicator (comm.cu); null on a single GPU
};

namespace fitslam {

inline fitslam_status fail(fitslam_ctx *c, fitslam_status s, const std::string &msg) {
  if (c) c->err = msg;
  return s;
}
#define FS_CUDA(ctx, expr)                                                                         \
  do {                                                                                             \
    cudaError_t _e = (expr);                                                                       \
    if (_e != cudaSuccess) {                                                                       \
      return ::fitslam::fail((ctx), FITSLAM_E_CUDA,                                                \
                             std::string(#expr) + ": " + cudaGetErrorString(_e));                  \
    }                                                                                              \
  } while (0)

inline MapView map_view(const fitslam_ctx *c) {
  MapView m;
  m.data = c->map.p;
  m.W = c->W;
  m.H = c->H;
  m.pitch = c->pitch;
  m.res = c->res;
  m.ox = c->ox;
  m.oy = c->oy;
  return m;
}

fitslam_status ensure_pinned(fitslam_ctx *c, size_t bytes);
// candidate sharding.  In the cycle, rank r owns the frontiers f with f % nranks == r: ownership is then known
// the moment a frontier exists, so a rank can score its share while the search is still running; every rank writes its
// 16-byte records into a zeroed buffer of one record per frontier, and an all-reduce (integer sum: one owner, zeros
// elsewhere) merges them.  The stand-alone scoring calls (fitslam_score_arrival / _fisher on a caller's list) use
// contiguous blocks instead: rank r owns [r * B, min(n, (r + 1) * B)), B = ceil(n / nranks), and their records
// concatenate into an in-place all-gather buffer of nranks * B records
inline uint32_t shard_block(const fitslam_ctx *c, uint32_t n) { return (n + (uint32_t)c->nranks - 1) / (uint32_t)c->nranks; }
inline void shard_range(const fitslam_ctx *c, uint32_t n, uint32_t *first, uint32_t *count) {
  const uint32_t B = shard_block(c, n);
  const uint64_t lo = (uint64_t)c->rank * B;
  *first = (uint32_t)(lo < n ? lo : n);
  *count = lo < n ? (uint32_t)((lo + B < n ? lo + B : n) - lo) : 0u;
}
fitslam_status update_ray_shape(fitslam_ctx *c);

void destroy_frontier_state(fitslam_ctx *c);
void destroy_fisher_state(fitslam_ctx *c);
void destroy_select_state(fitslam_ctx *c);
void destroy_cycle_state(fitslam_ctx *c);
void destroy_comm_state(fitslam_ctx *c);
bool comm_active(const fitslam_ctx *c);
fitslam_status comm_broadcast(fitslam_ctx *c, void *dev, size_t bytes);  // rank 0 -> all, on the context stream

// ---- stage N state (select.cu), shared with cycle.cu
struct SelectMeta {
  double min_len, max_len, min_info;
  uint32_t n_eligible;     // achievable && !blacklisted
  uint32_t n_local_range;  // ... && path_length_m <= LOCAL_FRONTIER_SEARCH_RADIUS
  uint32_t err_key;        // 2 * index + (0: arrival utility > 1, 1: cost out of bounds); 0xffffffff = none
  uint32_t pad;
  fitslam_selection sel;
};
struct SelectState {
  DevBuf<double> len, head, wc, ua, ud;
  DevBuf<uint8_t> bl, ach;
  DevBuf<uint32_t> sorted, rank;
  DevBuf<SelectMeta> meta;
};
SelectState *sstate(fitslam_ctx *c);

// A slice of the resident frontier list for one scoring launch: frontiers first + [0, count) whose tag matches.  With
// `range` set, first / count come from device memory at kernel run time ({first, end}, count clamped to `cap`): that is
// how frontiers are scored while the search that produces them is still running, without a host round trip.
struct ScoreSlice {
  const uint32_t *range = nullptr;
  uint32_t first = 0, count = 0, cap = 0xFFFFFFFFu;
  uint32_t limit = 0xFFFFFFFFu;   // with `range`: frontiers at or above this index belong to a later launch
  const uint8_t *tags = nullptr;  // per frontier: which launch scores it (nullptr: no filter)
  uint32_t tag = 0;
  uint32_t mod = 1, rem = 0;      // only frontiers f with f % mod == rem (the cycle's sharding: rank `rem` of `mod`)
  // Compact work list of a Fisher launch (optional): make_poses_kernel appends every frontier of the slice that passes the
  // filters above, so that the Fisher grid holds work only (CTAs of slots outside the range, of another tag or of another
  // rank used to start and exit: up to 12 of them per scored pose).  *work_count = frontiers appended; the launch covers the
  // first work_cap entries of work_list; the ones beyond it are retagged TAG_FINAL (scored by the launches after the search).
  uint32_t *work_count = nullptr, *work_list = nullptr;
  uint32_t work_cap = 0;
  uint8_t *retag = nullptr;
};
#ifdef __CUDACC__
__device__ __forceinline__ uint32_t slice_slots(const ScoreSlice &s) {  // slots of the slice (owned or not, any tag)
  if (!s.range) return s.count;
  const uint32_t first = s.range[0], end = min(s.range[1], s.limit);
  return end > first ? min(end - first, s.cap) : 0u;
}
__device__ __forceinline__ bool slice_frontier(const ScoreSlice &s, uint32_t slot, uint32_t &f) {
  uint32_t first = s.first, count = s.count;
  if (s.range) {
    first = s.range[0];
    const uint32_t end = min(s.range[1], s.limit);
    count = end > first ? min(end - first, s.cap) : 0u;
  }
  if (slot >= count) return false;
  f = first + slot;
  if (s.mod > 1u && f % s.mod != s.rem) return false;
  return !s.tags || s.tags[f] == s.tag;
}
#endif
#ifdef __CUDACC__
// slot k of a launch that reads its frontiers from the work list
__device__ __forceinline__ bool slice_work_frontier(const ScoreSlice &s, uint32_t slot, uint32_t &f) {
  if (slot >= min(*s.work_count, s.work_cap)) return false;
  f = s.work_list[slot];
  return true;
}
#endif
static constexpr uint32_t TAG_FINAL = 255;     // frontier produced after the search ended: scored by the host-sized launches
static constexpr uint32_t TAG_HOST = 254;      // frontier whose median sort the host redoes after the search (then TAG_FINAL)
static constexpr uint32_t EARLY_SLICE_CAP = 512;  // frontiers one early launch covers; what exceeds it is left to the final pass
// called by the frontier search once per finished batch of frontiers, on a side stream (cycle.cu scores them there)
struct EarlyHook {
  fitslam_status (*fn)(fitslam_ctx *c, cudaStream_t stream, const ScoreSlice &slice, void *user);
  void *user;
  uint32_t limit;  // the scorer's buffers hold this many frontiers: the ones at or above it are left to the launches after the search
};

// stage entry points (each .cu file)
fitslam_status frontier_search_device(fitslam_ctx *c, double rx, double ry, const EarlyHook *hook = nullptr);
uint32_t frontier_capacity(fitslam_ctx *c);          // frontiers the resident list can hold (0 on failure)
const uint8_t *frontier_tags_device(fitslam_ctx *c);  // per frontier of the last search: TAG_FINAL or the early launch that scored it
const uint32_t *frontier_cells_device(fitslam_ctx *c, uint32_t *n_cells, uint32_t *n_comp, uint32_t *levels, uint32_t *ambiguous);
void frontier_wavefront_device(fitslam_ctx *c, uint32_t *widest_level, uint32_t *ctas);  // of the last search's outer wavefront
fitslam_status fisher_score_device(fitslam_ctx *c, const fitslam_pose *d_poses, uint32_t n, uint32_t first, uint32_t count,
                                   float *d_info, uint8_t *d_safe, uint32_t *d_nvis, ScoreRecord *d_rec,
                                   const fitslam_arrival *d_arrival, uint32_t poses_per_frontier);
bool fisher_ready(const fitslam_ctx *c);
fitslam_status plan_euclidean_device(fitslam_ctx *c, uint32_t n, const fitslam_frontier *d_fr, const fitslam_arrival *d_arr,
                                     const double *d_goal_yaw, double rx, double ry, double ryaw);
fitslam_status select_device(fitslam_ctx *c, uint32_t n, const fitslam_arrival *d_arr, const uint8_t *d_ach,
                             const uint8_t *d_bl, const double *d_len, const double *d_head);
fitslam_status raycast_limits(fitslam_ctx *c);
fitslam_status raycast_score(fitslam_ctx *c, const fitslam_frontier *d_frontiers, uint32_t n,
                             const double poly[4], fitslam_arrival *d_out, uint16_t *d_info_ray,
                             int32_t *d_win_sums, ScoreRecord *d_records);
fitslam_status raycast_score_slice(fitslam_ctx *c, cudaStream_t stream, const fitslam_frontier *d_frontiers, const double poly[4],
                                   fitslam_arrival *d_out, ScoreRecord *d_records, const ScoreSlice &slice);
fitslam_status fisher_score_slice(fitslam_ctx *c, cudaStream_t stream, const fitslam_pose *d_poses, uint32_t poses_per_frontier,
                                  const ScoreSlice &slice, float *d_info, uint8_t *d_safe, uint32_t *d_nvis, ScoreRecord *d_rec,
                                  const fitslam_arrival *d_arrival);

}  // namespace fitslam
