// raycast.cu -- stage R: per-(frontier, yaw-window) raycast information gain.
//
// Replaces FrontierCostCalculator::setArrivalInformationForFrontier (DEP/src/CostCalculator.cpp:23-121),
// getTracedCells + bresenham2D (DEP/src/Helpers.cpp:7-96), RayTracedCells (DEP/include/.../Helpers.hpp:20-111),
// isRobotFootprintInLethal (DEP/src/Helpers.cpp:135-155, with the bounds check of RE/src/Helpers.cpp:126-149)
// and setMaxArrivalInformation (DEP/src/CostCalculator.cpp:123-191).
//
// Layout: one thread group (a whole number of warps) per frontier, one lane per ray.  The (2*half+1)^2 costmap
// window around the goal cell is staged into shared memory by the TMA engine, one 1-D bulk copy per window row
// (cp.async.bulk.shared::cluster.global, SASS UBLKCP, completion on an mbarrier; the tensor-map form UTMALDG raises
// "illegal instruction" on this pool's B200 boxes even through libcu++'s own wrapper), rays walk it
// with Bresenham out of shared memory,
// per-ray unknown-cell counts are prefix-summed with warp shuffles, the k-ray windows are differenced from the
// prefix and the first-max argmax is a warp-shuffle reduction.  All outputs are integers (bit-exact); the only
// floating point is the IEEE double arithmetic that locates the end cell of each ray (no FMA contraction:
// this file is compiled with -fmad=false).
#include "ctx.h"

#include <cmath>
#include <cstring>

namespace fitslam {

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t phase) {
  uint32_t done = 0;
  while (!done) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done)
        : "r"(smem_u32(bar)), "r"(phase)
        : "memory");
  }
}
// 1-D bulk copy global -> shared through the TMA engine; src/dst 16-byte aligned, bytes a multiple of 16
__device__ __forceinline__ void tma_bulk_load(void *dst, const void *src, uint32_t bytes, uint64_t *bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
               "l"(src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}

struct RayParams {
  MapView map;
  RayShape shape;
  const double *ray_dxdy;  // [2*n_rays]
  double poly[4];          // polygon_xy_min_max
  double hi_x, hi_y;       // getOriginX() + getSizeInMetersX(), likewise y
  double max_length_d;     // (double)max_length, the `double max_length` argument of getTracedCells
  double min_arrival_info_gt;
  double delta_theta, half_fov;
  uint32_t tile_w, tile_h;  // staged window: tile_w = round_up(2*half+1, 16), tile_h = 2*half+1
  uint32_t tpf;             // threads per frontier (multiple of 32)
  uint32_t fpb;             // frontiers per block
  uint32_t tile_stride;     // bytes between the windows of two frontiers (multiple of 128)
  uint32_t n;               // frontiers in the list
  ScoreSlice slice;         // this launch scores the frontiers of the slice (a rank's shard, or a batch scored early)
};

// Costmap2D::worldToMap (nav2_costmap_2d, ROS 2 Humble)
__device__ __forceinline__ bool world_to_map(const MapView &m, double wx, double wy, uint32_t &mx, uint32_t &my) {
  if (wx < m.ox || wy < m.oy) return false;
  mx = (uint32_t)((wx - m.ox) / m.res);
  my = (uint32_t)((wy - m.oy) / m.res);
  return mx < m.W && my < m.H;
}

// One ray: getTracedCells + bresenham2D + RayTracedCells(obstacle 240..254, trace 255..255).
// `cell(off)` returns the cost at a linear offset in whatever address space the caller staged; COUNT_ALL is the
// (260,260,0,255) gatherer of setMaxArrivalInformation, which pushes every visited cell.
template <bool COUNT_ALL, typename CellFn>
__device__ __forceinline__ int trace_ray(int x0, int y0, int x1, int y1, int stride, int off0, double max_length_d,
                                         CellFn cell) {
  const int dx = x1 - x0, dy = y1 - y0;
  const double dist = sqrt((double)(dx * dx + dy * dy));  // std::hypot of two small integers (exact radicand)
  const unsigned abs_dx = (unsigned)abs(dx), abs_dy = (unsigned)abs(dy);
  const int offset_dx = dx > 0 ? 1 : -1;
  const int offset_dy = (dy > 0 ? 1 : -1) * stride;
  const double scale = (dist == 0.0) ? 1.0 : fmin(1.0, max_length_d / dist);
  unsigned abs_da, abs_db;
  int offset_a, offset_b;
  if (abs_dx >= abs_dy) {
    abs_da = abs_dx; abs_db = abs_dy; offset_a = offset_dx; offset_b = offset_dy;
  } else {
    abs_da = abs_dy; abs_db = abs_dx; offset_a = offset_dy; offset_b = offset_dx;
  }
  int error_b = (int)(abs_da / 2);
  const unsigned max_len = (unsigned)(scale * (double)abs_da);
  const unsigned end = min(max_len, abs_da);
  if (COUNT_ALL) return (int)end + 1;
  int count = 0;
  int off = off0;
  for (unsigned i = 0; i <= end; ++i) {
    const int cost = cell(off);
    if (cost == 255) ++count;
    else if (cost >= 240) break;  // obstacle band 240..254: nothing is pushed after the first hit
    off += offset_a;
    error_b += (int)abs_db;
    if ((unsigned)error_b >= abs_da) {
      off += offset_b;
      error_b -= (int)abs_da;
    }
  }
  return count;
}

template <bool USE_TMA>
__global__ void __launch_bounds__(1024)
raycast_kernel(const RayParams P, const fitslam_frontier *__restrict__ fr,
               fitslam_arrival *__restrict__ out, uint16_t *__restrict__ info_ray_out, int32_t *__restrict__ win_out,
               ScoreRecord *__restrict__ rec_out) {
  extern __shared__ __align__(128) uint8_t smem[];
  const uint32_t g = threadIdx.x / P.tpf;  // frontier slot inside the block
  const uint32_t t = threadIdx.x - g * P.tpf;
  const uint32_t lane = threadIdx.x & 31u;
  const uint32_t wig = t >> 5;  // warp inside the group
  const uint32_t nwg = P.tpf >> 5;
  // shared layout: [fpb tiles][fpb mbarriers][fpb * (tpf prefix ints)][fpb * 32 warp scratch (int64)][fpb flags]
  uint8_t *tile = smem + (size_t)g * P.tile_stride;
  uint64_t *bars = reinterpret_cast<uint64_t *>(smem + (USE_TMA ? (size_t)P.fpb * P.tile_stride : 0));
  int32_t *prefix_all = reinterpret_cast<int32_t *>(bars + P.fpb);
  int32_t *prefix = prefix_all + (size_t)g * (P.tpf + 1);
  long long *wscr_all = reinterpret_cast<long long *>(prefix_all + (size_t)P.fpb * (P.tpf + 1) + ((P.fpb * (P.tpf + 1)) & 1));
  long long *wscr = wscr_all + (size_t)g * 32;
  int32_t *flags_all = reinterpret_cast<int32_t *>(wscr_all + (size_t)P.fpb * 32);
  int32_t *gflag = flags_all + g * 2;  // [0] any ray failed, [1] footprint in lethal

  const uint32_t slot = blockIdx.x * P.fpb + g;
  uint32_t f = 0u;
  const bool live = slice_frontier(P.slice, slot, f);

  double sx = 0.0, sy = 0.0;
  int fsize = 0;
  uint32_t x0 = 0, y0 = 0;
  bool goal_ok = false;
  if (live) {
    sx = fr[f].goal_x;
    sy = fr[f].goal_y;
    fsize = fr[f].size;
    goal_ok = world_to_map(P.map, sx, sy, x0, y0);
  }
  const int half = (int)P.shape.half;
  if (t == 0) {
    gflag[0] = 0;
    gflag[1] = 0;
    if (USE_TMA) {
      mbar_init(&bars[g], 1);
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
  }
  __syncthreads();
  // staged window: rows [wy0, wy0 + wrows) x columns [wx0, wx0 + tile_w) of the map, wx0 16-byte aligned and
  // clamped into the pitched row so every bulk copy stays inside the allocation
  int wx0 = 0, wy0 = 0, wrows = 0;
  if (USE_TMA && live && goal_ok) {
    wx0 = ((int)x0 - half) & ~15;
    wx0 = max(0, min(wx0, (int)P.map.pitch - (int)P.tile_w));
    wy0 = max(0, (int)y0 - half);
    wrows = min((int)P.map.H - 1, (int)y0 + half) - wy0 + 1;
    if (t == 0) mbar_expect_tx(&bars[g], (uint32_t)wrows * P.tile_w);
    for (int r = (int)t; r < wrows; r += (int)P.tpf)
      tma_bulk_load(tile + r * (int)P.tile_w, P.map.data + (size_t)(wy0 + r) * P.map.pitch + wx0, P.tile_w, &bars[g]);
  }

  // ---- ray end cell (DEP/src/CostCalculator.cpp:42-48), overlapped with the tile load
  bool ray_ok = true;
  uint32_t x1 = 0, y1 = 0;
  const bool has_ray = live && t < P.shape.n_rays;
  if (has_ray) {
    double wx = sx + P.ray_dxdy[2 * t];
    double wy = sy + P.ray_dxdy[2 * t + 1];
    wx = fmax(P.poly[0], fmax(P.map.ox, fmin(P.poly[2], fmin(P.hi_x, wx))));
    wy = fmax(P.poly[1], fmax(P.map.oy, fmin(P.poly[3], fmin(P.hi_y, wy))));
    ray_ok = world_to_map(P.map, wx, wy, x1, y1) && goal_ok;
    if (!ray_ok) atomicOr(&gflag[0], 1);
  }
  if (USE_TMA && live && goal_ok) mbar_wait(&bars[g], 0);

  int count = 0;
  if (has_ray && ray_ok) {
    if (USE_TMA) {
      const int stride = (int)P.tile_w;
      count = trace_ray<false>((int)x0, (int)y0, (int)x1, (int)y1, stride, ((int)y0 - wy0) * stride + ((int)x0 - wx0),
                               P.max_length_d, [&](int off) { return (int)tile[off]; });
    } else {
      const uint8_t *__restrict__ m = P.map.data;
      const int stride = (int)P.map.pitch;
      count = trace_ray<false>((int)x0, (int)y0, (int)x1, (int)y1, stride, (int)(y0 * P.map.pitch + x0),
                               P.max_length_d, [&](int off) { return (int)__ldg(m + off); });
    }
  }

  // ---- isRobotFootprintInLethal: disc of radius r cells around the goal cell, any cost == 254
  if (live && goal_ok) {
    const int r = (int)P.shape.radius_cells;
    const int side = 2 * r + 1;
    bool lethal = false;
    for (int i = (int)t; i < side * side; i += (int)P.tpf) {
      const int dy = i / side - r, dx = i - (i / side) * side - r;
      if (dx * dx + dy * dy > r * r) continue;
      const int x = (int)x0 + dx, y = (int)y0 + dy;
      if (x < 0 || y < 0 || x >= (int)P.map.W || y >= (int)P.map.H) continue;
      int cost;
      if (USE_TMA) {
        cost = tile[(y - wy0) * (int)P.tile_w + (x - wx0)];
      } else {
        cost = __ldg(P.map.data + (size_t)y * P.map.pitch + x);
      }
      lethal |= (cost == 254);
    }
    if (__any_sync(0xffffffffu, lethal) && lane == 0) atomicOr(&gflag[1], 1);
  }

  // ---- inclusive prefix of information_along_ray over the group's warps
  int incl = count;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const int v = __shfl_up_sync(0xffffffffu, incl, d);
    if ((int)lane >= d) incl += v;
  }
  if (lane == 31) wscr[wig] = incl;
  __syncthreads();
  int base = 0;
  for (uint32_t w = 0; w < wig; ++w) base += (int)wscr[w];
  incl += base;
  if (t == 0) prefix[0] = 0;
  prefix[t + 1] = incl;
  __syncthreads();

  // ---- box filter of k rays (no wrap-around) and first-max argmax (DEP/src/CostCalculator.cpp:87-107)
  long long key = -1;
  int wsum = 0;
  if (live && t < P.shape.n_windows) {
    wsum = prefix[t + P.shape.k] - prefix[t];
    key = ((long long)wsum << 32) | (long long)(0xffffffffu - t);
  }
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) {
    const long long o = __shfl_xor_sync(0xffffffffu, key, d);
    key = o > key ? o : key;
  }
  if (lane == 0) wscr[wig] = key;
  __syncthreads();
  if (!live) return;
  const bool failed = gflag[0] != 0;
  if (info_ray_out && t < P.shape.n_rays) info_ray_out[(size_t)f * P.shape.n_rays + t] = failed ? 0 : (uint16_t)count;
  if (win_out && t < P.shape.n_windows) win_out[(size_t)f * P.shape.n_windows + t] = failed ? 0 : wsum;
  if (t == 0) {
    long long best = wscr[0];
    for (uint32_t w = 1; w < nwg; ++w) best = wscr[w] > best ? wscr[w] : best;
    fitslam_arrival a;
    memset(&a, 0, sizeof(a));
    a.achievable = 1;  // Frontier(): is_achievable = true (DEP/src/Frontier.cpp:17)
    if (failed) {
      a.ray_failed = 1;  // setArrivalInformation(0.0); setGoalOrientation(0.0); return  (:50-55)
    } else {
      const int max_value = (int)(best >> 32);
      const int max_index = (int)(0xffffffffu - (uint32_t)(best & 0xffffffffll));
      a.information = max_value;
      a.max_index = max_index;
      a.footprint_in_lethal = gflag[1] ? 1 : 0;
      if (gflag[1] && fsize < 10) a.achievable = 0;                      // :77-82
      if ((double)max_value < P.min_arrival_info_gt) a.achievable = 0;  // :112-118
      a.theta_s_star = ((double)max_index * P.delta_theta) + P.half_fov;  // :119
    }
    out[f] = a;
    if (rec_out) {
      ScoreRecord r;
      r.information = a.information;
      r.max_index = a.max_index;
      r.fisher = 0.0f;
      r.flags = (a.achievable ? 1u : 0u) | (a.footprint_in_lethal ? 2u : 0u) | (a.ray_failed ? 4u : 0u);
      rec_out[f] = r;
    }
  }
}

// setMaxArrivalInformation (DEP/src/CostCalculator.cpp:123-191): the fan from (0,0) with the count-everything
// gatherer.  One block; out[0] = 1 if every ray traced, out[1] = max window sum.
__global__ void limits_kernel(const RayParams P, int32_t *out) {
  __shared__ int32_t info[1025];
  __shared__ int32_t fail_flag;
  if (threadIdx.x == 0) fail_flag = 0;
  __syncthreads();
  const uint32_t t = threadIdx.x;
  if (t < P.shape.n_rays) {
    uint32_t x0, y0, x1, y1;
    const double wx = 0.0 + P.ray_dxdy[2 * t];
    const double wy = 0.0 + P.ray_dxdy[2 * t + 1];
    const bool ok = world_to_map(P.map, wx, wy, x1, y1) && world_to_map(P.map, 0.0, 0.0, x0, y0);
    if (!ok) {
      atomicOr(&fail_flag, 1);
    } else {
      info[t] = trace_ray<true>((int)x0, (int)y0, (int)x1, (int)y1, 1, 0, P.max_length_d, [](int) { return 0; });
    }
  }
  __syncthreads();
  if (t == 0) {
    if (fail_flag) {
      out[0] = 0;
      out[1] = 0;
    } else {
      int best = 0;
      bool first = true;
      for (uint32_t i = 0; i < P.shape.n_windows; ++i) {
        int s = 0;
        for (uint32_t j = 0; j < P.shape.k; ++j) s += info[i + j];
        if (first || s > best) best = s;
        first = false;
      }
      out[0] = 1;
      out[1] = best;
    }
  }
}

static RayParams make_params(fitslam_ctx *c, const double poly[4], uint32_t n) {
  RayParams P;
  memset(&P, 0, sizeof(P));
  P.map = map_view(c);
  P.shape = c->shape;
  P.ray_dxdy = c->ray_dxdy.p;
  const double none[4] = {-1e300, -1e300, 1e300, 1e300};
  const double *pl = poly ? poly : none;
  for (int i = 0; i < 4; ++i) P.poly[i] = pl[i];
  // Costmap2D::getSizeInMetersX() = (size_x - 1 + 0.5) * resolution
  P.hi_x = c->ox + ((c->W - 1 + 0.5) * c->res);
  P.hi_y = c->oy + ((c->H - 1 + 0.5) * c->res);
  P.max_length_d = (double)c->shape.max_length;
  P.min_arrival_info_gt = c->min_arrival_info_gt;
  P.delta_theta = c->prm.delta_theta;
  P.half_fov = c->prm.camera_fov / 2;
  P.tile_h = 2 * c->shape.half + 1;
  P.tile_w = (P.tile_h + 15u + 15u) & ~15u;  // + up to 15 columns of slack from aligning the row start down to 16 B
  P.tpf = (std::max(c->shape.n_rays, 32u) + 31u) & ~31u;
  P.tile_stride = (P.tile_w * P.tile_h + 127u) & ~127u;
  P.n = n;
  shard_range(c, n, &P.slice.first, &P.slice.count);
  return P;
}

static size_t smem_bytes(const RayParams &P, bool tma) {
  size_t s = tma ? (size_t)P.fpb * P.tile_stride : 0;
  s += (size_t)P.fpb * 8;                                                     // mbarriers
  s += ((size_t)P.fpb * (P.tpf + 1) + ((P.fpb * (P.tpf + 1)) & 1)) * 4;       // prefix
  s += (size_t)P.fpb * 32 * 8;                                                // warp scratch
  s += (size_t)P.fpb * 2 * 4;                                                 // flags
  return s;
}

fitslam_status raycast_limits(fitslam_ctx *c) {
  if (c->limits_set) return FITSLAM_OK;  // arrival_info_limits_set_ (DEP/src/CostCalculator.cpp:125-129)
  RayParams P = make_params(c, nullptr, 0);
  FS_CUDA(c, c->scratch.reserve(64));
  int32_t *d_out = reinterpret_cast<int32_t *>(c->scratch.p);
  limits_kernel<<<1, (c->shape.n_rays + 31u) & ~31u, 0, c->stream>>>(P, d_out);
  ++c->launches;
  FS_CUDA(c, cudaGetLastError());
  int32_t h[2] = {0, 0};
  FS_CUDA(c, cudaMemcpyAsync(h, d_out, sizeof(h), cudaMemcpyDeviceToHost, c->stream));
  FS_CUDA(c, cudaStreamSynchronize(c->stream));
  c->limits_max_value = h[1];
  if (h[0]) {
    c->limits_set = true;
    c->max_arrival_info_gt = h[1] * c->prm.max_arrival_factor;                 // :186
    c->min_arrival_info_gt = c->prm.min_arrival_factor * c->max_arrival_info_gt;  // :188
  }
  return FITSLAM_OK;
}

static fitslam_status raycast_launch(fitslam_ctx *c, cudaStream_t stream, RayParams &P, uint32_t slots, const fitslam_frontier *d_fr,
                                     fitslam_arrival *d_out, uint16_t *d_info_ray, int32_t *d_win, ScoreRecord *d_rec) {
  const bool tma = c->use_tma && c->pitch >= P.tile_w;
  // frontiers per block: fill ~256 threads, bounded by shared memory
  uint32_t fpb = std::max(1u, 256u / P.tpf);
  while (fpb > 1 && (size_t)fpb * P.tile_stride > 96 * 1024) --fpb;
  P.fpb = fpb;
  const size_t smem = smem_bytes(P, tma);
  if (slots == 0) return FITSLAM_OK;
  const uint32_t grid = (slots + fpb - 1) / fpb;
  const uint32_t block = fpb * P.tpf;
  if (c->ray_smem_attr < smem) {  // (set before the early launches run concurrently: the value only grows)
    FS_CUDA(c, cudaFuncSetAttribute(raycast_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    FS_CUDA(c, cudaFuncSetAttribute(raycast_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    c->ray_smem_attr = smem;
  }
  if (tma) raycast_kernel<true><<<grid, block, smem, stream>>>(P, d_fr, d_out, d_info_ray, d_win, d_rec);
  else raycast_kernel<false><<<grid, block, smem, stream>>>(P, d_fr, d_out, d_info_ray, d_win, d_rec);
  ++c->launches;
  FS_CUDA(c, cudaGetLastError());
  return FITSLAM_OK;
}

fitslam_status raycast_score(fitslam_ctx *c, const fitslam_frontier *d_fr, uint32_t n, const double poly[4],
                             fitslam_arrival *d_out, uint16_t *d_info_ray, int32_t *d_win, ScoreRecord *d_rec) {
  if (n == 0) return FITSLAM_OK;
  RayParams P = make_params(c, poly, n);
  return raycast_launch(c, c->stream, P, P.slice.count, d_fr, d_out, d_info_ray, d_win, d_rec);
}

// the same for a slice of the resident list (cycle.cu: batches scored while the search runs, then the rest)
fitslam_status raycast_score_slice(fitslam_ctx *c, cudaStream_t stream, const fitslam_frontier *d_fr, const double poly[4],
                                   fitslam_arrival *d_out, ScoreRecord *d_rec, const ScoreSlice &slice) {
  RayParams P = make_params(c, poly, 0);
  P.slice = slice;
  return raycast_launch(c, stream, P, slice.range ? slice.cap : slice.count, d_fr, d_out, nullptr, nullptr, d_rec);
}

}  // namespace fitslam

// ================================================================================================ roadmap edges
namespace fitslam {

// FrontierRoadMap::isConnectable (RE/src/planners/FrontierRoadmap.cpp:671-694): one thread per node pair walks the
// Bresenham line of getTracedCells (RE/src/Helpers.cpp:30-91, bresenham2D :7-28) over the costmap.  The cells of a
// Bresenham line are distinct, so RayTracedCells' duplicate check (Helpers.hpp:41-46) never fires and the walk needs
// no list: obstacle = any cost in 253..254 (hasHitObstacle), unknown = count of 255 (getNumUnknown).
__global__ void connectable_kernel(const MapView M, const double *__restrict__ seg, uint32_t n, double max_length_d,
                                   double unknown_limit, uint8_t *__restrict__ ok_out, int32_t *__restrict__ unk_out) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const double sx = seg[4 * i], sy = seg[4 * i + 1], wx = seg[4 * i + 2], wy = seg[4 * i + 3];
  uint32_t x0, y0, x1, y1;
  if (!world_to_map(M, wx, wy, x1, y1) || !world_to_map(M, sx, sy, x0, y0)) {  // "Not world to map": not connectable
    ok_out[i] = 0;
    if (unk_out) unk_out[i] = -1;
    return;
  }
  const int dx = (int)(x1 - x0), dy = (int)(y1 - y0);
  // std::hypot(int, int): the integer sum of squares is exact in double, and its IEEE square root is the correctly
  // rounded hypotenuse
  const double dist = sqrt((double)((long long)dx * dx + (long long)dy * dy));
  const uint32_t abs_dx = (uint32_t)abs(dx), abs_dy = (uint32_t)abs(dy);
  const long long step_x = dx > 0 ? 1 : -1, step_y = (dy > 0 ? 1 : -1) * (long long)M.pitch;  // sign(0) = -1 (Helpers.hpp:92-95)
  const double scale = (dist == 0.0) ? 1.0 : fmin(1.0, max_length_d / dist);
  const bool x_major = abs_dx >= abs_dy;
  const uint32_t abs_da = x_major ? abs_dx : abs_dy, abs_db = x_major ? abs_dy : abs_dx;
  const long long off_a = x_major ? step_x : step_y, off_b = x_major ? step_y : step_x;
  const uint32_t end = min((uint32_t)(scale * abs_da), abs_da);
  int error_b = (int)(abs_da / 2);
  long long off = (long long)y0 * M.pitch + x0;
  bool hit = false;
  int unknown = 0;
  for (uint32_t k = 0; k <= end; ++k) {  // `end` steps, then the final at(offset)
    const uint32_t cost = M.data[off];
    hit |= cost >= 253u && cost <= 254u;
    unknown += cost == 255u ? 1 : 0;
    off += off_a;
    error_b += (int)abs_db;
    if ((uint32_t)error_b >= abs_da) {
      off += off_b;
      error_b -= (int)abs_da;
    }
  }
  ok_out[i] = (!hit && !((double)unknown > unknown_limit)) ? 1 : 0;
  if (unk_out) unk_out[i] = unknown;
}

}  // namespace fitslam

extern "C" fitslam_status fitslam_roadmap_connectable(fitslam_ctx *c, const double *segments, uint32_t n,
                                                      double radius_to_decide_edges, uint8_t *connectable,
                                                      int32_t *num_unknown) {
  using namespace fitslam;
  if (!c || (!segments && n) || (!connectable && n)) return fail(c, FITSLAM_E_INVALID, "roadmap_connectable: null argument");
  if (!c->have_map) return fail(c, FITSLAM_E_NO_COSTMAP, "no costmap set");
  if (!(radius_to_decide_edges >= 0.0)) return fail(c, FITSLAM_E_INVALID, "roadmap_connectable: bad radius");
  if (n == 0) return FITSLAM_OK;
  cudaSetDevice(c->device);
  cudaPointerAttributes at;
  const bool seg_dev = cudaPointerGetAttributes(&at, segments) == cudaSuccess && at.type == cudaMemoryTypeDevice;
  cudaGetLastError();
  const double *d_seg = segments;
  if (!seg_dev) {
    FS_CUDA(c, c->edge_seg.reserve(4 * (size_t)n));
    FS_CUDA(c, cudaMemcpyAsync(c->edge_seg.p, segments, 4 * (size_t)n * sizeof(double), cudaMemcpyHostToDevice, c->stream));
    d_seg = c->edge_seg.p;
  }
  // [n int32 unknown counts][n bytes flags]
  FS_CUDA(c, c->edge_out.reserve((size_t)n + ((size_t)n + 3) / 4));
  int32_t *d_unk = c->edge_out.p;
  uint8_t *d_ok = reinterpret_cast<uint8_t *>(c->edge_out.p + n);
  // max_connection_length_ = RADIUS_TO_DECIDE_EDGES * 1.5 (FrontierRoadmap.cpp:32); unsigned int max_length = that / resolution (:675)
  const unsigned int max_length = (unsigned int)((radius_to_decide_edges * 1.5) / c->res);
  const double unknown_limit = radius_to_decide_edges / c->res * 0.3;  // :685
  connectable_kernel<<<(n + 127) / 128, 128, 0, c->stream>>>(map_view(c), d_seg, n, (double)max_length, unknown_limit, d_ok, d_unk);
  ++c->launches;
  FS_CUDA(c, cudaGetLastError());
  FS_CUDA(c, cudaMemcpyAsync(connectable, d_ok, n, cudaMemcpyDefault, c->stream));
  if (num_unknown) FS_CUDA(c, cudaMemcpyAsync(num_unknown, d_unk, (size_t)n * sizeof(int32_t), cudaMemcpyDefault, c->stream));
  FS_CUDA(c, cudaStreamSynchronize(c->stream));
  return FITSLAM_OK;
}
