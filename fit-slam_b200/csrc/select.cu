// select.cu -- stage N: cost normalisation (U1 utility block) and goal selection.
//
// Replaces FrontierCostsManager::assignCosts, U1 block (DEP/src/FrontierCostsManager.cpp:121-205), the running
// min/max of recomputeNormalizationFactors (DEP/src/CostCalculator.cpp:512-520), setPlanForFrontierEuclidean
// (DEP/src/CostCalculator.cpp:445-485) and FullPathOptimizer::getFilteredFrontiersN
// (DEP/src/FullPathOptimizer.cpp:157-227).
//
// The U1 block is plain IEEE double arithmetic (+ - * /), evaluated here in the reference's order with FMA
// contraction off, so weighted_cost is bit-exact.  The std::sort by path_length_m becomes a rank-by-counting pass
// (every eligible frontier counts the eligible frontiers that sort before it; ties broken by index) and the
// sequential local/global walk collapses to arithmetic on two counts (eligible, eligible within the local radius).
#include "ctx.h"

#include <cfloat>
#include <cmath>
#include <cstring>

namespace fitslam {

SelectState *sstate(fitslam_ctx *c) {
  if (!c->sel) c->sel = new SelectState();
  return c->sel;
}

void destroy_select_state(fitslam_ctx *c) {
  if (!c->sel) return;
  SelectState *s = c->sel;
  s->len.release(); s->head.release(); s->wc.release(); s->ua.release(); s->ud.release();
  s->bl.release(); s->ach.release(); s->sorted.release(); s->rank.release(); s->meta.release();
  delete s;
  c->sel = nullptr;
}

struct SelectParams {
  double alpha, beta, vx, wz, max_arrival_info_gt, local_radius;
  uint32_t n, n_local_area;
};

__device__ __forceinline__ double warp_min(double v) {
  for (int o = 16; o > 0; o >>= 1) v = fmin(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ double warp_max(double v) {
  for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

__global__ void init_achievable_kernel(const fitslam_arrival *__restrict__ arr, uint8_t *ach, uint32_t n) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) ach[i] = arr[i].achievable;
}
// the caller's path lengths decide too: a frontier the planner could not reach carries DBL_MAX (CostCalculator.cpp:449-466)
__global__ void init_achievable_len_kernel(const fitslam_arrival *__restrict__ arr, const double *__restrict__ len, uint8_t *ach, uint32_t n) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) ach[i] = arr[i].achievable && len[i] != DBL_MAX;
}

// setPlanForFrontierEuclidean for every frontier (DEP/src/CostCalculator.cpp:445-485); robot_yaw is an input, and so is
// atan2(goal - robot) per frontier: the host evaluates it with its own libm (glibc's atan2 is not correctly rounded, so no
// device implementation can promise the reference's last bit; cycle.cu overlaps the F calls with the scoring kernels).
__global__ void plan_euclidean_kernel(const fitslam_frontier *__restrict__ fr, const double *__restrict__ goal_yaw_in, uint8_t *ach,
                                      double *len, double *head, uint32_t n, double rx, double ry, double robot_yaw_in) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  if (!ach[i]) {
    len[i] = DBL_MAX;
    head[i] = DBL_MAX;
    return;
  }
  const double dx = rx - fr[i].goal_x, dy = ry - fr[i].goal_y;
  const double l = sqrt(dx * dx + dy * dy);
  if (l < 0.5) {
    ach[i] = 0;
    len[i] = DBL_MAX;
    head[i] = DBL_MAX;
    return;
  }
  double robot_yaw = robot_yaw_in;
  if (robot_yaw < 0) robot_yaw = robot_yaw + (M_PI * 2);
  double goal_yaw = goal_yaw_in[i];
  if (goal_yaw < 0) goal_yaw = goal_yaw + (M_PI * 2);
  double h = fabs(robot_yaw - goal_yaw);
  if (h > M_PI) h = (2 * M_PI) - h;
  len[i] = l;
  head[i] = h;
}

// One block.  Pass 1: min/max over the achievable, non-blacklisted frontiers + the two counts.  Pass 2: U1 per frontier.
__global__ void __launch_bounds__(1024)
normalise_kernel(const SelectParams P, const fitslam_arrival *__restrict__ arr, const uint8_t *__restrict__ ach,
                 const uint8_t *__restrict__ bl, const double *__restrict__ len, const double *__restrict__ head,
                 double *__restrict__ wc, double *__restrict__ ua, double *__restrict__ ud, SelectMeta *meta) {
  __shared__ double s_min[32], s_max[32], s_inf[32];
  __shared__ uint32_t s_e, s_l, s_err;
  if (threadIdx.x == 0) { s_e = 0; s_l = 0; s_err = 0xffffffffu; }
  __syncthreads();
  double mn = DBL_MAX, mx = -1.0, mi = DBL_MAX;
  uint32_t e = 0, l = 0;
  for (uint32_t i = threadIdx.x; i < P.n; i += blockDim.x) {
    if ((bl && bl[i]) || !ach[i]) continue;
    mn = fmin(mn, len[i]);
    mx = fmax(mx, len[i]);
    mi = fmin(mi, (double)arr[i].information);
    ++e;
    if (len[i] <= P.local_radius) ++l;
  }
  mn = warp_min(mn); mx = warp_max(mx); mi = warp_min(mi);
  const uint32_t lane = threadIdx.x & 31u, wid = threadIdx.x >> 5;
  if (lane == 0) { s_min[wid] = mn; s_max[wid] = mx; s_inf[wid] = mi; }
  if (e) atomicAdd(&s_e, e);
  if (l) atomicAdd(&s_l, l);
  __syncthreads();
  mn = s_min[0]; mx = s_max[0]; mi = s_inf[0];
  for (uint32_t w = 1; w < blockDim.x / 32; ++w) { mn = fmin(mn, s_min[w]); mx = fmax(mx, s_max[w]); mi = fmin(mi, s_inf[w]); }
  const double denom = mx / P.vx + M_PI / P.wz;
  const bool plan_flat = (denom - (mn / P.vx + 0.0 / P.wz)) == 0.0;
  const bool arrival_flat = (P.max_arrival_info_gt - mi) == 0.0;
  for (uint32_t i = threadIdx.x; i < P.n; i += blockDim.x) {
    if ((bl && bl[i]) || !ach[i]) {
      wc[i] = DBL_MAX;  // setWeightedCost(std::numeric_limits<double>::max()), costs -69.8 / -1.8 (:131-136)
      ua[i] = -69.8;
      ud[i] = -1.8;
      continue;
    }
    const double a = arrival_flat ? 0.0 : (double)arr[i].information / P.max_arrival_info_gt;
    if (a > 1.0) {
      atomicMin(&s_err, 2u * i);
      continue;
    }
    double p = plan_flat ? 1.0 : (len[i] / P.vx + head[i] / P.wz) / denom;
    p = 1.0 - p;
    if (a > 1.0 || a < 0.0 || p > 1.0 || p < 0.0) {
      atomicMin(&s_err, 2u * i + 1u);
      continue;
    }
    double u = (P.alpha * a) + ((1.0 - P.alpha) * p);
    if (u == 0.0) u = 1e-16;
    wc[i] = 1 / (P.beta * u);
    ua[i] = a;
    ud[i] = p;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    meta->min_len = mn; meta->max_len = mx; meta->min_info = mi;
    meta->n_eligible = s_e; meta->n_local_range = s_l; meta->err_key = s_err;
  }
}

// std::sort(all_frontiers, path_length_m ascending) as rank-by-counting over the eligible frontiers.
__global__ void __launch_bounds__(256)
rank_kernel(const uint8_t *__restrict__ ach, const uint8_t *__restrict__ bl, const double *__restrict__ len, uint32_t n,
            uint32_t *__restrict__ sorted) {
  __shared__ double s_len[256];
  __shared__ uint8_t s_ok[256];
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  const bool mine = i < n && ach[i] && !(bl && bl[i]);
  const double li = mine ? len[i] : 0.0;
  uint32_t r = 0;
  for (uint32_t base = 0; base < n; base += 256) {
    const uint32_t j = base + threadIdx.x;
    s_ok[threadIdx.x] = j < n && ach[j] && !(bl && bl[j]);
    s_len[threadIdx.x] = j < n ? len[j] : 0.0;
    __syncthreads();
    if (mine) {
      const uint32_t m = min(256u, n - base);
      for (uint32_t t = 0; t < m; ++t) {
        const double lj = s_len[t];
        r += (s_ok[t] && (lj < li || (lj == li && base + t < i))) ? 1u : 0u;
      }
    }
    __syncthreads();
  }
  if (mine) sorted[r] = i;
}

// The same ranks for long lists: the comparisons of one frontier are split over gridDim.y CTAs (each counts the smaller
// elements of its part of the list, partial ranks added up in `rank`), then a scatter.  One CTA per 256 frontiers alone is
// 85 CTAs at C4's 21 791 frontiers, each walking the whole list: 0.3 ms on a third of the GPU.
__global__ void __launch_bounds__(256)
rank_count_kernel(const uint8_t *__restrict__ ach, const uint8_t *__restrict__ bl, const double *__restrict__ len, uint32_t n,
                  uint32_t part, uint32_t *rank) {
  __shared__ double s_len[256];
  __shared__ uint8_t s_ok[256];
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  const bool mine = i < n && ach[i] && !(bl && bl[i]);
  const double li = mine ? len[i] : 0.0;
  uint32_t r = 0;
  const uint32_t j0 = blockIdx.y * part, j1 = min(n, j0 + part);
  for (uint32_t base = j0; base < j1; base += 256) {
    const uint32_t j = base + threadIdx.x;
    s_ok[threadIdx.x] = j < j1 && ach[j] && !(bl && bl[j]);
    s_len[threadIdx.x] = j < j1 ? len[j] : 0.0;
    __syncthreads();
    if (mine) {
      const uint32_t m = min(256u, j1 - base);
      for (uint32_t t = 0; t < m; ++t) {
        const double lj = s_len[t];
        r += (s_ok[t] && (lj < li || (lj == li && base + t < i))) ? 1u : 0u;
      }
    }
    __syncthreads();
  }
  if (mine && r) atomicAdd(rank + i, r);
}
__global__ void rank_scatter_kernel(const uint8_t *__restrict__ ach, const uint8_t *__restrict__ bl, const uint32_t *__restrict__ rank,
                                    uint32_t n, uint32_t *__restrict__ sorted) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n && ach[i] && !(bl && bl[i])) sorted[rank[i]] = i;
}

// The local/global walk of getFilteredFrontiersN (:186-226) in closed form.
__global__ void selection_kernel(const SelectParams P, const uint32_t *__restrict__ sorted, SelectMeta *meta) {
  fitslam_selection s;
  memset(&s, 0, sizeof(s));
  s.closest_global = -1;
  const uint32_t E = meta->n_eligible, L = meta->n_local_range, n = P.n_local_area;
  s.n_sorted = E;
  if (E > 0) {
    const bool need_to_pop = (E == L) && (E <= n);
    uint32_t n_local = min(L, n);
    if (need_to_pop && n_local) --n_local;
    uint32_t n_global = E - L, gb = n_global ? L : E;
    const uint32_t closest = E > L ? sorted[L] : sorted[E - 1];
    if (n_local == 0 && n_global == 0) { n_global = 1; gb = 0; }
    s.n_local = n_local;
    s.n_global = n_global;
    s.global_begin = gb;
    s.closest_global = (int32_t)closest;
  }
  meta->sel = s;
}

// Enqueue the stage on the context stream.  ach/len/head/bl are device arrays (bl may be null).
fitslam_status select_device(fitslam_ctx *c, uint32_t n, const fitslam_arrival *d_arr, const uint8_t *d_ach,
                             const uint8_t *d_bl, const double *d_len, const double *d_head) {
  SelectState *s = sstate(c);
  FS_CUDA(c, s->wc.reserve(n));
  FS_CUDA(c, s->ua.reserve(n));
  FS_CUDA(c, s->ud.reserve(n));
  FS_CUDA(c, s->sorted.reserve(n));
  FS_CUDA(c, s->meta.reserve(1));
  SelectParams P;
  P.alpha = c->prm.alpha; P.beta = c->prm.beta; P.vx = c->prm.vx_max; P.wz = c->prm.wz_max;
  P.max_arrival_info_gt = c->max_arrival_info_gt;
  P.local_radius = c->prm.local_frontier_search_radius;
  P.n = n;
  P.n_local_area = (uint32_t)std::max(0, c->prm.num_frontiers_in_local_area);
  normalise_kernel<<<1, 1024, 0, c->stream>>>(P, d_arr, d_ach, d_bl, d_len, d_head, s->wc.p, s->ua.p, s->ud.p, s->meta.p);
  const uint32_t parts = std::min<uint32_t>(n / 2048u, 16u);
  if (parts >= 2) {
    FS_CUDA(c, s->rank.reserve(n));
    FS_CUDA(c, cudaMemsetAsync(s->rank.p, 0, (size_t)n * sizeof(uint32_t), c->stream));
    const uint32_t part = ((n + parts - 1u) / parts + 255u) / 256u * 256u;
    rank_count_kernel<<<dim3((n + 255) / 256, (n + part - 1u) / part), 256, 0, c->stream>>>(d_ach, d_bl, d_len, n, part, s->rank.p);
    rank_scatter_kernel<<<(n + 255) / 256, 256, 0, c->stream>>>(d_ach, d_bl, s->rank.p, n, s->sorted.p);
    ++c->launches;
  } else {
    rank_kernel<<<(n + 255) / 256, 256, 0, c->stream>>>(d_ach, d_bl, d_len, n, s->sorted.p);
  }
  selection_kernel<<<1, 1, 0, c->stream>>>(P, s->sorted.p, s->meta.p);
  c->launches += 3;
  FS_CUDA(c, cudaGetLastError());
  return FITSLAM_OK;
}

fitslam_status plan_euclidean_device(fitslam_ctx *c, uint32_t n, const fitslam_frontier *d_fr, const fitslam_arrival *d_arr,
                                     const double *d_goal_yaw, double rx, double ry, double ryaw) {
  SelectState *s = sstate(c);
  FS_CUDA(c, s->ach.reserve(n));
  FS_CUDA(c, s->len.reserve(n));
  FS_CUDA(c, s->head.reserve(n));
  init_achievable_kernel<<<(n + 255) / 256, 256, 0, c->stream>>>(d_arr, s->ach.p, n);
  plan_euclidean_kernel<<<(n + 255) / 256, 256, 0, c->stream>>>(d_fr, d_goal_yaw, s->ach.p, s->len.p, s->head.p, n, rx, ry, ryaw);
  c->launches += 2;
  FS_CUDA(c, cudaGetLastError());
  return FITSLAM_OK;
}

}  // namespace fitslam

using namespace fitslam;

extern "C" fitslam_status fitslam_normalise_select(fitslam_ctx *c, const double *path_length_m, const double *path_heading,
                                                   const uint8_t *blacklisted, uint32_t n, double *weighted_cost,
                                                   double *arrival_utility, double *distance_utility, uint8_t *achievable,
                                                   uint32_t *sorted, fitslam_selection *sel) {
  if (!c || !path_length_m || !path_heading) return fail(c, FITSLAM_E_INVALID, "normalise_select: null argument");
  if (n == 0) return fail(c, FITSLAM_E_NO_FRONTIERS, "No frontiers found from frontier search.");
  if (n != c->n_frontiers || c->arrival.cap < n)
    return fail(c, FITSLAM_E_INVALID, "normalise_select: n differs from the list scored by the last fitslam_score_arrival");
  cudaSetDevice(c->device);
  SelectState *s = sstate(c);
  FS_CUDA(c, s->len.reserve(n));
  FS_CUDA(c, s->head.reserve(n));
  FS_CUDA(c, s->bl.reserve(n));
  FS_CUDA(c, s->ach.reserve(n));
  FS_CUDA(c, cudaMemcpyAsync(s->len.p, path_length_m, (size_t)n * 8, cudaMemcpyDefault, c->stream));
  FS_CUDA(c, cudaMemcpyAsync(s->head.p, path_heading, (size_t)n * 8, cudaMemcpyDefault, c->stream));
  if (blacklisted) FS_CUDA(c, cudaMemcpyAsync(s->bl.p, blacklisted, n, cudaMemcpyDefault, c->stream));
  init_achievable_len_kernel<<<(n + 255) / 256, 256, 0, c->stream>>>(c->arrival.p, s->len.p, s->ach.p, n);
  ++c->launches;
  fitslam_status st = select_device(c, n, c->arrival.p, s->ach.p, blacklisted ? s->bl.p : nullptr, s->len.p, s->head.p);
  if (st != FITSLAM_OK) return st;
  SelectMeta m;
  FS_CUDA(c, cudaMemcpyAsync(&m, s->meta.p, sizeof(m), cudaMemcpyDeviceToHost, c->stream));
  if (weighted_cost) FS_CUDA(c, cudaMemcpyAsync(weighted_cost, s->wc.p, (size_t)n * 8, cudaMemcpyDefault, c->stream));
  if (arrival_utility) FS_CUDA(c, cudaMemcpyAsync(arrival_utility, s->ua.p, (size_t)n * 8, cudaMemcpyDefault, c->stream));
  if (distance_utility) FS_CUDA(c, cudaMemcpyAsync(distance_utility, s->ud.p, (size_t)n * 8, cudaMemcpyDefault, c->stream));
  if (achievable) FS_CUDA(c, cudaMemcpyAsync(achievable, s->ach.p, n, cudaMemcpyDefault, c->stream));
  FS_CUDA(c, cudaStreamSynchronize(c->stream));
  if (sorted && m.n_eligible) FS_CUDA(c, cudaMemcpy(sorted, s->sorted.p, (size_t)m.n_eligible * 4, cudaMemcpyDefault));
  if (sel) *sel = m.sel;
  if (m.err_key != 0xffffffffu) {
    const bool arrival = (m.err_key & 1u) == 0;
    return fail(c, arrival ? FITSLAM_E_ARRIVAL_UTILITY : FITSLAM_E_COST_OUT_OF_BOUNDS,
                std::string(arrival ? "ARRIVAL UTILITY ERROR! MORE THAN 1" : "Cost out of bounds") + " (frontier " +
                    std::to_string(m.err_key >> 1) + ")");
  }
  return FITSLAM_OK;
}
