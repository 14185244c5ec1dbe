// frontier.cu -- stage S: frontier extraction.
//
// Replaces FrontierSearch::searchFrom / buildNewFrontier / isNewFrontierCell / getAllFrontiers
// (DEP/src/FrontierSearch.cpp:21-254), getCentroidOfCells + SortByMedianFunctor
// (DEP/include/frontier_exploration/FrontierSearch.hpp:85-184) and nearestFreeCell (DEP/src/Helpers.cpp:285-330).
//
// The reference is two nested FIFO wavefronts; what it computes is ("stage S"):
//   1. the static frontier mask M (isNewFrontierCell without the visited flag)                 -> mask_kernel
//   2. the 8-connected components of M                                                         -> union-find kernels
//   3. which components are found, by which seed cell, and in which order: the first (dequeue rank of an
//      expanded reachable cell, neighbour slot) that touches the component                      -> rank_bfs_kernel
//   4. per found component, the cells in inner-wavefront order from the seed, cut into runs of
//      max_frontier_cluster_size + 1                                                           -> replay_kernel
//   5. per run, centroid / offset rule / libstdc++ std::sort by angle / median cell             -> chunk_kernel
// Only step 3 is order-sensitive at map scale; it is a level-synchronous breadth-first pass that keeps every level in
// exact FIFO order (children claim with atomicMin on a key that encodes the parent's queue position and the
// neighbour slot; an order-preserving compaction builds the next level).
#include "ctx.h"

#include <cmath>
#include <cstring>

namespace fitslam {

static constexpr uint32_t KEY_INF = 0xFFFFFFFFu;
static constexpr uint32_t KEY_VISITED = 0xFFFFFFFEu;  // replay_kernel: mask cell already flagged
static constexpr int BFS_THREADS = 1024;
static constexpr int MAX_CHUNK = 64;  // max_frontier_cluster_size + 1 <= 64

struct Counters {
  uint32_t n_mask;         // cells in the frontier mask
  uint32_t n_comp;         // components touched by the search
  uint32_t n_cells;        // cells of those components (every_frontier list)
  uint32_t n_chunks;       // emitted runs (FrontierSearch.cpp:155-199), kept or not
  uint32_t n_frontiers;    // runs with size > min_frontier_cluster_size
  uint32_t bfs_levels;
  uint32_t start_cell;
  uint32_t overflow;       // bit0 mask list, bit1 bfs queue, bit2 component table
  uint32_t ambiguous;      // chunk sorts with two angles within 4 ulp, or a comparator walk off the array
  uint32_t expanded;       // reachable cells dequeued by the outer wavefront
  uint32_t pad[6];
};

struct SearchParams {
  MapView map;
  uint32_t wpr;            // mask words per row
  uint32_t lethal;         // lethal_threshold
  uint32_t robot_cell;     // getIndex(mx, my) of the robot
  double robot_x, robot_y;
  double dist_limit;       // max_frontier_distance + max_frontier_cluster_size * resolution * 1.414
  int32_t min_size, max_size;
  uint32_t mask_cap, queue_cap, comp_cap;
};

struct FrontierState {
  DevBuf<uint32_t> claim;     // per cell: outer-wavefront key (reachable cells) / first-touch key (mask cells)
  DevBuf<uint32_t> slot;      // per cell: index into the mask-cell list (valid where the mask bit is set)
  DevBuf<uint32_t> bits;      // frontier mask, wpr words per row
  DevBuf<uint32_t> mcell;     // mask-cell list (unordered)
  DevBuf<uint32_t> parent;    // union-find over list indices
  DevBuf<uint32_t> rkey;      // per root: min touch key of the component
  DevBuf<uint32_t> rsize;     // per root: cell count
  DevBuf<uint32_t> queue;     // 2 x queue_cap
  DevBuf<uint32_t> ckey, cseed, csize, corder;       // component table (unordered) + discovery order
  DevBuf<uint32_t> ccell0, cchunk0, cfront0;         // per component (discovery order): offsets
  DevBuf<uint32_t> cells;     // every_frontier list, discovery order
  DevBuf<uint32_t> chunk_comp;                       // per emitted run: component ordinal
  DevBuf<Counters> counters;
  DevBuf<double> centroids;   // per frontier: centroid after the offset rule (stage dump)
  double search_distance = -1.0;  // < 0: use params.max_frontier_distance
  uint32_t mask_cap = 0, queue_cap = 0, comp_cap = 0;
};

// ------------------------------------------------------------------------------------------------ helpers
__device__ __forceinline__ bool mask_bit(const uint32_t *bits, uint32_t wpr, uint32_t mx, uint32_t my) {
  return (bits[(size_t)my * wpr + (mx >> 5)] >> (mx & 31u)) & 1u;
}
__device__ __forceinline__ uint32_t ld_cg(const uint32_t *p) { return __ldcg(p); }

// ------------------------------------------------------------------------------------------------ K1: mask
// isNewFrontierCell on the static map (DEP/src/FrontierSearch.cpp:218-249, predicates FrontierSearch.hpp:129-142):
// cost == 255, no 4-neighbour with lethal_threshold <= cost < 255, at least one 4-neighbour with cost < lethal_threshold.
// One thread per 32-cell mask word; rows are read as 16-byte vectors (pitch is a multiple of 16).
__global__ void __launch_bounds__(256) mask_kernel(const SearchParams S, uint32_t *__restrict__ bits,
                                                   uint32_t *__restrict__ slot, uint32_t *__restrict__ mcell,
                                                   Counters *__restrict__ cnt) {
  const uint32_t W = S.map.W, H = S.map.H, pitch = S.map.pitch;
  const uint64_t n_words = (uint64_t)S.wpr * H;
  const uint64_t gid = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  uint32_t word = 0, my = 0, mx0 = 0;
  if (gid < n_words) {
    my = (uint32_t)(gid / S.wpr);
    mx0 = (uint32_t)(gid - (uint64_t)my * S.wpr) * 32u;
    const uint8_t *row = S.map.data + (size_t)my * pitch;
    // centre row bytes mx0-1 .. mx0+32, up/down rows mx0 .. mx0+31 (the allocation is padded past the last row)
    uint8_t c[34], u[32], d[32];
    const bool full = mx0 + 32u <= pitch;
    if (full) {
      const uint4 *rc = reinterpret_cast<const uint4 *>(row + mx0);
      uint4 a = __ldg(rc), b = __ldg(rc + 1);
      memcpy(c + 1, &a, 16);
      memcpy(c + 17, &b, 16);
    } else {
      for (int i = 0; i < 32; ++i) c[1 + i] = (mx0 + i < pitch) ? __ldg(row + mx0 + i) : 0;
    }
    c[0] = mx0 > 0 ? __ldg(row + mx0 - 1) : 0;
    c[33] = (mx0 + 32u < W) ? __ldg(row + mx0 + 32) : 0;
    const bool has_up = my > 0, has_dn = my + 1 < H;
    if (full && has_up) {
      const uint4 *r = reinterpret_cast<const uint4 *>(row - pitch + mx0);
      uint4 a = __ldg(r), b = __ldg(r + 1);
      memcpy(u, &a, 16);
      memcpy(u + 16, &b, 16);
    } else {
      for (int i = 0; i < 32; ++i) u[i] = (has_up && mx0 + i < pitch) ? __ldg(row - pitch + mx0 + i) : 0;
    }
    if (full && has_dn) {
      const uint4 *r = reinterpret_cast<const uint4 *>(row + pitch + mx0);
      uint4 a = __ldg(r), b = __ldg(r + 1);
      memcpy(d, &a, 16);
      memcpy(d + 16, &b, 16);
    } else {
      for (int i = 0; i < 32; ++i) d[i] = (has_dn && mx0 + i < pitch) ? __ldg(row + pitch + mx0 + i) : 0;
    }
    const uint32_t thr = S.lethal;
#pragma unroll
    for (int i = 0; i < 32; ++i) {
      const uint32_t mx = mx0 + i;
      if (mx >= W || c[1 + i] != 255) continue;
      bool fr = false, le = false;
      if (mx > 0) { const uint32_t v = c[i]; fr |= v < thr; le |= (v >= thr && v != 255); }
      if (mx + 1 < W) { const uint32_t v = c[2 + i]; fr |= v < thr; le |= (v >= thr && v != 255); }
      if (has_up) { const uint32_t v = u[i]; fr |= v < thr; le |= (v >= thr && v != 255); }
      if (has_dn) { const uint32_t v = d[i]; fr |= v < thr; le |= (v >= thr && v != 255); }
      if (fr && !le) word |= 1u << i;
    }
    bits[gid] = word;
  }
  // append the set cells to the list: one atomicAdd per warp
  const uint32_t lane = threadIdx.x & 31u;
  const uint32_t n = __popc(word);
  uint32_t incl = n;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t v = __shfl_up_sync(0xffffffffu, incl, o);
    if ((int)lane >= o) incl += v;
  }
  const uint32_t total = __shfl_sync(0xffffffffu, incl, 31);
  uint32_t base = 0;
  if (total) {
    if (lane == 31) base = atomicAdd(&cnt->n_mask, total);
    base = __shfl_sync(0xffffffffu, base, 31);
    uint32_t pos = base + incl - n;
    while (word) {
      const int i = __ffs(word) - 1;
      word &= word - 1;
      const uint32_t cell = my * W + mx0 + i;
      if (pos < S.mask_cap) {
        mcell[pos] = cell;
        slot[cell] = pos;
      } else {
        atomicOr(&cnt->overflow, 1u);
      }
      ++pos;
    }
  }
}

// ------------------------------------------------------------------------------------------------ K2: ranked wavefront
// Block-wide exclusive scan of 0/1 flags over BFS_THREADS threads; returns this thread's offset, *total = sum.
__device__ __forceinline__ uint32_t block_scan_flags(bool flag, uint32_t *warp_sums, uint32_t *total) {
  const uint32_t lane = threadIdx.x & 31u, wid = threadIdx.x >> 5;
  const uint32_t ballot = __ballot_sync(0xffffffffu, flag);
  const uint32_t excl = __popc(ballot & ((1u << lane) - 1u));
  if (lane == 0) warp_sums[wid] = __popc(ballot);
  __syncthreads();
  if (wid == 0) {
    uint32_t v = warp_sums[lane];
    uint32_t incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
      if ((int)lane >= o) incl += t;
    }
    warp_sums[lane] = incl - v;
    if (lane == 31) warp_sums[32] = incl;
  }
  __syncthreads();
  const uint32_t r = warp_sums[wid] + excl;
  *total = warp_sums[32];
  __syncthreads();
  return r;
}

// neighbour `s` of cell (mx, my) in the reference's order: -1, +1, -W, +W, -1-W, -1+W, +1-W, +1+W
// (DEP/src/Helpers.cpp:185-255).  Returns false where the reference's bounds tests drop the neighbour.
__device__ __forceinline__ bool nbr_of(uint32_t mx, uint32_t my, uint32_t W, uint32_t H, int s, uint32_t &nx, uint32_t &ny) {
  const int dx = (s == 0 || s == 4 || s == 5) ? -1 : ((s == 1 || s == 6 || s == 7) ? 1 : 0);
  const int dy = (s == 2 || s == 4 || s == 6) ? -1 : ((s == 3 || s == 5 || s == 7) ? 1 : 0);
  const int x = (int)mx + dx, y = (int)my + dy;
  if (x < 0 || y < 0 || x >= (int)W || y >= (int)H) return false;
  nx = (uint32_t)x;
  ny = (uint32_t)y;
  return true;
}

// One level-synchronous pass in exact FIFO order.  NB = 4: the outer wavefront of searchFrom
// (DEP/src/FrontierSearch.cpp:58-94); NB = 8: nearestFreeCell (DEP/src/Helpers.cpp:285-330), every cell traversable,
// stops at the first dequeued cell with cost < lethal_threshold.
template <int NB>
__device__ void ranked_wavefront(const SearchParams &S, const uint32_t *__restrict__ bits, uint32_t *claim,
                                 uint32_t *q0, uint32_t *q1, uint32_t start, uint32_t *sh, Counters *cnt,
                                 uint32_t *found_cell) {
  constexpr int LOG_NB = NB == 4 ? 2 : 3;
  const uint32_t W = S.map.W, H = S.map.H;
  uint32_t *warp_sums = sh;       // 33 words
  uint32_t *s_min = sh + 40;      // NB == 8: first queue position with a free cell
  uint32_t n = 1, base = 1, levels = 0, expanded = 0;
  uint32_t *q = q0, *qn = q1;
  if (threadIdx.x == 0) {
    q[0] = start;
    claim[start] = 0;
  }
  __syncthreads();
  while (n > 0) {
    if (NB == 8) {  // dequeue order test: first cell of this level with cost < val wins
      if (threadIdx.x == 0) *s_min = KEY_INF;
      __syncthreads();
      for (uint32_t i = threadIdx.x; i < n; i += BFS_THREADS) {
        const uint32_t c = q[i];
        const uint32_t my = c / W, mx = c - my * W;
        if (S.map.data[(size_t)my * S.map.pitch + mx] < S.lethal) atomicMin(s_min, i);
      }
      __syncthreads();
      if (*s_min != KEY_INF) {
        if (threadIdx.x == 0) *found_cell = q[*s_min];
        __syncthreads();
        break;
      }
    }
    const uint32_t total = n << LOG_NB;
    expanded += n;
    // phase A: every (queue position, slot) pair bids for its neighbour with key base + pos * NB + slot
    for (uint32_t j = threadIdx.x; j < total; j += BFS_THREADS) {
      const uint32_t c = q[j >> LOG_NB];
      const uint32_t my = c / W, mx = c - my * W;
      uint32_t nx, ny;
      if (!nbr_of(mx, my, W, H, (int)(j & (NB - 1)), nx, ny)) continue;
      const uint32_t nb = ny * W + nx;
      if (NB == 8) {
        if (ld_cg(claim + nb) >= base) atomicMin(claim + nb, base + j);
      } else {
        const uint32_t m = S.map.data[(size_t)ny * S.map.pitch + nx];
        if (m < 254u) {  // map_[nbr] < LETHAL_OBSTACLE && !visited_flag[nbr]
          if (ld_cg(claim + nb) >= base) atomicMin(claim + nb, base + j);
        } else if (m == 255u && mask_bit(bits, S.wpr, nx, ny)) {
          atomicMin(claim + nb, base + j);  // first touch of a frontier-mask cell (isNewFrontierCell branch)
        }
      }
    }
    __threadfence();
    __syncthreads();
    // phase B: winners, in key order, form the next level
    uint32_t n_next = 0;
    for (uint32_t j0 = 0; j0 < total; j0 += BFS_THREADS) {
      const uint32_t j = j0 + threadIdx.x;
      bool push = false;
      uint32_t nb = 0;
      if (j < total) {
        const uint32_t c = q[j >> LOG_NB];
        const uint32_t my = c / W, mx = c - my * W;
        uint32_t nx, ny;
        if (nbr_of(mx, my, W, H, (int)(j & (NB - 1)), nx, ny)) {
          nb = ny * W + nx;
          if (NB == 8) {
            push = ld_cg(claim + nb) == base + j;
          } else {
            const uint32_t m = S.map.data[(size_t)ny * S.map.pitch + nx];
            if (m < 254u && ld_cg(claim + nb) == base + j) {
              // visited now; enqueued only inside the search radius (FrontierSearch.cpp:72-76)
              const double wx = S.map.ox + (nx + 0.5) * S.map.res;
              const double wy = S.map.oy + (ny + 0.5) * S.map.res;
              const double ddx = S.robot_x - wx, ddy = S.robot_y - wy;
              push = sqrt(ddx * ddx + ddy * ddy) < S.dist_limit;
            }
          }
        }
      }
      uint32_t tile_total;
      const uint32_t off = block_scan_flags(push, warp_sums, &tile_total);
      if (push) {
        if (n_next + off < S.queue_cap) qn[n_next + off] = nb;
        else atomicOr(&cnt->overflow, 2u);
      }
      n_next += tile_total;
    }
    __threadfence();
    __syncthreads();
    base += total;
    n = min(n_next, S.queue_cap);
    uint32_t *t = q;
    q = qn;
    qn = t;
    ++levels;
  }
  if (threadIdx.x == 0 && NB == 4) {
    cnt->bfs_levels = levels;
    cnt->expanded = expanded;
  }
  // NB == 8: undo the claims so the outer wavefront starts from a clean array (Chebyshev ball of radius `levels`)
  if (NB == 8) {
    __syncthreads();
    const uint32_t sy = start / W, sx = start - sy * W;
    const int r = (int)levels + 1;
    const int side = 2 * r + 1;
    for (long i = threadIdx.x; i < (long)side * side; i += BFS_THREADS) {
      const int y = (int)sy - r + (int)(i / side), x = (int)sx - r + (int)(i % side);
      if (x >= 0 && y >= 0 && x < (int)W && y < (int)H) claim[(size_t)y * W + x] = KEY_INF;
    }
    __threadfence();
    __syncthreads();
  }
}

__global__ void __launch_bounds__(BFS_THREADS) rank_bfs_kernel(const SearchParams S, const uint32_t *__restrict__ bits,
                                                               uint32_t *claim, uint32_t *queue, Counters *cnt) {
  __shared__ uint32_t sh[64];
  __shared__ uint32_t s_start;
  uint32_t *q0 = queue, *q1 = queue + S.queue_cap;
  // nearestFreeCell(clear, pos, lethal_threshold_): trivial when the robot stands on a free cell
  const uint32_t ry = S.robot_cell / S.map.W, rx = S.robot_cell - ry * S.map.W;
  if (threadIdx.x == 0) s_start = S.robot_cell;
  __syncthreads();
  if (S.map.data[(size_t)ry * S.map.pitch + rx] >= S.lethal) {
    ranked_wavefront<8>(S, bits, claim, q0, q1, S.robot_cell, sh, cnt, &s_start);
    __syncthreads();
  }
  const uint32_t start = s_start;  // falls back to pos when no cell qualifies (FrontierSearch.cpp:47-54)
  if (threadIdx.x == 0) cnt->start_cell = start;
  ranked_wavefront<4>(S, bits, claim, q0, q1, start, sh, cnt, nullptr);
}

}  // namespace fitslam
