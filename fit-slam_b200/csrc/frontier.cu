// frontier.cu -- stage S: frontier extraction.
//
// Replaces FrontierSearch::searchFrom / buildNewFrontier / isNewFrontierCell / getAllFrontiers
// (DEP/src/FrontierSearch.cpp:21-254), getCentroidOfCells + SortByMedianFunctor
// (DEP/include/frontier_exploration/FrontierSearch.hpp:85-184) and nearestFreeCell (DEP/src/Helpers.cpp:285-330).
//
// The reference is two nested FIFO wavefronts; what it computes is ("stage S"):
//   1. the static frontier mask M (isNewFrontierCell without the visited flag) and the traversable mask
//      (cost < 254)                                                                             -> mask_kernel
//   2. the 8-connected components of M                                                          -> union-find kernels
//   3. which components are found, by which seed cell, and in which order: the first (dequeue rank of an
//      expanded reachable cell, neighbour slot) that touches the component                      -> rank_bfs_kernel
//   4. per found component, the cells in inner-wavefront order from the seed, cut into runs of
//      max_frontier_cluster_size + 1                                                            -> replay_kernel
//   5. per run, centroid / offset rule / libstdc++ std::sort by angle / median cell             -> chunk_kernel
// Steps 1, 2 and 5 are order-free.  Steps 3 and 4 must reproduce FIFO order exactly, because the chunk boundaries
// and therefore sizes and goal cells depend on it: they are level-synchronous passes in which a cell
// decides locally whether it is the first in queue order to reach a neighbour (it reads the ranks of the neighbour's
// other neighbours; no atomics, no sort) and an order-preserving block scan appends the winners to the queue.
#include "ctx.h"
#include "atan2_cr.h"

#include <cooperative_groups.h>

#include <algorithm>
#include <chrono>
#include <cfloat>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>

namespace fitslam {

// `claim` encoding, one uint32 per map cell:
//   0xFFFFFFFF            untouched
//   0xFFFFFFFE            traversable cell visited by the outer wavefront but beyond the search radius (not enqueued)
//   [0, 2^29)             traversable cell: its dequeue rank in the outer wavefront
//   0x40000000 | r        frontier-mask cell: its position r in the inner wavefront of its component
//   0x80000000 | key      frontier-mask cell: smallest (outer rank * 4 + slot) that touched it
static constexpr uint32_t KEY_INF = 0xFFFFFFFFu;
static constexpr uint32_t KEY_FAR = 0xFFFFFFFEu;
static constexpr uint32_t INNER_BIT = 0x40000000u;
static constexpr uint32_t TOUCH_BIT = 0x80000000u;
static constexpr uint32_t MAX_RANK = (1u << 29) - 2;
static constexpr int BFS_THREADS = 1024;
static constexpr int OW_THREADS = 1024;      // shared-memory outer wavefront: one thread per queue position
static constexpr int REPLAY_THREADS = 128;
static constexpr int MAX_CHUNK = 64;  // max_frontier_cluster_size + 1 <= 64
// shared-memory wavefronts (the default path; the global-memory kernels below them are the fallback for levels or
// components that do not fit)
static constexpr uint32_t OW_QCAP = 8192;    // outer wavefront: widest level the shared-memory queue holds (128 KB; the rest of the SM's 256 KB stays L1)
static constexpr uint32_t OW_SMEM = 2 * OW_QCAP * 8;
// The outer wavefront publishes its progress OW_PARTS - 1 times at most, at the level counts of SearchParams::ow_part_end (the
// last part runs to the end).  After part k the components it touched have their final seed (later touches carry larger keys), so
// their inner wavefronts replay on side streams, behind a one-thread kernel that polls the flag, while the wavefront
// goes on: the long thin components that bound the replay time are usually touched long before it ends.  (Separate
// launches per part were tried first: once the side streams fill the other SMs, the next part waits for a whole SM.)
static constexpr int OW_LARGE_STREAMS = 4;
#ifndef FS_OW_PARTS
#define FS_OW_PARTS 16
#endif
static constexpr int OW_PARTS = FS_OW_PARTS;          // levels per part: search radius in cells / OW_PARTS (SearchParams::ow_part_levels)
static constexpr uint32_t IW_CELLS = 4096;   // inner wavefront: largest component replayed out of shared memory
static constexpr uint32_t IW_SMALL = 128;    // size classes: <= IW_SMALL cells (many warps per SM), <= IW_CELLS cells (two per SM)
static constexpr uint32_t IW_BITS_SMALL = 1024, IW_BITS_LARGE = 8192;  // words of the component's mask window held in shared memory
static constexpr uint32_t iw_smem(uint32_t cells, uint32_t bit_words) { return (2 * 2 * cells + cells + bit_words) * 4; }
static constexpr uint32_t HKEY_EMPTY = 0xFFFFFFFFu;

struct Counters {
  uint32_t n_mask;         // cells in the frontier mask
  uint32_t n_comp;         // components touched by the search
  uint32_t n_cells;        // cells of those components (every_frontier list)
  uint32_t n_frontiers;    // runs with size > min_frontier_cluster_size
  uint32_t bfs_levels;
  uint32_t start_cell;
  uint32_t overflow;       // bit0 mask list, bit1 bfs queue, bit2 component table
  uint32_t ambiguous;      // chunk sorts handed to the host (see chunk_kernel): entries of the fix list
  uint32_t expanded;       // reachable cells dequeued by the outer wavefront
  uint32_t visited;        // queue entries + far cells (cleanup bound)
  uint32_t dirty_queue;    // entries of queue[] whose claim words (and their 4 neighbours') the cleanup must reset
  uint32_t ow_fallback;    // the shared-memory outer wavefront met a level it cannot hold: 2 = repeat with the spill instantiation, 1 = with rank_bfs_kernel
  uint32_t pad0[3];
  uint32_t widest_level;   // cells in the widest level of the shared-memory outer wavefront
  uint32_t part_flag[OW_PARTS];        // part k of the outer wavefront is complete (published by the kernel, polled by wait_part_kernel)
  uint32_t part_end[OW_PARTS];         // cells dequeued when part k ended: touch keys below 4 * part_end[k] are final after it
  uint32_t comp_done;                  // components ordered and scanned so far
  uint32_t coll_c0[OW_PARTS], coll_c1[OW_PARTS];  // components [c0, c1) of collection k (the one after the last part: OW_PARTS - 1)
  uint32_t coll_fr[OW_PARTS][2];       // frontiers [first, end) of collection k
  uint32_t cell_carry, front_carry;    // running totals of comp_scan_kernel across collections
  uint32_t pad[1];
  unsigned long long prof[8];  // FITSLAM_PROF: [4] passes, [5] cycles of the outer wavefront kernel
};

struct SearchParams {
  MapView map;
  uint32_t wpr;            // mask words per row
  uint32_t lethal;         // lethal_threshold
  uint32_t robot_cell;     // getIndex(mx, my) of the robot
  double robot_x, robot_y;
  double dist_limit;       // max_frontier_distance + max_frontier_cluster_size * resolution * 1.414
  int64_t near_cells_sq;   // (|dx| + 1)^2 + (|dy| + 1)^2 < near_cells_sq (in cells, from the robot cell) => surely inside dist_limit
  int64_t far_cells_sq;    // max(|dx| - 1, 0)^2 + max(|dy| - 1, 0)^2 > far_cells_sq => surely outside
  int32_t min_size, max_size;
  uint32_t mask_cap, queue_cap, comp_cap, cell_cap, front_cap;
  uint32_t use_smem;       // 1: shared-memory wavefronts first, global-memory kernels only for what does not fit
  uint32_t ow_wide;        // 1: levels wider than ow_qcap spill to the global queue (maps up to 16384 cells a side)
  uint32_t ow_qcap;        // widest level the shared-memory outer wavefront keeps (<= OW_QCAP; FITSLAM_OW_QCAP lowers it in tests)
  uint32_t ow_part_levels; // levels per part of the outer wavefront where the schedule below is not used (first search on a map)
  uint32_t ow_part_end[OW_PARTS];  // part k of the outer wavefront ends when this many levels are done (ascending; see enqueue_search)
  uint32_t early_cap;      // frontiers one early scoring launch covers (what exceeds it waits for the launches after the search)
  uint32_t early_limit;    // frontier indices at or above this are not scored early (EarlyHook::limit)
  uint32_t fix_all;        // FITSLAM_FIX_ALL=1: every chunk's median sort goes to the host (test of that path)
};

struct FrontierState {
  DevBuf<uint32_t> claim;     // per cell, see the encoding above; all KEY_INF between calls
  DevBuf<uint32_t> slot;      // per cell: index into the mask-cell list (valid where the mask bit is set)
  DevBuf<uint32_t> bits;      // frontier mask, wpr words per row
  DevBuf<uint32_t> trav;      // traversable mask (cost < 254), same layout
  DevBuf<uint2> tm;           // {trav word, frontier-mask word} interleaved: one 8-byte load classifies a neighbour
  DevBuf<uint32_t> ebits;     // expandable cells (traversable and inside the search radius), same layout
  int ow_cluster = -1;        // CTAs per cluster of ow_cluster_kernel (0: not available, use the single-CTA kernel); -1: not probed yet
  uint32_t last_ow_ctas = 0;  // CTAs the last search's outer wavefront ran on (1, or the cluster size)
  uint32_t ow_cluster_min = 1500;  // ... used when the last search's widest level had at least this many cells (FITSLAM_OW_CLUSTER_MIN)
  DevBuf<uint16_t> rk;        // outer wavefront: class of a cell, then its position inside its level (ow_rk_* layout); the search disc's
                              // tiles are rewritten before every search
  uint64_t ow_cells = 0;      // padded cells for which rk is initialised
  DevBuf<uint32_t> mcell;     // mask-cell list (unordered)
  DevBuf<uint32_t> parent;    // union-find over list indices
  DevBuf<uint32_t> rkey;      // per root: min touch key of the component
  DevBuf<uint32_t> rsize;     // per root: cell count
  DevBuf<uint32_t> rbox;      // per root: bounding box of the component (4 words: min x, min y, max x, max y)
  DevBuf<uint32_t> cbox, obox;  // the same per found component, append order / discovery order
  DevBuf<uint32_t> queue;     // outer wavefront, all levels back to back in dequeue order
  DevBuf<uint32_t> ckey, cseed, csize;               // component table, unordered (append order)
  DevBuf<uint32_t> oseed, osize, ocell0, ofront0;    // component table in discovery order + offsets
  DevBuf<uint32_t> cells;     // every_frontier list, discovery order
  DevBuf<double> centroids;   // per frontier: centroid after the offset rule (stage dump)
  DevBuf<uint8_t> ftag;       // per frontier: 0 = not built yet, else the launch that scores it (ctx.h: TAG_FINAL, TAG_HOST, early tags)
  DevBuf<uint32_t> fix_list;  // frontiers whose median sort the host redoes with its own libm (chunk_kernel), in append order
  DevBuf<uint32_t> fix_stage; // per fix-list entry: frontier, cell count, cells / the host's answers on the way back
  std::vector<uint32_t> fix_host;
  DevBuf<Counters> counters;
  double search_distance = -1.0;  // < 0: use params.max_frontier_distance
  uint64_t claim_cells = 0;   // cells for which claim is known to be all KEY_INF
  uint32_t mask_cap = 0, queue_cap = 0, comp_cap = 0, cell_cap = 0, front_cap = 0;
  Counters last{};
  cudaEvent_t ev_bfs[7] = {};  // outer wavefront: [0] start, [1] last part done, [3] second stream joined, [4] fallback kernels done; [2] replay done
  cudaEvent_t ev_part[OW_PARTS] = {}, ev_coll[OW_PARTS] = {};  // part k of the outer wavefront done; tables of collection k done
  cudaEvent_t ev_join = nullptr, ev_join_large[OW_LARGE_STREAMS] = {};  // side streams drained
  cudaEvent_t ev_chunk[OW_PARTS][2] = {}, ev_score[1 + OW_LARGE_STREAMS] = {};  // frontiers of collection k, size class c built; scoring streams drained
  cudaStream_t stream2 = nullptr;     // collections that overlap the outer wavefront: tables, small inner wavefronts
  cudaStream_t stream_large[OW_LARGE_STREAMS] = {};  // their large inner wavefronts, collection k on stream k % OW_LARGE_STREAMS: one
                                      // warp walks a thin component for up to milliseconds, and must not hold up the next collection
  cudaStream_t score_stream[1 + OW_LARGE_STREAMS] = {};  // low priority: early scoring (cycle.cu's hook), [0] small class, [1 + i] behind stream_large[i]
  uint32_t ow_parts = OW_PARTS;    // FITSLAM_OW_PARTS=1: one launch, no overlap (ablation)
  double host_enqueue_ms = 0;      // FITSLAM_PROF: host time spent enqueuing the side work of a search
  bool have_ev = false;
  bool attr_set = false;
  int force_global = -1;      // FITSLAM_WAVEFRONT=global selects the global-memory wavefront kernels (ablation / fallback test)
  bool prefer_wide = false;   // the last search met a level wider than the shared-memory queue: start with the spill instantiation
  float bfs_ms = 0, replay_ms = 0;
};

static FrontierState *fr_state(fitslam_ctx *c) {
  if (!c->fs) c->fs = new FrontierState();
  return c->fs;
}

void destroy_frontier_state(fitslam_ctx *c) {
  if (!c->fs) return;
  FrontierState *s = c->fs;
  s->claim.release(); s->slot.release(); s->bits.release(); s->trav.release(); s->tm.release(); s->rk.release(); s->ebits.release(); s->mcell.release();
  s->parent.release(); s->rkey.release(); s->rsize.release(); s->queue.release();
  s->rbox.release(); s->cbox.release(); s->obox.release();
  s->ckey.release(); s->cseed.release(); s->csize.release();
  if (s->have_ev) {
    for (auto &e : s->ev_bfs) cudaEventDestroy(e);
    for (auto &e : s->ev_part) cudaEventDestroy(e);
    for (auto &e : s->ev_coll) cudaEventDestroy(e);
    cudaEventDestroy(s->ev_join);
    for (auto &e : s->ev_join_large) cudaEventDestroy(e);
    for (auto &row : s->ev_chunk)
      for (auto &e : row) cudaEventDestroy(e);
    for (auto &e : s->ev_score) cudaEventDestroy(e);
    cudaStreamDestroy(s->stream2);
    for (auto &st : s->stream_large) cudaStreamDestroy(st);
    for (auto &st : s->score_stream) cudaStreamDestroy(st);
  }
  s->oseed.release(); s->osize.release(); s->ocell0.release(); s->ofront0.release();
  s->cells.release(); s->centroids.release(); s->ftag.release(); s->counters.release(); s->fix_list.release(); s->fix_stage.release();
  delete s;
  c->fs = nullptr;
}

// ------------------------------------------------------------------------------------------------ helpers
__device__ __forceinline__ bool get_bit(const uint32_t *__restrict__ bits, uint32_t wpr, uint32_t mx, uint32_t my) {
  return (__ldg(bits + (size_t)my * wpr + (mx >> 5)) >> (mx & 31u)) & 1u;
}
__device__ __forceinline__ uint32_t ld_cg(const uint32_t *p) { return __ldcg(p); }
__device__ __forceinline__ void prefetch_l1(const void *p) { asm volatile("prefetch.global.L1 [%0];" ::"l"(p)); }

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

// ------------------------------------------------------------------------------------------------ K1: mask
// Which of the 32 cells (mx0 .. mx0 + 31, my) would be enqueued by the outer wavefront (FrontierSearch.cpp:72-76):
// distance from the robot to the cell centre < max_frontier_distance + max_frontier_cluster_size * resolution * 1.414.
// The exact double expression runs only for words that straddle the rim.
__device__ __forceinline__ uint32_t in_radius_word(const SearchParams &S, uint32_t mx0, uint32_t my) {
  const int64_t rx = S.robot_cell % S.map.W, ry = S.robot_cell / S.map.W;
  const int64_t dy = llabs((int64_t)my - ry);
  const int64_t da = llabs((int64_t)mx0 - rx), db = llabs((int64_t)mx0 + 31 - rx);
  const int64_t dfar = max(da, db), dnear = ((int64_t)mx0 <= rx && rx <= (int64_t)mx0 + 31) ? 0 : min(da, db);
  if ((dfar + 1) * (dfar + 1) + (dy + 1) * (dy + 1) < S.near_cells_sq) return 0xFFFFFFFFu;
  const int64_t nx = max(dnear - 1, (int64_t)0), ny = max(dy - 1, (int64_t)0);
  if (nx * nx + ny * ny > S.far_cells_sq) return 0u;
  uint32_t w = 0u;
  const double ddy = S.robot_y - (S.map.oy + (my + 0.5) * S.map.res);
  for (int i = 0; i < 32; ++i) {
    const double ddx = S.robot_x - (S.map.ox + ((mx0 + i) + 0.5) * S.map.res);
    if (sqrt(ddx * ddx + ddy * ddy) < S.dist_limit) w |= 1u << i;
  }
  return w;
}

// isNewFrontierCell on the static map (DEP/src/FrontierSearch.cpp:218-249, predicates FrontierSearch.hpp:129-142):
// cost == 255, no 4-neighbour with lethal_threshold <= cost < 255, at least one 4-neighbour with cost < lethal_threshold.
//
// Every cell's three predicates are evaluated ONCE, four cells at a time on the bytes packed in a 32-bit word, and kept as
// one bit per cell: F (cost < lethal_threshold), U (cost == 255), T (cost < 254); L = neither F nor U.  The neighbour tests are
// then shifts and ORs of whole 32-cell words.  A thread owns one 32-cell column of words over MASK_ROWS consecutive rows and
// slides down them, so a row of the map is read once (plus the two halo rows of the strip), in 16-byte vectors (pitch is a
// multiple of 16).  The first version read the three rows of every word and compared byte by byte: 0.38 TB/s on this part.
static constexpr int MASK_ROWS = 8;
struct RowBits { uint32_t F, L, U, T, edge; };  // edge: bit 0 / 1 = F / L of the cell left of the word, bit 2 / 3 = right of it

// bit 7 of every byte: byte != 0
__device__ __forceinline__ uint32_t swar_nonzero(uint32_t v) { return (((v & 0x7F7F7F7Fu) + 0x7F7F7F7Fu) | v) & 0x80808080u; }
// bit 7 of every byte: byte < the byte of t4 (unsigned)
__device__ __forceinline__ uint32_t swar_less(uint32_t x, uint32_t t4) {
  const uint32_t d = (x | 0x80808080u) - (t4 & 0x7F7F7F7Fu);  // bit 7: low 7 bits of x >= low 7 bits of t (no borrow across bytes)
  return ((~x & t4) | (~(x ^ t4) & ~d)) & 0x80808080u;
}
// bits 7, 15, 23, 31 -> bits 0 .. 3
__device__ __forceinline__ uint32_t swar_pack(uint32_t m) { return (((m >> 7) * 0x01020408u) >> 24) & 0xFu; }

__device__ __forceinline__ RowBits mask_row_bits(const SearchParams &S, int y, uint32_t mx0, uint32_t thr4) {
  RowBits r = {0u, 0u, 0u, 0u, 0u};
  const uint32_t W = S.map.W, pitch = S.map.pitch;
  if (y < 0 || y >= (int)S.map.H) return r;  // no such row: no neighbours there
  const uint8_t *row = S.map.data + (size_t)y * pitch;
  uint32_t x[8];
  if (mx0 + 32u <= pitch) {
    const uint4 *rc = reinterpret_cast<const uint4 *>(row + mx0);
    const uint4 a = __ldg(rc), b = __ldg(rc + 1);
    x[0] = a.x; x[1] = a.y; x[2] = a.z; x[3] = a.w; x[4] = b.x; x[5] = b.y; x[6] = b.z; x[7] = b.w;
  } else {
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      uint32_t v = 0;
      for (int j = 0; j < 4; ++j) {
        const uint32_t mx = mx0 + 4u * k + j;
        if (mx < pitch) v |= (uint32_t)__ldg(row + mx) << (8 * j);
      }
      x[k] = v;
    }
  }
  uint32_t F = 0, U = 0, T = 0;
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    const uint32_t v = x[k];
    F |= swar_pack(swar_less(v, thr4)) << (4 * k);
    U |= swar_pack(~swar_nonzero(~v) & 0x80808080u) << (4 * k);                 // byte == 255
    T |= swar_pack(swar_nonzero(~(v | 0x01010101u))) << (4 * k);               // byte < 254
  }
  const uint32_t valid = (mx0 + 32u <= W) ? 0xFFFFFFFFu : ((1u << (W - mx0)) - 1u);  // (mx0 < W always)
  if (S.lethal > 255u) F = 0xFFFFFFFFu;  // (a threshold no byte reaches: every cell counts as free, none as lethal)
  r.F = F & valid;
  r.U = U & valid;
  r.T = T & valid;
  r.L = ~(F | U) & valid;
  const uint32_t thr = S.lethal;
  if (mx0 > 0) {
    const uint32_t v = __ldg(row + mx0 - 1);
    r.edge |= (v < thr ? 1u : 0u) | ((v >= thr && v != 255u) ? 2u : 0u);
  }
  if (mx0 + 32u < W) {
    const uint32_t v = __ldg(row + mx0 + 32);
    r.edge |= (v < thr ? 4u : 0u) | ((v >= thr && v != 255u) ? 8u : 0u);
  }
  return r;
}

__global__ void __launch_bounds__(256) mask_kernel(const SearchParams S, uint32_t *__restrict__ bits,
                                                   uint32_t *__restrict__ trav, uint32_t *__restrict__ ebits, uint2 *__restrict__ tm,
                                                   uint32_t *__restrict__ slot, uint32_t *__restrict__ mcell,
                                                   Counters *__restrict__ cnt) {
  const uint32_t W = S.map.W, H = S.map.H;
  const uint32_t strips = (H + MASK_ROWS - 1u) / MASK_ROWS;
  const uint64_t gid = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  uint32_t words[MASK_ROWS];
#pragma unroll
  for (int k = 0; k < MASK_ROWS; ++k) words[k] = 0u;
  uint32_t mx0 = 0, y0 = 0, n = 0;
  if (gid < (uint64_t)S.wpr * strips) {
    const uint32_t strip = (uint32_t)(gid / S.wpr), wx = (uint32_t)(gid - (uint64_t)strip * S.wpr);
    mx0 = wx * 32u;
    y0 = strip * MASK_ROWS;
    const uint32_t thr4 = min(S.lethal, 255u) * 0x01010101u;  // (above 255: see mask_row_bits)
    RowBits up = mask_row_bits(S, (int)y0 - 1, mx0, thr4), mid = mask_row_bits(S, (int)y0, mx0, thr4);
#pragma unroll
    for (int k = 0; k < MASK_ROWS; ++k) {
      const uint32_t my = y0 + k;
      if (my >= H) break;
      const RowBits dn = mask_row_bits(S, (int)my + 1, mx0, thr4);
      const uint32_t Fn = up.F | dn.F | (mid.F << 1) | (mid.edge & 1u) | (mid.F >> 1) | ((mid.edge & 4u) << 29);
      const uint32_t Ln = up.L | dn.L | (mid.L << 1) | ((mid.edge >> 1) & 1u) | (mid.L >> 1) | ((mid.edge & 8u) << 28);
      const uint32_t word = mid.U & Fn & ~Ln, tword = mid.T;
      const uint64_t at = (uint64_t)my * S.wpr + wx;
      bits[at] = word;
      trav[at] = tword;
      ebits[at] = tword ? (tword & in_radius_word(S, mx0, my)) : 0u;
      tm[((size_t)(my >> 3) * S.wpr + wx) * 8u + (my & 7u)] = make_uint2(tword, word);  // tm_index layout
      words[k] = word;
      n += __popc(word);
      up = mid;
      mid = dn;
    }
  }
  // append the set cells to the list: one atomicAdd per warp
  const uint32_t lane = threadIdx.x & 31u;
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
#pragma unroll
    for (int k = 0; k < MASK_ROWS; ++k) {
      uint32_t word = words[k];
      while (word) {
        const int i = __ffs(word) - 1;
        word &= word - 1;
        const uint32_t cell = (y0 + k) * W + mx0 + i;
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
}

// ------------------------------------------------------------------------------------------------ K2: union-find
// 8-connected components of the frontier mask, over mask-list indices (lock-free min-root union).
__device__ __forceinline__ uint32_t uf_find(uint32_t *parent, uint32_t i) {
  uint32_t p = ld_cg(parent + i);
  while (p != i) {
    const uint32_t gp = ld_cg(parent + p);
    if (gp != p) parent[i] = gp;  // path halving; racing writers only ever store an ancestor
    i = p;
    p = gp;
  }
  return i;
}
__device__ __forceinline__ void uf_union(uint32_t *parent, uint32_t a, uint32_t b) {
  bool done;
  do {
    a = uf_find(parent, a);
    b = uf_find(parent, b);
    if (a < b) {
      const uint32_t old = atomicMin(parent + b, a);
      done = (old == b);
      b = old;
    } else if (b < a) {
      const uint32_t old = atomicMin(parent + a, b);
      done = (old == a);
      a = old;
    } else {
      done = true;
    }
  } while (!done);
}
__global__ void uf_init_kernel(uint32_t *parent, uint32_t *rkey, uint32_t *rsize, uint32_t *rbox, const Counters *cnt, uint32_t cap) {
  const uint32_t n = min(cnt->n_mask, cap);
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    parent[i] = i;
    rkey[i] = KEY_INF;
    rsize[i] = 0;
    reinterpret_cast<uint4 *>(rbox)[i] = make_uint4(0xFFFFFFFFu, 0xFFFFFFFFu, 0u, 0u);
  }
}
// each mask cell links to its mask neighbours in the forward half of the 8-neighbourhood
__global__ void uf_merge_kernel(const SearchParams S, const uint32_t *__restrict__ bits, const uint32_t *__restrict__ slot,
                                const uint32_t *__restrict__ mcell, uint32_t *parent, const Counters *cnt) {
  if (cnt->overflow) return;  // slot[] is incomplete when the mask list overflowed; the host grows the list and reruns
  const uint32_t n = min(cnt->n_mask, S.mask_cap);
  const uint32_t W = S.map.W, H = S.map.H;
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const uint32_t c = mcell[i];
    const uint32_t my = c / W, mx = c - my * W;
    if (mx + 1 < W && get_bit(bits, S.wpr, mx + 1, my)) uf_union(parent, i, slot[c + 1]);
    if (my + 1 < H) {
      if (mx > 0 && get_bit(bits, S.wpr, mx - 1, my + 1)) uf_union(parent, i, slot[c + W - 1]);
      if (get_bit(bits, S.wpr, mx, my + 1)) uf_union(parent, i, slot[c + W]);
      if (mx + 1 < W && get_bit(bits, S.wpr, mx + 1, my + 1)) uf_union(parent, i, slot[c + W + 1]);
    }
  }
}
// flatten: parent[i] = root; rsize[root] = number of cells.  The walk is read-only: a path-halving store from a
// thread passing through i could land after i's own "parent[i] = root" and leave a non-root there.
__global__ void uf_compress_kernel(uint32_t *parent, uint32_t *rsize, uint32_t *rbox, const uint32_t *__restrict__ mcell, uint32_t W,
                                   const Counters *cnt, uint32_t cap) {
  const uint32_t n = min(cnt->n_mask, cap);
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    uint32_t r = i, p = ld_cg(parent + r);
    while (p != r) {  // every value ever stored in parent[x] is an ancestor of x, so this ends at the root
      r = p;
      p = ld_cg(parent + r);
    }
    parent[i] = r;
    atomicAdd(rsize + r, 1u);
    const uint32_t c = mcell[i], my = c / W, mx = c - my * W;
    atomicMin(rbox + 4 * r, mx);
    atomicMin(rbox + 4 * r + 1, my);
    atomicMax(rbox + 4 * r + 2, mx);
    atomicMax(rbox + 4 * r + 3, my);
  }
}

// ------------------------------------------------------------------------------------------------ K3: ranked wavefronts
// Block-wide exclusive scan of small per-thread counts; returns this thread's offset, *total = sum.
// warp_sums needs 33 words.  Ends with a barrier, so shared scratch can be reused right away.
__device__ __forceinline__ uint32_t block_scan(uint32_t v, uint32_t *warp_sums, uint32_t *total) {
  const uint32_t lane = threadIdx.x & 31u, wid = threadIdx.x >> 5, nw = blockDim.x >> 5;
  uint32_t incl = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
    if ((int)lane >= o) incl += t;
  }
  if (lane == 31) warp_sums[wid] = incl;
  __syncthreads();
  if (wid == 0) {
    const uint32_t w = lane < nw ? warp_sums[lane] : 0;
    uint32_t wi = w;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t t = __shfl_up_sync(0xffffffffu, wi, o);
      if ((int)lane >= o) wi += t;
    }
    warp_sums[lane] = wi - w;
    if (lane == 31) warp_sums[32] = wi;
  }
  __syncthreads();
  const uint32_t r = warp_sums[wid] + incl - v;
  *total = warp_sums[32];
  __syncthreads();
  return r;
}

// nearestFreeCell (DEP/src/Helpers.cpp:285-330): 8-connected FIFO wavefront over every cell, stops at the first
// dequeued cell with cost < lethal_threshold.  Rare path (the robot normally stands on free space), so it uses the
// plain claim-with-atomicMin formulation: key = level base + queue position * 8 + slot.  Leaves claim[] clean.
__device__ void nearest_free_cell(const SearchParams &S, uint32_t *claim, uint32_t *q0, uint32_t *q1, uint32_t start,
                                  uint32_t *sh, Counters *cnt, uint32_t *found_cell) {
  const uint32_t W = S.map.W, H = S.map.H;
  uint32_t *warp_sums = sh;   // 33 words
  uint32_t *s_min = sh + 40;
  uint32_t n = 1, base = 1, levels = 0;
  uint32_t *q = q0, *qn = q1;
  if (threadIdx.x == 0) {
    q[0] = start;
    claim[start] = 0;
  }
  __syncthreads();
  while (n > 0) {
    if (threadIdx.x == 0) *s_min = KEY_INF;
    __syncthreads();
    for (uint32_t i = threadIdx.x; i < n; i += blockDim.x) {
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
    const uint32_t total = n << 3;
    for (uint32_t j = threadIdx.x; j < total; j += blockDim.x) {
      const uint32_t c = q[j >> 3];
      const uint32_t my = c / W, mx = c - my * W;
      uint32_t nx, ny;
      if (!nbr_of(mx, my, W, H, (int)(j & 7), nx, ny)) continue;
      const uint32_t nb = ny * W + nx;
      if (ld_cg(claim + nb) >= base) atomicMin(claim + nb, base + j);
    }
    __threadfence();
    __syncthreads();
    uint32_t n_next = 0;
    for (uint32_t j0 = 0; j0 < total; j0 += blockDim.x) {
      const uint32_t j = j0 + threadIdx.x;
      bool push = false;
      uint32_t nb = 0;
      if (j < total) {
        const uint32_t c = q[j >> 3];
        const uint32_t my = c / W, mx = c - my * W;
        uint32_t nx, ny;
        if (nbr_of(mx, my, W, H, (int)(j & 7), nx, ny)) {
          nb = ny * W + nx;
          push = ld_cg(claim + nb) == base + j;
        }
      }
      uint32_t tile_total;
      const uint32_t off = block_scan(push ? 1u : 0u, warp_sums, &tile_total);
      if (push) {
        if (n_next + off < S.queue_cap / 2) qn[n_next + off] = nb;
        else atomicOr(&cnt->overflow, 2u);
      }
      n_next += tile_total;
    }
    __threadfence();
    __syncthreads();
    base += total;
    n = min(n_next, S.queue_cap / 2);
    uint32_t *t = q; q = qn; qn = t;
    ++levels;
  }
  // undo the claims (Chebyshev ball of radius levels + 1 around the start)
  __syncthreads();
  const uint32_t sy = start / W, sx = start - sy * W;
  const int r = (int)levels + 1;
  const long side = 2 * r + 1;
  for (long i = threadIdx.x; i < side * side; i += blockDim.x) {
    const int y = (int)sy - r + (int)(i / side), x = (int)sx - r + (int)(i % side);
    if (x >= 0 && y >= 0 && x < (int)W && y < (int)H) claim[(size_t)y * W + x] = KEY_INF;
  }
  __threadfence();
  __syncthreads();
}

// The outer wavefront of searchFrom (DEP/src/FrontierSearch.cpp:58-94), one CTA, levels in exact FIFO order.
// A dequeued cell p with rank r looks at its 4 neighbours in the reference's order.  A traversable, unvisited neighbour c
// is visited by p iff no other neighbour of c has a rank in [level base, r): those are exactly the cells dequeued
// before p in this level that would have visited c first.  Winners are appended in (rank, slot) order by a block scan,
// which is the order the reference pushes them.  Frontier-mask neighbours only record the smallest touching key.
__global__ void __launch_bounds__(BFS_THREADS)
rank_bfs_kernel(const SearchParams S, const uint32_t *__restrict__ bits, const uint32_t *__restrict__ trav, uint32_t *claim,
                uint32_t *queue, Counters *cnt) {
  __shared__ uint32_t sh[48];
  __shared__ uint32_t s_start;
  if (cnt->overflow) return;
  if (S.use_smem && cnt->ow_fallback != 1u) return;  // a shared-memory wavefront finished the search
  const uint32_t W = S.map.W, H = S.map.H;
  const uint32_t ry = S.robot_cell / W, rx = S.robot_cell - ry * W;
  if (threadIdx.x == 0) s_start = S.robot_cell;  // falls back to pos when no cell qualifies (FrontierSearch.cpp:47-54)
  __syncthreads();
  if (S.map.data[(size_t)ry * S.map.pitch + rx] >= S.lethal) {
    nearest_free_cell(S, claim, queue, queue + S.queue_cap / 2, S.robot_cell, sh, cnt, &s_start);
    __syncthreads();
  }
  // The current level's cells are mirrored in shared memory (first SQ entries), so that the common narrow level costs
  // two global round trips (neighbour bits + claims, then the claims of the neighbours' neighbours) instead of three.
  constexpr uint32_t SQ = 2048;
  __shared__ uint32_t sq[2][SQ];
  uint32_t base = 0, n = 1, levels = 0, cur = 0;
  {
    const uint32_t start = s_start;
    if (threadIdx.x == 0) {
      cnt->start_cell = start;
      queue[0] = start;
      claim[start] = 0;
      sq[0][0] = start;
    }
  }
  __syncthreads();
  const int dxs[4] = {-1, 1, 0, 0}, dys[4] = {0, 0, -1, 1};
  while (n > 0) {
    const uint32_t next_base = base + n;
    uint32_t n_next = 0;
    for (uint32_t tile = 0; tile < n; tile += BFS_THREADS) {
      const uint32_t i = tile + threadIdx.x;
      uint32_t win[4];
      uint32_t n_win = 0, far_mask = 0;
      uint32_t c = 0, mx = 0, my = 0;
      if (i < n) {
        const uint32_t rank = base + i;
        c = i < SQ ? sq[cur][i] : ld_cg(queue + rank);
        my = c / W;
        mx = c - my * W;
        // ---- batch 1: the four neighbours' class bits and claims, all loads independent
        uint32_t nb[4], cl[4];
        bool ok[4], tr[4], mk[4];
#pragma unroll
        for (int s = 0; s < 4; ++s) {
          const int x = (int)mx + dxs[s], y = (int)my + dys[s];
          ok[s] = x >= 0 && y >= 0 && x < (int)W && y < (int)H;
          nb[s] = ok[s] ? (uint32_t)y * W + (uint32_t)x : c;
        }
#pragma unroll
        for (int s = 0; s < 4; ++s) {
          const uint32_t ny = nb[s] / W, nx = nb[s] - ny * W;
          tr[s] = get_bit(trav, S.wpr, nx, ny);
          mk[s] = get_bit(bits, S.wpr, nx, ny);
          cl[s] = ld_cg(claim + nb[s]);
        }
        // ---- batch 2: ranks of each candidate's other neighbours (own cell included: its rank is never < rank)
        uint32_t gr[4][4];
#pragma unroll
        for (int s = 0; s < 4; ++s) {
          const bool cand = ok[s] && tr[s] && cl[s] == KEY_INF;
          const uint32_t ny = nb[s] / W, nx = nb[s] - ny * W;
#pragma unroll
          for (int t = 0; t < 4; ++t) {
            const int x = (int)nx + dxs[t], y = (int)ny + dys[t];
            const bool in = cand && x >= 0 && y >= 0 && x < (int)W && y < (int)H;
            gr[s][t] = in ? ld_cg(claim + (uint32_t)y * W + (uint32_t)x) : KEY_INF;
          }
        }
#pragma unroll
        for (int s = 0; s < 4; ++s) {
          if (!ok[s]) continue;
          if (tr[s]) {
            if (cl[s] != KEY_INF) continue;  // visited_flag[nbr]
            bool first = true;
#pragma unroll
            for (int t = 0; t < 4; ++t)
              if (gr[s][t] >= base && gr[s][t] < rank) first = false;
            if (!first) continue;
            // visited now; enqueued only inside the search radius (FrontierSearch.cpp:72-76)
            const uint32_t ny = nb[s] / W, nx = nb[s] - ny * W;
            const double wx = S.map.ox + (nx + 0.5) * S.map.res;
            const double wy = S.map.oy + (ny + 0.5) * S.map.res;
            const double ddx = S.robot_x - wx, ddy = S.robot_y - wy;
            if (sqrt(ddx * ddx + ddy * ddy) < S.dist_limit) win[n_win++] = nb[s];
            else far_mask |= 1u << s;
          } else if (mk[s]) {
            atomicMin(claim + nb[s], TOUCH_BIT | (rank * 4u + (uint32_t)s));  // isNewFrontierCell branch: first touch
          }
        }
      }
      uint32_t tile_total;
      const uint32_t off = block_scan(n_win, sh, &tile_total);
      const uint32_t pos = n_next + off;
      for (uint32_t k = 0; k < n_win; ++k) {
        queue[next_base + pos + k] = win[k];
        if (pos + k < SQ) sq[cur ^ 1][pos + k] = win[k];
        claim[win[k]] = next_base + pos + k;
      }
      if (far_mask) {
        for (int s = 0; s < 4; ++s)
          if (far_mask & (1u << s)) claim[(uint32_t)((int)my + dys[s]) * W + (uint32_t)((int)mx + dxs[s])] = KEY_FAR;
      }
      n_next += tile_total;
      __syncthreads();
    }
    base = next_base;
    n = n_next;
    cur ^= 1;
    ++levels;
  }
  if (threadIdx.x == 0) {
    cnt->bfs_levels = levels;
    cnt->expanded = base;
    cnt->dirty_queue = base;
  }
}

// ---- shared-memory wavefronts ------------------------------------------------------------------------------------
// Multiplicative hash of a cell index into a table of 2^bits slots.
__device__ __forceinline__ uint32_t hslot(uint32_t cell, uint32_t bits) { return (cell * 0x9E3779B1u) >> (32u - bits); }

// class words {traversable, frontier mask} of 32 cells: rows are interleaved in groups of 8 so that the three rows a
// cell needs share a 64-byte chunk, and the ring of a wavefront level stays inside L1 from one level to the next
__host__ __device__ __forceinline__ size_t tm_index(uint32_t wx, uint32_t y, uint32_t wpr) {
  return ((size_t)(y >> 3) * wpr + wx) * 8u + (y & 7u);
}
// ---- the outer wavefront, default path ---------------------------------------------------------------------------
// Same decision rule as rank_bfs_kernel above (a dequeued cell p visits a traversable, unvisited neighbour c iff no
// other neighbour of c was dequeued before p in this level), with everything that does not depend on the order moved
// out of the loop, and the loop cut down to what the profile says bounds it: instructions per level (one SM issues
// about two warp instructions per cycle on this integer code) and distinct L1 lines per load (profiles/README.md).
//   * the state is 2 bytes per cell: ow_prep_kernel (all SMs, once per search) writes the cell's class, the wavefront
//     replaces it with the cell's position inside its level when the cell is enqueued.
//       OW_OPEN    expandable, not visited: inside the map, cost < 254, inside the search radius (FrontierSearch.cpp:67,74)
//       OW_MASK    frontier-mask cell (isNewFrontierCell, the `else if` branch)
//       OW_BLOCKED anything else
//       < OW_MAX_LEVEL: position in its level
//     The loop needs no map bytes, no bounds tests, no radius arithmetic and no visited flags: a dequeued cell reads its
//     four neighbours; OPEN ones are its candidates, MASK ones get its touch key.
//   * a cell c that is still OPEN while level d is expanded has no neighbour on a level below d (it would have visited
//     c), and the 4-connected grid is bipartite, so any neighbour of c with a position IS a level-d cell: no level
//     stamps.  A thread reads the positions of the 8 cells at distance 2 (those next to a candidate; each diagonal serves
//     two candidates) and its cell visits c iff none of c's three other neighbours has a smaller position.
//   * the array is tiled (8 x 8 cells = one 128-byte line) and padded by OW_PAD cells of BLOCKED; a level's cells lie on
//     diagonals, so a warp's load touches ~5 lines instead of 32.  The index separates into f(x) + g(y): a queue entry
//     carries the cell's index, and a step to a neighbour is an add plus a select for the tile edge.
// One CTA: every position it reads was stored by the same CTA before an earlier barrier, so plain L1-cached loads are
// coherent.  Two barriers per level (one more per 1024 positions beyond the first).
static constexpr uint32_t OW_PAD = 8;
static constexpr uint32_t OW_OPEN = 0xFFFFu, OW_MASK = 0xFFFEu, OW_BLOCKED = 0xFFFDu;
static constexpr uint32_t OW_MAX_LEVEL = 0xFF00u;  // widest level whose positions stay below the class values

struct OwLayout {
  uint32_t txr;            // tiles (8 x 8 cells) per tile row
  uint32_t rk_yt;          // txr * 64 - 64: see ow_rk_fy
  uint32_t Wp, Hp;         // padded size in cells
};
// idx = fx(X) + fy(Y) with X = x + OW_PAD, Y = y + OW_PAD
__host__ __device__ __forceinline__ uint32_t ow_rk_fx(uint32_t X) { return X + (X >> 3) * 56u; }                   // (X >> 3) * 64 + (X & 7)
__host__ __device__ __forceinline__ uint32_t ow_rk_fy(uint32_t Y, uint32_t yt) { return Y * 8u + (Y >> 3) * yt; }    // (Y >> 3) * txr * 64 + (Y & 7) * 8
static OwLayout ow_layout(uint32_t W, uint32_t H) {
  OwLayout L;
  L.Wp = (W + 2u * OW_PAD + 15u) / 16u * 16u;
  L.Hp = (H + 2u * OW_PAD + 7u) / 8u * 8u;
  L.txr = L.Wp / 8u;
  L.rk_yt = L.txr * 64u - 64u;
  return L;
}

// 32 bits of row y of a 1-bit-per-cell array, columns xs .. xs + 31 (xs >= -32); cells outside the map read 0
__device__ __forceinline__ uint32_t row_bits(const uint32_t *__restrict__ a, uint32_t wpr, uint32_t H, int y, int xs) {
  if (y < 0 || y >= (int)H) return 0u;
  const int w0 = xs >> 5;
  const uint32_t *row = a + (size_t)y * wpr;
  const uint32_t lo = (w0 >= 0 && w0 < (int)wpr) ? __ldg(row + w0) : 0u;
  const uint32_t hi = (w0 + 1 >= 0 && w0 + 1 < (int)wpr) ? __ldg(row + w0 + 1) : 0u;
  return __funnelshift_r(lo, hi, (uint32_t)xs & 31u);
}

// Classes for the tiles of the search disc's bounding box (pairs of 8 x 8-cell tiles [2 * tx0, 2 * (tx0 + ntx)) x
// [ty0, ty0 + nty), padded coordinates).  One thread per 16 cells of a row: two 16-byte stores.  The wavefront never
// leaves the box: cells are enqueued only inside the disc.
__global__ void __launch_bounds__(256)
ow_prep_kernel(const SearchParams S, const OwLayout L, const uint32_t *__restrict__ ebits, const uint32_t *__restrict__ mbits,
               uint16_t *__restrict__ rk, uint32_t tx0, uint32_t ty0, uint32_t ntx, uint32_t nty) {
  const uint64_t n = (uint64_t)ntx * nty * 8u;
  const uint32_t H = S.map.H;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
    const uint32_t r = (uint32_t)(i / ntx), bx = tx0 + (uint32_t)(i - (uint64_t)r * ntx);
    const uint32_t Y = ty0 * 8u + r, X0 = bx * 16u;
    const int y = (int)Y - (int)OW_PAD, xs = (int)X0 - (int)OW_PAD;  // bit j of the row words below = column xs + j
    const uint32_t em = row_bits(ebits, S.wpr, H, y, xs), mm = row_bits(mbits, S.wpr, H, y, xs);  // (0 outside the map)
    uint32_t out[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const uint32_t a = ((em >> (2 * k)) & 1u) ? OW_OPEN : (((mm >> (2 * k)) & 1u) ? OW_MASK : OW_BLOCKED);
      const uint32_t b = ((em >> (2 * k + 1)) & 1u) ? OW_OPEN : (((mm >> (2 * k + 1)) & 1u) ? OW_MASK : OW_BLOCKED);
      out[k] = a | (b << 16);
    }
    uint16_t *t = rk + ((size_t)(Y >> 3) * L.txr + 2u * bx) * 64u + (Y & 7u) * 8u;  // first of the two tiles' rows
    *reinterpret_cast<uint4 *>(t) = make_uint4(out[0], out[1], out[2], out[3]);
    *reinterpret_cast<uint4 *>(t + 64) = make_uint4(out[4], out[5], out[6], out[7]);
  }
}

struct OwQueues {
  // entries: .x = x | y << 14 | candidate directions << 28 (the OPEN neighbours), .y = directions of the MASK neighbours.
  // Whoever appends a cell has read the cell's other three neighbours to decide (they are cells at distance 2 from the
  // appender), so the classes of all four travel with the entry and a dequeued cell never reads its own neighbours.
  uint2 *q;        // shared memory: two queues of OW_QCAP entries, this level's (`cur`) and the next one's
  uint2 *g;        // the same in global memory (two of `gcap` entries), used while a level is wider than the shared-memory queue
  uint32_t gcap;
  bool from_global, to_global;
};

__device__ __forceinline__ void ow_bar() { asm volatile("bar.sync 0;" ::: "memory"); }
#ifdef FITSLAM_OW_PROF
// cycle stamps of thread 0 (warp 0 works in every pass); `dep` orders the stamp behind the value it times
__device__ __forceinline__ uint32_t ow_clock(uint32_t dep) { uint32_t t; asm volatile("mov.u32 %0, %%clock;" : "=r"(t) : "r"(dep) : "memory"); return t; }
#define OW_STAMP(k, dep) do { if (threadIdx.x == 0) { const uint32_t _t = ow_clock(dep); pf.acc[k] += _t - pf.last; pf.last = _t; } } while (0)
#else
#define OW_STAMP(k, dep) do { } while (0)
#endif
struct OwProf { uint32_t last; uint32_t acc[8]; };

// The decision of one dequeued cell (both wavefront kernels).  In: its queue entry, its position `pos` in the level, the
// state array through `ld` (plain loads in the single-CTA kernel, ld.global.cg in the cluster kernel).  Out: which
// neighbours it visits (`win`), and for each of them the state index (`ci`) and the queue entry (`ent`).
struct OwCell {
  uint32_t win;
  uint32_t ci[4];
  uint2 ent[4];
};
template <typename Load>
__device__ __forceinline__ void ow_decide(const OwLayout &L, const uint2 e, uint32_t pos, Load ld, OwCell &out) {
  const uint32_t cand = e.x >> 28, xy = e.x & 0x0FFFFFFFu;
  const uint32_t i = ow_rk_fx((xy & 0x3FFFu) + OW_PAD) + ow_rk_fy((xy >> 14) + OW_PAD, L.rk_yt);
  // index steps to the cells at distance 1 and 2: the index separates into f(x) + g(y), so a step along x costs a fixed
  // amount plus 56 when it crosses a tile edge, a step along y 8 plus (tile row - 64) when it does
  const uint32_t tx = i & 7u, ty = i & 56u;
  const uint32_t xa = tx < 2u ? 58u : 2u, xb = tx < 1u ? 57u : 1u, xd = tx > 6u ? 57u : 1u, xe = tx > 5u ? 58u : 2u;
  const uint32_t ya = (ty < 16u ? L.rk_yt : 0u) + 16u, yb = (ty < 8u ? L.rk_yt : 0u) + 8u, yd = (ty > 48u ? L.rk_yt : 0u) + 8u,
                 ye = (ty > 40u ? L.rk_yt : 0u) + 16u;
  const uint32_t i0 = i - xb, i1 = i + xd, i2 = i - yb, i3 = i + yd;  // the four neighbours: -x, +x, -y, +y
  // state of the cells at distance 2 that are next to a candidate (all loads independent): a position if the cell is on
  // this level, else its class, which is above every position
  const uint32_t l2 = (cand & 1u) ? ld(i - xa) : OW_BLOCKED, r2 = (cand & 2u) ? ld(i + xe) : OW_BLOCKED;
  const uint32_t u2 = (cand & 4u) ? ld(i - ya) : OW_BLOCKED, d2 = (cand & 8u) ? ld(i + ye) : OW_BLOCKED;
  const uint32_t ul = (cand & 5u) ? ld(i0 - yb) : OW_BLOCKED, ur = (cand & 6u) ? ld(i1 - yb) : OW_BLOCKED;
  const uint32_t dl = (cand & 9u) ? ld(i0 + yd) : OW_BLOCKED, dr = (cand & 10u) ? ld(i1 + yd) : OW_BLOCKED;
  // a level-d neighbour of the candidate that was dequeued before this cell visited it first
  uint32_t win = 0u;
  if ((cand & 1u) && min(l2, min(ul, dl)) > pos) win |= 1u;
  if ((cand & 2u) && min(r2, min(ur, dr)) > pos) win |= 2u;
  if ((cand & 4u) && min(u2, min(ul, ur)) > pos) win |= 4u;
  if ((cand & 8u) && min(d2, min(dl, dr)) > pos) win |= 8u;
  out.win = win;
  out.ci[0] = i0; out.ci[1] = i1; out.ci[2] = i2; out.ci[3] = i3;
  // entries of the visited cells: their OPEN neighbours are their candidates, their MASK neighbours get their touch key
  // (the fourth neighbour is this cell)
  const uint32_t ol2 = l2 == OW_OPEN, or2 = r2 == OW_OPEN, ou2 = u2 == OW_OPEN, od2 = d2 == OW_OPEN;
  const uint32_t oul = ul == OW_OPEN, our = ur == OW_OPEN, odl = dl == OW_OPEN, odr = dr == OW_OPEN;
  const uint32_t ml2 = l2 == OW_MASK, mr2 = r2 == OW_MASK, mu2 = u2 == OW_MASK, md2 = d2 == OW_MASK;
  const uint32_t mul = ul == OW_MASK, mur = ur == OW_MASK, mdl = dl == OW_MASK, mdr = dr == OW_MASK;
  out.ent[0] = make_uint2((xy - 1u) | ((ol2 | (oul << 2) | (odl << 3)) << 28), ml2 | (mul << 2) | (mdl << 3));
  out.ent[1] = make_uint2((xy + 1u) | (((or2 << 1) | (our << 2) | (odr << 3)) << 28), (mr2 << 1) | (mur << 2) | (mdr << 3));
  out.ent[2] = make_uint2((xy - (1u << 14)) | ((oul | (our << 1) | (ou2 << 2)) << 28), mul | (mur << 1) | (mu2 << 2));
  out.ent[3] = make_uint2((xy + (1u << 14)) | ((odl | (odr << 1) | (od2 << 3)) << 28), mdl | (mdr << 1) | (md2 << 3));
}
// frontier-mask neighbours record the first touch (isNewFrontierCell branch, FrontierSearch.cpp:78-81)
__device__ __forceinline__ void ow_touch(uint32_t *claim, uint32_t W, const uint2 e, uint32_t rank) {
  if (!e.y) return;
  const uint32_t key = TOUCH_BIT | (rank << 2);
  uint32_t *at = claim + (size_t)((e.x >> 14) & 0x3FFFu) * W + (e.x & 0x3FFFu);
  if (e.y & 1u) atomicMin(at - 1, key);
  if (e.y & 2u) atomicMin(at + 1, key | 1u);
  if (e.y & 4u) atomicMin(at - W, key | 2u);
  if (e.y & 8u) atomicMin(at + W, key | 3u);
}
// entry of the start cell: the classes of its four neighbours
__device__ __forceinline__ uint2 ow_start_entry(const OwLayout &L, const uint16_t *rk, uint32_t sx, uint32_t sy, uint32_t *si) {
  const uint32_t i = ow_rk_fx(sx + OW_PAD) + ow_rk_fy(sy + OW_PAD, L.rk_yt);
  const uint32_t v[4] = {rk[ow_rk_fx(sx + OW_PAD - 1u) + ow_rk_fy(sy + OW_PAD, L.rk_yt)], rk[ow_rk_fx(sx + OW_PAD + 1u) + ow_rk_fy(sy + OW_PAD, L.rk_yt)],
                         rk[ow_rk_fx(sx + OW_PAD) + ow_rk_fy(sy + OW_PAD - 1u, L.rk_yt)], rk[ow_rk_fx(sx + OW_PAD) + ow_rk_fy(sy + OW_PAD + 1u, L.rk_yt)]};
  uint32_t cand = 0u, mk = 0u;
  for (int k = 0; k < 4; ++k) { cand |= (v[k] == OW_OPEN ? 1u : 0u) << k; mk |= (v[k] == OW_MASK ? 1u : 0u) << k; }
  *si = i;
  return make_uint2(sx | (sy << 14) | (cand << 28), mk);
}

// one pass over up to OW_THREADS queue positions starting at `tile`; returns the number of cells it appended (valid in
// the warps that had positions to work on: warp 0 always has).  Warps whose positions all lie beyond the level only
// take part in the barrier: with ~400 cells per level two thirds of the warps are idle, and the instructions they would
// spend on an empty scan come out of the issue slots of the busy ones.
// WIDE: the level or its successor may not fit the shared-memory queue (a separate instantiation keeps the common
// narrow pass free of those branches).
template <bool WIDE>
__device__ __forceinline__ uint32_t ow_pass(const SearchParams &S, const OwLayout &L, uint16_t *rk, uint32_t *claim, const OwQueues &Q,
                                           uint32_t cur, uint32_t tile, uint32_t n, uint32_t base, uint32_t n_next,
                                           uint32_t *ws, OwProf &pf) {
  const uint32_t lane = threadIdx.x & 31u, wid = threadIdx.x >> 5;
  if (tile + (wid << 5) >= n) {  // warp-uniform
    ow_bar();
    return 0u;
  }
  const uint32_t pos = tile + threadIdx.x;
  uint2 e = make_uint2(0u, 0u);  // lanes beyond the level: no candidates, no mask neighbours
  if (pos < n) e = (WIDE && Q.from_global) ? __ldcg(Q.g + cur * Q.gcap + pos) : Q.q[cur * OW_QCAP + pos];
  OW_STAMP(0, e.x);
  OwCell d;
  ow_decide(L, e, pos, [&](uint32_t at) { return (uint32_t)rk[at]; }, d);
  ow_touch(claim, S.map.W, e, base + pos);
  OW_STAMP(2, d.win);
  // block-wide exclusive scan of the winner counts (0..4 per thread): three ballots per warp, one barrier
  const uint32_t nw = __popc(d.win);
  const uint32_t b0 = __ballot_sync(0xffffffffu, nw & 1u), b1 = __ballot_sync(0xffffffffu, nw & 2u), b2 = __ballot_sync(0xffffffffu, nw & 4u);
  const uint32_t lt = (1u << lane) - 1u;
  const uint32_t excl = __popc(b0 & lt) + 2u * __popc(b1 & lt) + 4u * __popc(b2 & lt);
  if (lane == 0) ws[wid] = __popc(b0) + 2u * __popc(b1) + 4u * __popc(b2);
  OW_STAMP(3, excl);
  ow_bar();
  OW_STAMP(4, excl);
  const uint32_t aw = min((n - tile + 31u) >> 5, (uint32_t)(OW_THREADS / 32));  // warps with positions in this pass
  const uint32_t wv = lane < aw ? ws[lane] : 0u;
  const uint32_t total = __reduce_add_sync(0xffffffffu, wv);
  const uint32_t wbase = __reduce_add_sync(0xffffffffu, lane < wid ? wv : 0u);
  uint32_t dst = n_next + wbase + excl;
  // append the winners in (position, slot) order
#pragma unroll
  for (int k = 0; k < 4; ++k)
    if (d.win & (1u << k)) {
      rk[d.ci[k]] = (uint16_t)dst;
      if (dst < S.ow_qcap) Q.q[(cur ^ 1u) * OW_QCAP + dst] = d.ent[k];
      if (WIDE && Q.to_global) Q.g[(cur ^ 1u) * Q.gcap + dst] = d.ent[k];
      ++dst;
    }
  OW_STAMP(5, dst);
  return total;
}

// WIDE = false: every level fits the shared-memory queue; a level that does not ends the kernel with ow_fallback = 2 (the
// host repeats the search with the WIDE = true instantiation, which spills such levels to a global-memory queue, and
// starts with that one from then on while the map keeps producing them) or 1 (maps above 16384 cells a side or levels
// above OW_MAX_LEVEL cells: rank_bfs_kernel).
template <bool WIDE>
__global__ void __launch_bounds__(OW_THREADS, 1)
ow_kernel(const SearchParams S, const OwLayout L, uint16_t *rk, uint32_t *claim, uint32_t *queue, Counters *cnt, uint32_t n_parts) {
  extern __shared__ uint2 ow_sh[];
  __shared__ uint32_t sh48[48];
  __shared__ uint32_t s_start, s_n;
  __shared__ uint32_t s_wsum[64];
  const long long k0 = clock64();
  OwQueues Q;
  Q.q = ow_sh;
  Q.g = reinterpret_cast<uint2 *>(queue);
  Q.gcap = S.queue_cap / 4;
  Q.from_global = false;
  Q.to_global = false;
  // parts [from, n_parts - 1) are complete with `end` cells dequeued: every touch key below 4 * end is in memory (the
  // atomics of all threads precede the barrier before this call; the fence orders them before the flags)
  auto publish = [&](uint32_t from, uint32_t end) {
    if (threadIdx.x != 0) return;
    for (uint32_t k = from; k + 1 < n_parts; ++k) cnt->part_end[k] = end;
    __threadfence();
    for (uint32_t k = from; k + 1 < n_parts; ++k) atomicExch(&cnt->part_flag[k], 1u);
  };
  if (cnt->overflow) { publish(0, 0); return; }
  const uint32_t W = S.map.W;
  const uint32_t ry = S.robot_cell / W, rx = S.robot_cell - ry * W;
  uint32_t base = 0, n = 1, levels = 0, scans = 0, part = 0, widest = 1, cur = 0;
  if (threadIdx.x == 0) s_start = S.robot_cell;
  __syncthreads();
  if (S.map.data[(size_t)ry * S.map.pitch + rx] >= S.lethal) {
    nearest_free_cell(S, claim, queue, queue + S.queue_cap / 2, S.robot_cell, sh48, cnt, &s_start);
    __syncthreads();
  }
  const uint32_t start = s_start;
  if (threadIdx.x == 0) {
    cnt->start_cell = start;
    const uint32_t sy = start / W, sx = start - sy * W;
    uint32_t si;
    Q.q[0] = ow_start_entry(L, rk, sx, sy, &si);
    rk[si] = 0;  // visited_flag[bfs.front()] = true (FrontierSearch.cpp:56), whatever its class
  }
  __syncthreads();
  uint32_t next_part_level = S.ow_part_end[0];
  OwProf pf;
#ifdef FITSLAM_OW_PROF
  for (int k = 0; k < 8; ++k) pf.acc[k] = 0;
  pf.last = ow_clock(0);
#endif
  while (n > 0) {
    OW_STAMP(7, n);
    uint32_t n_next = 0;
    for (uint32_t tile = 0; tile < n; tile += OW_THREADS) {
      n_next += ow_pass<WIDE>(S, L, rk, claim, Q, cur, tile, n, base, n_next, s_wsum + (scans & 1u) * 32u, pf);
      ++scans;
    }
    if (threadIdx.x == 0) s_n = n_next;  // warp 0 works in every pass: its total is the level's
    ow_bar();
    OW_STAMP(6, n_next);
    base += n;
    n = s_n;
    ++levels;
    cur ^= 1u;
    widest = max(widest, n);
    if ((n > S.ow_qcap && !WIDE) || n > OW_MAX_LEVEL) {  // uniform: the next kernel reruns the search (claim[] holds touch
      if (threadIdx.x == 0) cnt->ow_fallback = (S.ow_wide && n <= OW_MAX_LEVEL) ? 2u : 1u;  // keys only, and it would write the same ones)
      publish(part, base);  // every level below `base` ranks is complete: those touches stand
      return;
    }
    if (WIDE) {
      Q.from_global = n > S.ow_qcap;               // this level did not fit the shared queue: read it from global memory
      Q.to_global = 3u * n + 4u > S.ow_qcap;       // a level has at most 3 children per cell: the next one may not fit
    }
    if (levels >= next_part_level && part + 1 < n_parts) {
      if (threadIdx.x == 0) {
        cnt->part_end[part] = base;
        __threadfence();
        atomicExch(&cnt->part_flag[part], 1u);
      }
      ++part;
      next_part_level = S.ow_part_end[min(part, (uint32_t)OW_PARTS - 1u)];
    }
  }
  publish(part, base);
  if (threadIdx.x == 0) {
    cnt->bfs_levels = levels;
    cnt->expanded = base;
    cnt->dirty_queue = 0;  // claim[] was written at frontier-mask cells only
    cnt->widest_level = widest;
    cnt->prof[4] = scans;
    cnt->prof[5] = clock64() - k0;
#ifdef FITSLAM_OW_PROF
    for (int k = 0; k < 8; ++k) printf("ow_prof[%d] = %u (%.1f per level)\n", k, pf.acc[k], (double)pf.acc[k] / levels);
#endif
  }
}

// ---- the outer wavefront on a thread-block cluster ----------------------------------------------------------------
// The same search with a level spread over the CTAs of one cluster (16 SMs; 8 where the non-portable size is refused).
// The single-CTA kernel above is bound by what one SM's L1 pipeline can look up per level (~7 cycles per dequeued cell:
// 80 ms for C4's 22.8 M cells); here every CTA takes a contiguous slice of the level's positions.  ONE cluster barrier
// per level:
//   * CTA r keeps the cells it appends (its winners, in (position, slot) order) in its own shared memory, and the state
//     array gets r * OWC_SLICE + k for its k-th winner at once.  The CTAs' slices of a level are consecutive position
//     ranges, so these codes grow with the position in the next level: they order the cells exactly as positions do,
//     which is all the decision rule asks of them (ow_decide compares, never indexes), and nobody has to know the other
//     CTAs' counts to write them.  The store is safe while other CTAs are still deciding: a level-d decision reads cells
//     at distance 2 from level-d cells through an unvisited candidate, i.e. cells of level d or d + 2 (the grid is
//     bipartite), and the cells written during level d are the level-(d + 1) ones.
//   * every CTA tells every CTA its count through distributed shared memory, then barrier.cluster (release / acquire: the
//     codes in global memory, the winners' entries and the counts in shared memory).
//   * the next level is cut into equal position ranges again; a CTA PULLS the entries of its range out of the winner lists
//     of the CTAs that produced them (ld.shared::cluster), position p belonging to the producer whose prefix of counts
//     covers it.  Balance is restored every level without a second barrier.
//   * the state array goes through L2 (ld.global.cg / plain stores): L1 is not coherent across CTAs.
// Touch keys need the exact dequeue rank: base + p, known to the puller.
static constexpr int OWC_THREADS = 512;
static constexpr uint32_t OWC_SLICE = 4080;             // winners one CTA can append per level; 16 * 4080 = OW_MAX_LEVEL codes
static constexpr uint32_t OWC_SMEM = 2 * OWC_SLICE * 8;
static_assert(16u * OWC_SLICE <= OW_MAX_LEVEL, "codes stay below the class values");

__global__ void __launch_bounds__(OWC_THREADS, 1)
ow_cluster_kernel(const SearchParams S, const OwLayout L, uint16_t *rk, uint32_t *claim, uint32_t *queue, Counters *cnt, uint32_t n_parts) {
  namespace cg = cooperative_groups;
  cg::cluster_group cluster = cg::this_cluster();
  const uint32_t CS = cluster.num_blocks(), rank = cluster.block_rank();
  extern __shared__ uint2 owc_sh[];         // two winner lists of OWC_SLICE entries: this level's (read by every CTA) and the next one's
  __shared__ uint32_t s_cnt[2][16];         // winners per CTA (written by the CTAs themselves), by level parity
  __shared__ uint32_t s_pref[16];           // inclusive prefix of this level's counts
  __shared__ uint32_t s_ws[2][OWC_THREADS / 32];
  __shared__ uint32_t sh48[48];
  __shared__ uint32_t s_start;
  const uint32_t lane = threadIdx.x & 31u, wid = threadIdx.x >> 5;
  const uint32_t W = S.map.W;
  const long long k0 = clock64();
  auto publish = [&](uint32_t from, uint32_t end) {  // (CTA 0, after a cluster barrier: every CTA's touches precede it)
    if (rank != 0 || threadIdx.x != 0) return;
    for (uint32_t k = from; k + 1 < n_parts; ++k) cnt->part_end[k] = end;
    __threadfence();
    for (uint32_t k = from; k + 1 < n_parts; ++k) atomicExch(&cnt->part_flag[k], 1u);
  };
  if (cnt->overflow) { publish(0, 0); return; }  // (uniform: written before this kernel)
  if (threadIdx.x < 16) s_cnt[0][threadIdx.x] = threadIdx.x == 0 ? 1u : 0u;  // level 0: the start cell, CTA 0's first winner
  if (rank == 0) {
    const uint32_t ry = S.robot_cell / W, rx = S.robot_cell - ry * W;
    if (threadIdx.x == 0) s_start = S.robot_cell;
    __syncthreads();
    if (S.map.data[(size_t)ry * S.map.pitch + rx] >= S.lethal) {
      nearest_free_cell(S, claim, queue, queue + S.queue_cap / 2, S.robot_cell, sh48, cnt, &s_start);
      __syncthreads();
    }
    if (threadIdx.x == 0) {
      const uint32_t start = s_start;
      cnt->start_cell = start;
      const uint32_t sy = start / W, sx = start - sy * W;
      uint32_t si;
      owc_sh[0] = ow_start_entry(L, rk, sx, sy, &si);
      rk[si] = 0;  // visited_flag[bfs.front()] = true (FrontierSearch.cpp:56): code 0 = CTA 0's winner 0
    }
  }
  cluster.sync();
  uint32_t base = 0, levels = 0, part = 0, widest = 1, cur = 0, scans = 0;
  uint32_t next_part_level = S.ow_part_end[0];
  bool bad = false;
  OwProf pf;
#ifdef FITSLAM_OW_PROF
  for (int k = 0; k < 8; ++k) pf.acc[k] = 0;
  pf.last = ow_clock(0);
#endif
  for (;;) {
    // this level: counts per producer (every warp: one load, a shuffle scan and two reductions)
    const uint32_t cv = lane < CS ? s_cnt[levels & 1u][lane] : 0u;
    const bool any_bad = __reduce_max_sync(0xffffffffu, cv) == 0xFFFFFFFFu;
    uint32_t incl = cv;
#pragma unroll
    for (int o = 1; o < 16; o <<= 1) {
      const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
      if ((int)lane >= o) incl += t;
    }
    const uint32_t n = any_bad ? 0u : __shfl_sync(0xffffffffu, incl, 15);
    if (any_bad) {  // uniform across the cluster: a CTA appended more than its list holds; the global-memory wavefront reruns the
      if (rank == 0 && threadIdx.x == 0) cnt->ow_fallback = 1u;  // search (claim[] holds touch keys only, and it would write the same ones)
      publish(part, base);
      return;
    }
    if (n == 0) break;
    widest = max(widest, n);
    if (lane < 16) s_pref[lane] = incl;  // (every warp stores the same values)
    __syncwarp();
    OW_STAMP(0, n);
    const uint32_t per = (n + CS - 1u) / CS;
    const uint32_t lo = rank * per, m = lo < n ? min(per, n - lo) : 0u;  // this CTA's positions: [lo, lo + m)
    uint2 *lnext = owc_sh + (cur ^ 1u) * OWC_SLICE;
    uint32_t n_local = 0;
    for (uint32_t tile = 0; tile < m; tile += OWC_THREADS) {
      uint32_t *ws = s_ws[scans & 1u];
      ++scans;
      const uint32_t pl = tile + threadIdx.x;
      OwCell d;
      d.win = 0u;
      if (tile + (wid << 5) < m) {  // warp-uniform
        uint2 e = make_uint2(0u, 0u);
        uint32_t code = 0u;
        const uint32_t p = lo + pl;
        if (pl < m) {
          uint32_t prod = 0;
#pragma unroll
          for (int j = 0; j < 15; ++j) prod += (p >= s_pref[j]) ? 1u : 0u;  // the producer whose prefix of counts covers p
          const uint32_t k = p - (prod ? s_pref[prod - 1u] : 0u);
          e = cluster.map_shared_rank(owc_sh + cur * OWC_SLICE, prod)[k];
          code = prod * OWC_SLICE + k;
        }
        ow_decide(L, e, code, [&](uint32_t at) { return (uint32_t)__ldcg(rk + at); }, d);
        ow_touch(claim, W, e, base + p);
      }
      const uint32_t wmask = d.win;
      // CTA-wide exclusive scan of the winner counts (0..4 per thread): three ballots per warp, one barrier
      const uint32_t nw = __popc(wmask);
      const uint32_t b0 = __ballot_sync(0xffffffffu, nw & 1u), b1 = __ballot_sync(0xffffffffu, nw & 2u), b2 = __ballot_sync(0xffffffffu, nw & 4u);
      const uint32_t lt = (1u << lane) - 1u;
      const uint32_t excl = __popc(b0 & lt) + 2u * __popc(b1 & lt) + 4u * __popc(b2 & lt);
      if (lane == 0) ws[wid] = __popc(b0) + 2u * __popc(b1) + 4u * __popc(b2);
      __syncthreads();
      const uint32_t wv = lane < OWC_THREADS / 32 ? ws[lane] : 0u;
      const uint32_t total = __reduce_add_sync(0xffffffffu, wv);
      const uint32_t wbase = __reduce_add_sync(0xffffffffu, lane < wid ? wv : 0u);
      uint32_t dst = n_local + wbase + excl;
      if (n_local + total > min(OWC_SLICE, S.ow_qcap)) bad = true;  // (uniform in the CTA; FITSLAM_OW_QCAP lowers the limit in tests)
      else {
#pragma unroll
        for (int k = 0; k < 4; ++k)
          if (wmask & (1u << k)) {
            lnext[dst] = d.ent[k];
            rk[d.ci[k]] = (uint16_t)(rank * OWC_SLICE + dst);
            ++dst;
          }
      }
      n_local += total;
    }
    OW_STAMP(1, n_local);
    // every CTA tells every CTA how many cells it appended
    if (threadIdx.x < CS) cluster.map_shared_rank(&s_cnt[(levels + 1u) & 1u][0], threadIdx.x)[rank] = bad ? 0xFFFFFFFFu : n_local;
    OW_STAMP(2, n_local);
    cluster.sync();  // release / acquire: the codes (global memory), the winners' entries and the counts (shared memory)
    OW_STAMP(3, n_local);
    base += n;
    ++levels;
    cur ^= 1u;
    if (levels >= next_part_level && part + 1 < n_parts) {
      if (rank == 0 && threadIdx.x == 0) {
        cnt->part_end[part] = base;
        __threadfence();
        atomicExch(&cnt->part_flag[part], 1u);
      }
      ++part;
      next_part_level = S.ow_part_end[min(part, (uint32_t)OW_PARTS - 1u)];
    }
  }
  publish(part, base);
  if (rank == 0 && threadIdx.x == 0) {
    cnt->bfs_levels = levels;
    cnt->expanded = base;
    cnt->dirty_queue = 0;  // claim[] was written at frontier-mask cells only
    cnt->widest_level = widest;
    cnt->prof[4] = scans;
    cnt->prof[5] = clock64() - k0;
#ifdef FITSLAM_OW_PROF
    for (int k = 0; k < 8; ++k) printf("owc_prof[%d] = %u (%.1f per level)\n", k, pf.acc[k], (double)pf.acc[k] / levels);
#endif
  }
}

// One thread on a side stream: returns when part `part` of the outer wavefront has been published, so that the kernels
// queued behind it see its touch keys.  Gives up (and flags the search as failed) if that never happens.
__global__ void wait_part_kernel(Counters *cnt, uint32_t part) {
  const long long t0 = clock64();
  while (atomicAdd(&cnt->part_flag[part], 0u) == 0u) {
    __nanosleep(500);
    if (clock64() - t0 > (60ll << 30)) {  // ~30 s: the wavefront kernel never ran
      atomicOr(&cnt->overflow, 32u);
      break;
    }
  }
  __threadfence();
}

// ------------------------------------------------------------------------------------------------ K4: found components
// Collections.  The kernels of this section run once after each part of the outer wavefront (`part` = its index, on the
// second stream, while the next part runs) and once more after the last one (`part` = -1, everything that is left).
// A collection after part k only looks at touch keys below 4 * part_end[k]: those were written before part k ended, and a
// component with such a key has its final minimum (keys grow with the dequeue rank).  Keys at or above the bound may be
// half-written by the part running concurrently and are left for the next collection.
__device__ __forceinline__ uint32_t collect_bound(const Counters *cnt, int part) {
  return part < 0 ? KEY_INF : (TOUCH_BIT | (cnt->part_end[part] << 2));
}
// per component: smallest touching key (= the seed's key, FrontierSearch.cpp:78-81)
__global__ void touch_reduce_kernel(const SearchParams S, const uint32_t *__restrict__ mcell, const uint32_t *__restrict__ parent,
                                    const uint32_t *claim, uint32_t *rkey, const Counters *cnt, int part) {
  if (cnt->overflow) return;
  const uint32_t n = min(cnt->n_mask, S.mask_cap), bound = collect_bound(cnt, part);
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const uint32_t k = ld_cg(claim + mcell[i]);
    if (k < bound && k < ld_cg(rkey + parent[i])) atomicMin(rkey + parent[i], k);
  }
}
// the cell whose key equals its component's minimum is the seed: append {key, seed, size} to the component table (once:
// the top bit of the root's size marks a component as collected)
__global__ void seed_collect_kernel(const SearchParams S, const uint32_t *__restrict__ mcell, const uint32_t *__restrict__ parent,
                                    const uint32_t *claim, const uint32_t *rkey, uint32_t *rsize, const uint32_t *__restrict__ rbox,
                                    uint32_t *ckey, uint32_t *cseed, uint32_t *csize, uint32_t *cbox, Counters *cnt, int part) {
  if (cnt->overflow & 3u) return;
  const uint32_t n = min(cnt->n_mask, S.mask_cap), bound = collect_bound(cnt, part);
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const uint32_t c = mcell[i];
    const uint32_t k = ld_cg(claim + c);
    if (k >= bound) continue;
    const uint32_t r = parent[i];
    if (ld_cg(rkey + r) != k) continue;
    const uint32_t sz = rsize[r];  // keys are unique: only this thread matches the component
    if (sz & 0x80000000u) continue;
    rsize[r] = sz | 0x80000000u;
    const uint32_t pos = atomicAdd(&cnt->n_comp, 1u);
    if (pos < S.comp_cap) {
      ckey[pos] = k;
      cseed[pos] = c;
      csize[pos] = sz;
      reinterpret_cast<uint4 *>(cbox)[pos] = reinterpret_cast<const uint4 *>(rbox)[r];
    } else {
      atomicOr(&cnt->overflow, 4u);
    }
  }
}
// discovery order = ascending seed key (keys are unique): rank by counting, scatter into the ordered table.  A collection
// orders its own components [comp_done, n_comp): every earlier collection's keys are smaller.
__global__ void __launch_bounds__(256)
comp_order_kernel(const SearchParams S, const uint32_t *__restrict__ ckey, const uint32_t *__restrict__ cseed,
                  const uint32_t *__restrict__ csize, const uint32_t *__restrict__ cbox, uint32_t *oseed, uint32_t *osize,
                  uint32_t *obox, const Counters *cnt) {
  __shared__ uint32_t sk[256];
  const uint32_t n = min(cnt->n_comp, S.comp_cap), c0 = min(cnt->comp_done, n);
  for (uint32_t i0 = c0 + blockIdx.x * 256; i0 < n; i0 += gridDim.x * 256) {
    const uint32_t i = i0 + threadIdx.x;
    const uint32_t mine = i < n ? ckey[i] : 0;
    uint32_t r = c0;
    for (uint32_t base = c0; base < n; base += 256) {
      sk[threadIdx.x] = base + threadIdx.x < n ? ckey[base + threadIdx.x] : KEY_INF;
      __syncthreads();
      if (i < n) {
        const uint32_t m = min(256u, n - base);
        for (uint32_t t = 0; t < m; ++t) r += sk[t] < mine ? 1u : 0u;
      }
      __syncthreads();
    }
    if (i < n) {
      oseed[r] = cseed[i];
      osize[r] = csize[i];
      reinterpret_cast<uint4 *>(obox)[r] = reinterpret_cast<const uint4 *>(cbox)[i];
    }
  }
}
// exclusive scans over the ordered components: first cell / first frontier of each; totals into the counters.
// run lengths (FrontierSearch.cpp:150-199): full runs of max+1 cells, then a remainder kept iff it is > min.
__global__ void __launch_bounds__(1024)
comp_scan_kernel(const SearchParams S, const uint32_t *__restrict__ osize, uint32_t *ocell0, uint32_t *ofront0, Counters *cnt, int coll) {
  __shared__ uint32_t sh[40];
  const uint32_t n = min(cnt->n_comp, S.comp_cap), c0 = min(cnt->comp_done, n);
  const uint32_t run = (uint32_t)S.max_size + 1u;
  uint32_t cell_carry = cnt->cell_carry, front_carry = cnt->front_carry;
  __syncthreads();  // every thread has read the counters thread 0 rewrites below
  for (uint32_t base = c0; base < n; base += 1024) {
    const uint32_t i = base + threadIdx.x;
    const uint32_t sz = i < n ? osize[i] : 0;
    const uint32_t rem = sz % run;
    const uint32_t nf = i < n ? ((int)run > S.min_size ? sz / run : 0u) + ((int)rem > S.min_size ? 1u : 0u) : 0u;
    uint32_t tc, tf;
    const uint32_t oc = block_scan(sz, sh, &tc);
    const uint32_t of = block_scan(nf, sh, &tf);
    if (i < n) {
      ocell0[i] = cell_carry + oc;
      ofront0[i] = front_carry + of;
    }
    cell_carry += tc;
    front_carry += tf;
  }
  if (threadIdx.x == 0) {
    cnt->coll_c0[coll] = c0;
    cnt->coll_c1[coll] = n;
    cnt->coll_fr[coll][0] = cnt->front_carry;
    cnt->coll_fr[coll][1] = front_carry;
    cnt->comp_done = n;
    cnt->cell_carry = cell_carry;
    cnt->front_carry = front_carry;
    cnt->n_cells = cell_carry;
    cnt->n_frontiers = front_carry;
    if (cell_carry > S.cell_cap) atomicOr(&cnt->overflow, 8u);
    if (front_carry > S.front_cap) atomicOr(&cnt->overflow, 16u);
  }
}

// ------------------------------------------------------------------------------------------------ K5: inner wavefronts
// buildNewFrontier's FIFO, 8-connected wavefront over the frontier mask from the seed (DEP/src/FrontierSearch.cpp:123-189),
// one CTA per found component, same first-in-queue-order rule as the outer wavefront.  The component's slice of the
// every_frontier list IS the queue: level after level, in acceptance order.
__global__ void __launch_bounds__(REPLAY_THREADS)
replay_kernel(const SearchParams S, const uint32_t *__restrict__ bits, uint32_t *claim, const uint32_t *__restrict__ oseed,
              const uint32_t *__restrict__ osize, const uint32_t *__restrict__ ocell0, uint32_t *cells, const Counters *cnt) {
  __shared__ uint32_t sh[40];
  const uint32_t W = S.map.W, H = S.map.H;
  const uint32_t n_comp = min(cnt->n_comp, S.comp_cap);
  if (cnt->overflow) return;
  for (uint32_t comp = blockIdx.x; comp < n_comp; comp += gridDim.x) {
    uint32_t *out = cells + ocell0[comp];
    const uint32_t seed = oseed[comp];
    if (threadIdx.x == 0) {
      out[0] = seed;
      claim[seed] = INNER_BIT;
    }
    __syncthreads();
    uint32_t base = 0, n = 1;
    while (n > 0) {
      const uint32_t next_base = base + n, items = n << 3;
      uint32_t n_next = 0;
      for (uint32_t tile = 0; tile < items; tile += REPLAY_THREADS) {
        const uint32_t j = tile + threadIdx.x;
        bool win = false;
        uint32_t nb = 0;
        if (j < items) {
          const uint32_t rank = base + (j >> 3);
          const uint32_t p = ld_cg(out + rank);
          const uint32_t my = p / W, mx = p - my * W;
          uint32_t nx, ny;
          if (nbr_of(mx, my, W, H, (int)(j & 7u), nx, ny) && get_bit(bits, S.wpr, nx, ny)) {
            nb = ny * W + nx;
            if ((ld_cg(claim + nb) & 0xC0000000u) != INNER_BIT) {  // !frontier_flag[nbr]
              win = true;
#pragma unroll
              for (int t = 0; t < 8; ++t) {
                uint32_t ox, oy;
                if (!nbr_of(nx, ny, W, H, t, ox, oy)) continue;
                const uint32_t o = oy * W + ox;
                if (o == p) continue;
                const uint32_t r = ld_cg(claim + o);
                if (r >= (INNER_BIT | base) && r < (INNER_BIT | rank)) win = false;
              }
            }
          }
        }
        uint32_t tile_total;
        const uint32_t off = block_scan(win ? 1u : 0u, sh, &tile_total);
        if (win) {
          const uint32_t dst = next_base + n_next + off;
          out[dst] = nb;
          claim[nb] = INNER_BIT | dst;
        }
        n_next += tile_total;
        __syncthreads();
      }
      base = next_base;
      n = n_next;
    }
    __syncthreads();
  }
}


// The inner wavefront of one component out of shared memory: one WARP per component (components of at most
// `max_cells` cells; the launch is repeated per size class so that small components pack many warps per SM).
// One hash table holds every cell the wavefront has accepted, with the bid that won it: rank(parent) * 8 + slot.  Bids
// grow from level to level, so "frontier_flag[nbr]" is "the entry's bid is below base * 8", and within a level the
// smallest bid wins, as in the outer wavefront.  The component's cells (packed y << 16 | x) are mirrored in shared
// memory in queue order, and the frontier-mask bits of the component's bounding box (+1 cell) are staged in shared
// memory when they fit, so a level costs a handful of shared-memory round trips and two __syncwarp.
__global__ void __launch_bounds__(32)
replay_warp_kernel(const SearchParams S, const uint32_t *__restrict__ bits, const uint2 *__restrict__ tm, const uint32_t *__restrict__ oseed,
                   const uint32_t *__restrict__ osize, const uint32_t *__restrict__ ocell0, const uint32_t *__restrict__ obox,
                   uint32_t *cells, const Counters *cnt, uint32_t min_cells, uint32_t max_cells, uint32_t bit_words, int coll) {
  extern __shared__ uint32_t iw_sh[];
  uint32_t *keys = iw_sh, *val = iw_sh + 2 * max_cells, *mir = iw_sh + 4 * max_cells, *bw = iw_sh + 5 * max_cells;
  const uint32_t W = S.map.W, H = S.map.H;
  const uint32_t n_comp = min(cnt->coll_c1[coll], S.comp_cap);  // this collection's components
  if (cnt->overflow) return;
  const uint32_t lane = threadIdx.x;
  for (uint32_t comp = cnt->coll_c0[coll] + blockIdx.x; comp < n_comp; comp += gridDim.x) {
    const uint32_t size = osize[comp];
    if (size < min_cells || size > max_cells) continue;
    uint32_t lg = 5;
    while ((1u << lg) < 2u * size) ++lg;
    const uint32_t slots = 1u << lg, smask = slots - 1u;
    for (uint32_t i = lane; i < slots; i += 32) { keys[i] = HKEY_EMPTY; val[i] = KEY_INF; }
    // mask window: rows y0-1 .. y1+1, words of columns x0-1 .. x1+1 (clamped to the map)
    const uint4 box = reinterpret_cast<const uint4 *>(obox)[comp];
    const uint32_t wy0 = box.y > 0 ? box.y - 1 : 0, wy1 = min(box.w + 1, H - 1);
    const uint32_t wx0 = (box.x > 0 ? box.x - 1 : 0) >> 5, wx1 = min(box.z + 1, W - 1) >> 5;
    const uint32_t ww = wx1 - wx0 + 1, wh = wy1 - wy0 + 1;
    const bool staged = (uint64_t)ww * wh <= bit_words;
    if (staged)
      for (uint32_t i = lane; i < ww * wh; i += 32) {
        const uint32_t r = i / ww, cw = i - r * ww;
        bw[i] = __ldg(bits + (size_t)(wy0 + r) * S.wpr + wx0 + cw);
      }
    uint32_t *out = cells + ocell0[comp];
    if (lane == 0) {
      const uint32_t seed = oseed[comp];
      const uint32_t sy = seed / W, sx = seed - sy * W;
      const uint32_t pk = (sy << 16) | sx;
      out[0] = seed;
      mir[0] = pk;
      const uint32_t h = hslot(pk, lg);
      keys[h] = pk;
      val[h] = 0;
    }
    __syncwarp();
    uint32_t base = 0, n = 1;
    while (n > 0) {
      const uint32_t next_base = base + n, items = n << 3;
      uint32_t n_next = 0;
      for (uint32_t tile = 0; tile < items; tile += 32) {
        const uint32_t j = tile + lane;
        bool cand = false;
        uint32_t nb = 0xFFFFFF00u + lane;  // distinct dummies: lanes without a candidate match nobody
        if (j < items) {
          const uint32_t rank = base + (j >> 3), s = j & 7u;
          const uint32_t p = mir[rank];
          // neighbour s in the reference's order: -1, +1, -W, +W, -1-W, -1+W, +1-W, +1+W (DEP/src/Helpers.cpp:185-255)
          const int ddx = (int)((0xA058u >> (2 * s)) & 3u) - 1, ddy = (int)((0x8885u >> (2 * s)) & 3u) - 1;
          const int x = (int)(p & 0xFFFFu) + ddx, y = (int)(p >> 16) + ddy;
          if (x >= 0 && y >= 0 && x < (int)W && y < (int)H) {
            // not staged (a long thin component with a large bounding box): the tiled class words, whose 128-byte lines
            // hold 8 rows, so a wavefront creeping along the component misses L1 once in several levels, not every level
            const uint32_t word = staged ? bw[((uint32_t)y - wy0) * ww + (((uint32_t)x >> 5) - wx0)]
                                         : __ldg(tm + tm_index((uint32_t)x >> 5, (uint32_t)y, S.wpr)).y;
            if ((word >> ((uint32_t)x & 31u)) & 1u) {
              const uint32_t c = ((uint32_t)y << 16) | (uint32_t)x;
              // frontier_flag[nbr]: the table holds every cell accepted so far (plain loads: only winners write it,
              // and they did so before this tile started)
              uint32_t h = hslot(c, lg);
              bool seen = false;
              while (true) {
                const uint32_t k = ((volatile uint32_t *)keys)[h];
                if (k == c) { seen = true; break; }
                if (k == HKEY_EMPTY) break;
                h = (h + 1u) & smask;
              }
              if (!seen) { cand = true; nb = c; }
            }
          }
        }
        // several lanes may reach the same cell in this tile; the lowest lane is the smallest (rank, slot) and wins.  The
        // 8 lanes of one dequeued cell reach 8 different cells, so a lane loses exactly when its cell is also a neighbour
        // of one of the (at most 3) cells dequeued by lower lanes of this tile: three shuffles instead of a MATCH, which
        // the profile showed to be a third of the level time
        bool win = cand;
        {
          const uint32_t pme = (j < items) ? mir[base + (j >> 3)] : 0xFFFFFFFFu;
          const int nx = (int)(nb & 0xFFFFu), ny = (int)(nb >> 16);
#pragma unroll
          for (int k = 0; k < 3; ++k) {
            const uint32_t pk = __shfl_sync(0xffffffffu, pme, 8 * k);
            const int dx = nx - (int)(pk & 0xFFFFu), dy = ny - (int)(pk >> 16);
            if ((int)(lane >> 3) > k && pk != 0xFFFFFFFFu && dx >= -1 && dx <= 1 && dy >= -1 && dy <= 1) win = false;
          }
        }
        const uint32_t bal = __ballot_sync(0xffffffffu, win);
        if (win) {
          const uint32_t dst = next_base + n_next + __popc(bal & ((1u << lane) - 1u));
          const uint32_t x = nb & 0xFFFFu, y = nb >> 16;
          out[dst] = y * W + x;
          mir[dst] = nb;
          uint32_t h = hslot(nb, lg);
          while (atomicCAS(&keys[h], HKEY_EMPTY, nb) != HKEY_EMPTY) h = (h + 1u) & smask;  // winners are distinct cells
        }
        n_next += __popc(bal);
        __syncwarp();
      }
      base = next_base;
      n = n_next;
    }
    __syncwarp();
  }
}

// The inner wavefront of a component above IW_SMALL cells: one warp walks buildNewFrontier's loop (FrontierSearch.cpp:123-189),
// one or two dequeued cells per step, lane s of a group of 8 takes neighbour s in the reference's order.  Components of this
// class are long and thin (arcs of sensor discs: 1-2 cells per level for thousands of levels), so what bounds them is the
// latency of a step: a shared-memory read of the queue head (the queue's tail is mirrored in a ring), a mask-bit load, the
// visited test and a ballot that appends the accepted neighbours in slot order.  While the last three levels hold at most 64
// cells the visited test is a search of the ring (see the loop); wider levels read the visited flag in the top bit of the
// cell's slot[] word (slot[] is rewritten by every search, and a cell belongs to one component only).
static constexpr uint32_t IL_RING = 2048;
static constexpr uint32_t IL_VISITED = 0x80000000u;
__global__ void __launch_bounds__(32)
replay_line_kernel(const SearchParams S, const uint32_t *__restrict__ bits, uint32_t *slot, const uint32_t *__restrict__ oseed,
                   const uint32_t *__restrict__ osize, const uint32_t *__restrict__ ocell0, uint32_t *cells, const Counters *cnt,
                   uint32_t min_cells, uint32_t max_cells, int coll) {
  __shared__ uint32_t ring[IL_RING];
  const uint32_t W = S.map.W, H = S.map.H;
  if (cnt->overflow) return;
  const uint32_t c_lo = coll < 0 ? 0u : cnt->coll_c0[coll], c_hi = min(coll < 0 ? cnt->n_comp : cnt->coll_c1[coll], S.comp_cap);
  const uint32_t lane = threadIdx.x, sl = lane & 7u;
  const bool act = lane < 8u;
  // neighbour sl in the reference's order: -1, +1, -W, +W, -1-W, -1+W, +1-W, +1+W (DEP/src/Helpers.cpp:185-255)
  const int ddx = (int)((0xA058u >> (2 * sl)) & 3u) - 1, ddy = (int)((0x8885u >> (2 * sl)) & 3u) - 1;
  const uint32_t mW = 0xFFFFFFFFu / W;  // floor(2^32 / W), or one less: the quotient estimate below is at most 2 short
  const uint32_t lt = (1u << lane) - 1u;
  for (uint32_t comp = c_lo + blockIdx.x; comp < c_hi; comp += gridDim.x) {
    const uint32_t size = osize[comp];
    if (size < min_cells || size > max_cells) continue;
    uint32_t *out = cells + ocell0[comp];
    if (lane == 0) {
      const uint32_t seed = oseed[comp];
      out[0] = seed;
      ring[0] = (seed % W) | ((seed / W) << 16);  // the ring holds x | y << 16 (W, H < 65536 as in replay_warp_kernel)
      slot[seed] |= IL_VISITED;
    }
    __syncwarp();
    uint32_t head = 0, tail = 1;
    // queue positions where the level before the dequeued cell's, that cell's level and the next one begin: a visited neighbour
    // of a level-L cell lies in levels L - 1 .. L + 1 (the 8-neighbourhood is symmetric), that is in [lv_prev, tail)
    uint32_t lv_prev = 0, lv_cur = 0, lv_end = 1;
#ifdef FITSLAM_IL_PROF
    long long pc[5] = {0, 0, 0, 0, 0}, t0 = clock64(), t1;
    uint32_t n_win = 0, n_fall = 0;
#define IL_T(k) { t1 = clock64(); pc[k] += t1 - t0; t0 = t1; }
#else
#define IL_T(k)
#endif
    while (head < tail) {
      if (head == lv_end) { lv_prev = lv_cur; lv_cur = lv_end; lv_end = tail; }
      const uint32_t wlen = tail - lv_prev;
      if (wlen <= 64u) {
        // Thin levels (the arcs this kernel is for).  The visited test is a search of the last three levels in the ring, 64
        // entries over the 32 lanes, instead of a load of slot[] (whose line the store below evicts from L1 every step: an L2
        // round trip per dequeued cell): each lane marks which cells of the 3 x 3 block around a dequeued cell its two entries
        // are, one warp-wide OR gives the block's visited cells.  And the step takes TWO cells, A = queue[head] on lanes 0-7 and
        // B = queue[head + 1] on lanes 8-15, when B is of A's level: a neighbour of B inside A's block that passes B's tests
        // passes A's too (same mask bit, same visited set), so A accepts it first and B's lane drops it; lane order is then
        // the reference's append order.  slot[]'s flag is still set, for the steps that take the other branch.
        const bool pair = head + 1u < lv_end;
        const uint32_t pa = ring[head & (IL_RING - 1u)], pb = ring[(head + 1u) & (IL_RING - 1u)];
        const uint32_t xa = pa & 0xFFFFu, ya = pa >> 16, xb = pb & 0xFFFFu, yb = pb >> 16;
        const bool second = (lane & 8u) != 0u;
        const bool act2 = lane < 8u || (lane < 16u && pair);
        const int nx = (int)(second ? xb : xa) + ddx, ny = (int)(second ? yb : ya) + ddy;
        const bool in = act2 && nx >= 0 && ny >= 0 && nx < (int)W && ny < (int)H;
        const uint32_t nb = (uint32_t)ny * W + (uint32_t)nx;
        IL_T(0)
        const uint32_t mw = in ? __ldg(bits + (size_t)ny * S.wpr + ((uint32_t)nx >> 5)) : 0u;
#ifdef FITSLAM_IL_PROF
        ++n_win;
#endif
        uint32_t m = 0;
#pragma unroll
        for (uint32_t k = 0; k < 2u; ++k) {
          if (lane + 32u * k < wlen) {
            const uint32_t e = ring[(lv_prev + 32u * k + lane) & (IL_RING - 1u)];
            const uint32_t ex = e & 0xFFFFu, ey = e >> 16;
            const uint32_t dxa = ex - xa + 1u, dya = ey - ya + 1u, dxb = ex - xb + 1u, dyb = ey - yb + 1u;
            if (dxa < 3u && dya < 3u) m |= 1u << (dya * 3u + dxa);
            if (dxb < 3u && dyb < 3u) m |= 512u << (dyb * 3u + dxb);
          }
        }
        const uint32_t seen = __reduce_or_sync(0xffffffffu, m);
        IL_T(1)
        const bool bit = (mw >> ((uint32_t)nx & 31u)) & 1u;
#ifdef FITSLAM_IL_PROF
        if (__ballot_sync(0xffffffffu, bit)) {}
        IL_T(2)
#endif
        const bool in_a = (uint32_t)(nx - (int)xa + 1) < 3u && (uint32_t)(ny - (int)ya + 1) < 3u;
        const bool acc = in && bit && !((seen >> ((second ? 9u : 0u) + (uint32_t)((ddy + 1) * 3 + (ddx + 1)))) & 1u) && !(second && in_a);
        const uint32_t bal = __ballot_sync(0xffffffffu, acc);
        if (acc) {
          const uint32_t at = tail + __popc(bal & lt);
          atomicOr(&slot[nb], IL_VISITED);
          out[at] = nb;
          ring[at & (IL_RING - 1u)] = (uint32_t)nx | ((uint32_t)ny << 16);
        }
        tail += __popc(bal);
        head += pair ? 2u : 1u;
        IL_T(3)
      } else {
        // (the ring holds positions [tail - IL_RING, tail): a level wider than that reads the list itself, through L2)
        uint32_t x, y;
        if (tail - head <= IL_RING) {
          const uint32_t pxy = ring[head & (IL_RING - 1u)];
          x = pxy & 0xFFFFu;
          y = pxy >> 16;
        } else {
          const uint32_t p = ld_cg(out + head);
          y = __umulhi(p, mW);
          x = p - y * W;
          while (x >= W) { x -= W; ++y; }
        }
        ++head;
        const int nx = (int)x + ddx, ny = (int)y + ddy;
        const bool in = act && nx >= 0 && ny >= 0 && nx < (int)W && ny < (int)H;
        const uint32_t nb = (uint32_t)ny * W + (uint32_t)nx;
        // both loads at once (the slot word of a cell that turns out not to be a mask cell is read and dropped): a step is one
        // memory round trip, not two
        const uint32_t mw = in ? __ldg(bits + (size_t)ny * S.wpr + ((uint32_t)nx >> 5)) : 0u;
        const uint32_t sw = in ? ld_cg(slot + nb) : IL_VISITED;
#ifdef FITSLAM_IL_PROF
        ++n_fall;
#endif
        const bool bit = (mw >> ((uint32_t)nx & 31u)) & 1u;
        const uint32_t sv = bit ? sw : IL_VISITED;  // isNewFrontierCell: a mask cell whose frontier_flag is not set yet
        const bool acc = !(sv & IL_VISITED);
        const uint32_t bal = __ballot_sync(0xffffffffu, acc);
        if (acc) {
          const uint32_t at = tail + __popc(bal & lt);
          slot[nb] = sv | IL_VISITED;
          out[at] = nb;
          ring[at & (IL_RING - 1u)] = (uint32_t)nx | ((uint32_t)ny << 16);
        }
        tail += __popc(bal);
      }
      __syncwarp();
      IL_T(4)
    }
#ifdef FITSLAM_IL_PROF
    if (lane == 0 && tail > 5000) printf("line comp %u cells %u window steps %u fallback %u cycles: head %lld window %lld maskwait %lld append %lld sync/fallback %lld\n", comp, tail, n_win, n_fall, pc[0], pc[1], pc[2], pc[3], pc[4]);
#endif
  }
}

// ------------------------------------------------------------------------------------------------ K6: runs -> frontiers
// SortByMedianFunctor (FrontierSearch.hpp:156-184) on precomputed angles
__device__ __forceinline__ bool median_less(double angle_a, double angle_b) {
  if (0 <= angle_a && angle_a <= M_PI / 2 && 3 * M_PI / 2 <= angle_b && angle_b <= 2 * M_PI) return false;
  if (0 <= angle_b && angle_b <= M_PI / 2 && 3 * M_PI / 2 <= angle_a && angle_a <= 2 * M_PI) return true;
  return angle_a < angle_b;
}

struct SortCtx {
  double *ang;     // angle of element k
  uint8_t *idx;    // permutation being sorted
  int n;
  bool bad;        // an unguarded loop left [0, n): the comparator is not a strict weak order on this input
};
#define SLESS(a, b) median_less(sc.ang[sc.idx[a]], sc.ang[sc.idx[b]])
__device__ __forceinline__ void s_swap(SortCtx &sc, int a, int b) { const uint8_t t = sc.idx[a]; sc.idx[a] = sc.idx[b]; sc.idx[b] = t; }

// libstdc++ __unguarded_linear_insert / __insertion_sort (bits/stl_algo.h)
__device__ void s_unguarded_linear_insert(SortCtx &sc, int last) {
  const uint8_t val = sc.idx[last];
  int next = last - 1;
  while (next >= 0 && median_less(sc.ang[val], sc.ang[sc.idx[next]])) {
    sc.idx[last] = sc.idx[next];
    last = next;
    --next;
  }
  if (next < 0 && last == 0) {
    // the guarded form never gets here with a consistent comparator: position 0 is <= val by construction
    sc.bad = true;
  }
  sc.idx[last] = val;
}
__device__ void s_insertion_sort(SortCtx &sc, int first, int last) {
  if (first == last) return;
  for (int i = first + 1; i != last; ++i) {
    if (SLESS(i, first)) {
      const uint8_t val = sc.idx[i];
      for (int k = i; k > first; --k) sc.idx[k] = sc.idx[k - 1];
      sc.idx[first] = val;
    } else {
      s_unguarded_linear_insert(sc, i);
    }
  }
}
// __move_median_to_first + __unguarded_partition
__device__ int s_partition_pivot(SortCtx &sc, int first, int last) {
  const int mid = first + (last - first) / 2;
  const int a = first + 1, b = mid, c = last - 1, result = first;
  if (SLESS(a, b)) {
    if (SLESS(b, c)) s_swap(sc, result, b);
    else if (SLESS(a, c)) s_swap(sc, result, c);
    else s_swap(sc, result, a);
  } else if (SLESS(a, c)) s_swap(sc, result, a);
  else if (SLESS(b, c)) s_swap(sc, result, c);
  else s_swap(sc, result, b);
  int lo = first + 1, hi = last;
  const int pivot = first;
  while (true) {
    while (lo < sc.n && SLESS(lo, pivot)) ++lo;
    --hi;
    while (hi >= 0 && SLESS(pivot, hi)) --hi;
    if (lo >= sc.n || hi < 0) { sc.bad = true; return min(max(lo, first + 1), last); }
    if (!(lo < hi)) return lo;
    s_swap(sc, lo, hi);
    ++lo;
  }
}
// __adjust_heap / __make_heap / __sort_heap: the depth-limit fallback (__partial_sort(first, last, last))
__device__ void s_adjust_heap(SortCtx &sc, int first, int hole, int len, uint8_t value) {
  const int top = hole;
  int child = hole;
  while (child < (len - 1) / 2) {
    child = 2 * (child + 1);
    if (median_less(sc.ang[sc.idx[first + child]], sc.ang[sc.idx[first + child - 1]])) --child;
    sc.idx[first + hole] = sc.idx[first + child];
    hole = child;
  }
  if ((len & 1) == 0 && child == (len - 2) / 2) {
    child = 2 * (child + 1);
    sc.idx[first + hole] = sc.idx[first + child - 1];
    hole = child - 1;
  }
  int parent = (hole - 1) / 2;  // __push_heap
  while (hole > top && median_less(sc.ang[sc.idx[first + parent]], sc.ang[value])) {
    sc.idx[first + hole] = sc.idx[first + parent];
    hole = parent;
    parent = (hole - 1) / 2;
  }
  sc.idx[first + hole] = value;
}
__device__ void s_heap_sort(SortCtx &sc, int first, int last) {
  const int len = last - first;
  if (len >= 2) {
    int parent = (len - 2) / 2;
    while (true) {
      const uint8_t v = sc.idx[first + parent];
      s_adjust_heap(sc, first, parent, len, v);
      if (parent == 0) break;
      --parent;
    }
  }
  for (int l = last; l - first > 1;) {
    --l;
    const uint8_t v = sc.idx[l];
    sc.idx[l] = sc.idx[first];
    s_adjust_heap(sc, first, 0, l - first, v);
  }
}
// std::sort: __introsort_loop (threshold 16, depth limit 2 * floor(log2 n)) + __final_insertion_sort
__device__ void s_std_sort(SortCtx &sc) {
  const int n = sc.n;
  if (n < 2) return;
  // the recursion `__introsort_loop(cut, last)` then `last = cut` is unrolled with an explicit stack of (first, last, depth)
  int stk_first[16], stk_last[16], stk_depth[16];
  int sp = 0;
  stk_first[0] = 0; stk_last[0] = n; stk_depth[0] = 2 * (31 - __clz(n)); sp = 1;
  while (sp > 0) {
    --sp;
    int first = stk_first[sp], last = stk_last[sp], depth = stk_depth[sp];
    // the reference recurses on the RIGHT part first and loops on the left part; a stack pops in LIFO order, so
    // process this range's loop and push right parts; order of independent sub-ranges does not affect the result
    while (last - first > 16) {
      if (depth == 0) {
        s_heap_sort(sc, first, last);
        break;
      }
      --depth;
      const int cut = s_partition_pivot(sc, first, last);
      if (sp < 16) { stk_first[sp] = cut; stk_last[sp] = last; stk_depth[sp] = depth; ++sp; }
      else sc.bad = true;
      last = cut;
    }
  }
  if (n > 16) {
    s_insertion_sort(sc, 0, 16);
    for (int i = 16; i < n; ++i) s_unguarded_linear_insert(sc, i);
  } else {
    s_insertion_sort(sc, 0, n);
  }
}

// One thread per kept run: getCentroidOfCells (FrontierSearch.hpp:85-126), std::sort with SortByMedianFunctor,
// goal = sorted[size / 2] (FrontierSearch.cpp:161-165, 194-199).
__global__ void __launch_bounds__(64)
chunk_kernel(const SearchParams S, const uint32_t *__restrict__ osize, const uint32_t *__restrict__ ocell0,
             const uint32_t *__restrict__ ofront0, const uint32_t *__restrict__ cells, fitslam_frontier *out,
             double *centroids, uint8_t *ftag, uint32_t *fix_list, Counters *cnt, int coll, uint32_t min_cells, uint32_t max_cells,
             uint32_t tag) {
  // coll >= 0: the frontiers of collection `coll` whose component is in the size class [min_cells, max_cells] (its inner
  // wavefronts have just been replayed on this stream), at most EARLY_SLICE_CAP of them; they get `tag`, and the launch
  // that scores them follows on the same stream.  coll < 0, after the search: every frontier not built yet, TAG_FINAL.
  if (cnt->overflow) return;
  const uint32_t c_lo = coll < 0 ? 0u : cnt->coll_c0[coll], c_hi = coll < 0 ? min(cnt->n_comp, S.comp_cap) : min(cnt->coll_c1[coll], S.comp_cap);
  const uint32_t f_lo = coll < 0 ? 0u : cnt->coll_fr[coll][0];
  const uint32_t f_hi = coll < 0 ? cnt->n_frontiers : min(min(cnt->coll_fr[coll][1], S.early_limit), f_lo + S.early_cap);
  if (f_hi > S.front_cap) return;
  const uint32_t run = (uint32_t)S.max_size + 1u;
  const uint32_t W = S.map.W;
  for (uint32_t f = f_lo + blockIdx.x * blockDim.x + threadIdx.x; f < f_hi; f += gridDim.x * blockDim.x) {
    if (coll < 0 && ftag[f] != 0) continue;
    // component of frontier f: last comp with ofront0[comp] <= f
    uint32_t lo = c_lo, hi = c_hi;
    while (hi - lo > 1) {
      const uint32_t mid = (lo + hi) >> 1;
      if (ofront0[mid] <= f) lo = mid; else hi = mid;
    }
    const uint32_t comp = lo, j = f - ofront0[comp];
    if (coll >= 0 && (osize[comp] < min_cells || osize[comp] > max_cells)) continue;
    const uint32_t first_cell = ocell0[comp] + j * run;
    const int n = (int)min(run, osize[comp] - j * run);
    const uint32_t *cl = cells + first_cell;
    double ang[MAX_CHUNK];
    uint8_t idx[MAX_CHUNK];
    double sumX = 0, sumY = 0;
    for (int k = 0; k < n; ++k) {
      const uint32_t c = cl[k];
      const uint32_t my = c / W, mx = c - my * W;
      sumX += S.map.ox + (mx + 0.5) * S.map.res;
      sumY += S.map.oy + (my + 0.5) * S.map.res;
    }
    double centerX = sumX / (double)n, centerY = sumY / (double)n;
    bool offset_centroid = false;
    double varX = 0, varY = 0;
    for (int k = 0; k < n; ++k) {
      const uint32_t c = cl[k];
      const uint32_t my = c / W, mx = c - my * W;
      const double dx = (S.map.ox + (mx + 0.5) * S.map.res) - centerX, dy = (S.map.oy + (my + 0.5) * S.map.res) - centerY;
      if (sqrt(dx * dx + dy * dy) < S.map.res * 3) offset_centroid = true;
      varX += fabs(dx);
      varY += fabs(dy);
    }
    const double distance_to_offset = S.map.res * 1.414 * 2;
    if (varX > varY && offset_centroid) centerY -= distance_to_offset;
    if (varX < varY && offset_centroid) centerX -= distance_to_offset;
    // Angles: correctly rounded (atan2_cr.h).  glibc's atan2, which the reference sorts by, returns the same value unless the
    // exact one lies within 0.03 ulp of a rounding boundary (`risky`), and then at most its neighbour.  A risky angle can
    // only change the outcome where a comparison hangs on its last bit: another angle within 2 ulp, or one of the
    // comparator's thresholds.  Such a chunk (in practice: collinear cells of a straight axis-aligned frontier, whose
    // angles all sit within a few ulp of pi / 2) is sorted again by the host with its own libm (fix_flagged_frontiers).
    unsigned long long risky = 0ull;
    for (int k = 0; k < n; ++k) {
      const uint32_t c = cl[k];
      const uint32_t my = c / W, mx = c - my * W;
      bool rk;
      double a = atan2cr::atan2_cr((S.map.oy + (my + 0.5) * S.map.res) - centerY, (S.map.ox + (mx + 0.5) * S.map.res) - centerX, &rk);
      if (a < 0) a = a + (2 * M_PI);
      ang[k] = a;
      idx[k] = (uint8_t)k;
      if (rk) risky |= 1ull << k;
    }
    SortCtx sc{ang, idx, n, false};
    s_std_sort(sc);
    bool to_host = sc.bad || S.fix_all;
    if (risky && !to_host) {
      const long long t0 = __double_as_longlong(M_PI / 2), t1 = __double_as_longlong(3 * M_PI / 2), t2 = __double_as_longlong(2 * M_PI);
      for (int k = 0; k < n && !to_host; ++k) {
        if (!((risky >> k) & 1ull)) continue;
        const long long a = __double_as_longlong(ang[k]);  // (angles are in [0, 2 pi]: bit distance = distance in ulp)
        if (llabs(a - t0) <= 2 || llabs(a - t1) <= 2 || llabs(a - t2) <= 2) to_host = true;
        for (int j = 0; j < n; ++j)
          if (j != k && llabs(a - __double_as_longlong(ang[j])) <= 2) to_host = true;
      }
    }
    if (to_host) {
      atomicAdd(&cnt->ambiguous, 1u);
      fix_list[atomicAdd(&cnt->pad0[0], 1u)] = f;  // (one entry per frontier at most: the list holds front_cap)
    }
    const uint32_t g = cl[idx[n / 2]];
    fitslam_frontier r;
    r.goal_my = g / W;
    r.goal_mx = g - r.goal_my * W;
    r.goal_x = S.map.ox + (r.goal_mx + 0.5) * S.map.res;
    r.goal_y = S.map.oy + (r.goal_my + 0.5) * S.map.res;
    r.size = n;
    r.component = comp;
    r.chunk = j;
    r.first_cell = first_cell;
    out[f] = r;
    ftag[f] = (uint8_t)(to_host ? TAG_HOST : tag);
    if (centroids) {
      centroids[2 * f] = centerX;
      centroids[2 * f + 1] = centerY;
    }
  }
}

// ---- the chunks whose median sort goes to the host (chunk_kernel)
static constexpr uint32_t FIX_WORDS = 2 + MAX_CHUNK;  // per entry: frontier, cell count, cells
__global__ void fix_gather_kernel(const uint32_t *__restrict__ fix_list, uint32_t n, const fitslam_frontier *__restrict__ fr,
                                  const uint32_t *__restrict__ cells, uint32_t *stage) {
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n * FIX_WORDS; i += gridDim.x * blockDim.x) {
    const uint32_t e = i / FIX_WORDS, k = i - e * FIX_WORDS;
    const uint32_t f = fix_list[e];
    const uint32_t sz = (uint32_t)fr[f].size;
    stage[i] = k == 0 ? f : (k == 1 ? sz : (k - 2 < sz ? cells[fr[f].first_cell + k - 2] : 0u));
  }
}
// stage on the way back: per entry {frontier, goal cell}
__global__ void fix_apply_kernel(const uint32_t *__restrict__ stage, uint32_t n, MapView map, fitslam_frontier *fr, uint8_t *ftag) {
  const uint32_t e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= n) return;
  const uint32_t f = stage[2 * e], g = stage[2 * e + 1];
  fr[f].goal_my = g / map.W;
  fr[f].goal_mx = g - fr[f].goal_my * map.W;
  fr[f].goal_x = map.ox + (fr[f].goal_mx + 0.5) * map.res;
  fr[f].goal_y = map.oy + (fr[f].goal_my + 0.5) * map.res;
  ftag[f] = (uint8_t)TAG_FINAL;  // scored by the launches that follow the search
}

// ------------------------------------------------------------------------------------------------ K7: clean claim[]
// every cell the search wrote is a queue entry, one of its 4 neighbours, or a frontier-mask cell
__global__ void cleanup_kernel(const SearchParams S, const uint32_t *__restrict__ queue, const uint32_t *__restrict__ mcell,
                               uint32_t *claim, const Counters *cnt) {
  const uint32_t W = S.map.W, H = S.map.H;
  const uint32_t nq = cnt->dirty_queue, nm = min(cnt->n_mask, S.mask_cap);
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < nq + nm; i += gridDim.x * blockDim.x) {
    if (i < nq) {
      const uint32_t c = queue[i];
      const uint32_t my = c / W, mx = c - my * W;
      claim[c] = KEY_INF;
      if (mx > 0) claim[c - 1] = KEY_INF;
      if (mx + 1 < W) claim[c + 1] = KEY_INF;
      if (my > 0) claim[c - W] = KEY_INF;
      if (my + 1 < H) claim[c + W] = KEY_INF;
    } else {
      claim[mcell[i - nq]] = KEY_INF;
    }
  }
}

}  // namespace fitslam

// ================================================================================================ host side
namespace fitslam {

static fitslam_status ensure_frontier_buffers(fitslam_ctx *c, FrontierState *s) {
  const uint64_t N = (uint64_t)c->W * c->H;
  if (N > MAX_RANK) return fail(c, FITSLAM_E_INVALID, "search_frontiers: more than 2^29 cells");
  const uint32_t wpr = (c->W + 31u) / 32u;
  FS_CUDA(c, s->bits.reserve((size_t)wpr * c->H));
  FS_CUDA(c, s->trav.reserve((size_t)wpr * c->H));
  FS_CUDA(c, s->tm.reserve((size_t)wpr * ((c->H + 7u) / 8u) * 8u));
  FS_CUDA(c, s->ebits.reserve((size_t)wpr * c->H));
  {
    const OwLayout L = ow_layout(c->W, c->H);
    const uint64_t pc = (uint64_t)L.Wp * L.Hp;
    if (s->rk.cap < pc || s->ow_cells != pc) {  // padding: OW_BLOCKED (0xFDFD is above OW_MAX_LEVEL and none of the other classes)
      FS_CUDA(c, s->rk.reserve(pc));
      FS_CUDA(c, cudaMemsetAsync(s->rk.p, 0xFD, s->rk.cap * sizeof(uint16_t), c->stream));
      s->ow_cells = pc;
    }
  }
  FS_CUDA(c, s->slot.reserve(N));
  FS_CUDA(c, s->queue.reserve(N + 8));
  FS_CUDA(c, s->counters.reserve(1));
  if (s->claim.cap < N || s->claim_cells != N) {
    FS_CUDA(c, s->claim.reserve(N));
    FS_CUDA(c, cudaMemsetAsync(s->claim.p, 0xFF, s->claim.cap * sizeof(uint32_t), c->stream));
    s->claim_cells = N;
  }
  if (s->mask_cap == 0) {
    s->mask_cap = (uint32_t)std::min<uint64_t>(N, std::max<uint64_t>(1u << 16, N / 32));
    s->comp_cap = s->mask_cap / 4 + 1024;
    s->cell_cap = s->mask_cap;
    s->front_cap = s->mask_cap / 2 + 1024;
  }
  FS_CUDA(c, s->mcell.reserve(s->mask_cap));
  FS_CUDA(c, s->parent.reserve(s->mask_cap));
  FS_CUDA(c, s->rkey.reserve(s->mask_cap));
  FS_CUDA(c, s->rsize.reserve(s->mask_cap));
  FS_CUDA(c, s->rbox.reserve(4 * (size_t)s->mask_cap));
  FS_CUDA(c, s->cbox.reserve(4 * (size_t)s->comp_cap));
  FS_CUDA(c, s->obox.reserve(4 * (size_t)s->comp_cap));
  FS_CUDA(c, s->ckey.reserve(s->comp_cap));
  FS_CUDA(c, s->cseed.reserve(s->comp_cap));
  FS_CUDA(c, s->csize.reserve(s->comp_cap));
  FS_CUDA(c, s->oseed.reserve(s->comp_cap));
  FS_CUDA(c, s->osize.reserve(s->comp_cap));
  FS_CUDA(c, s->ocell0.reserve(s->comp_cap));
  FS_CUDA(c, s->ofront0.reserve(s->comp_cap));
  FS_CUDA(c, s->cells.reserve(s->cell_cap));
  FS_CUDA(c, s->centroids.reserve(2 * (size_t)s->front_cap));
  FS_CUDA(c, s->ftag.reserve(s->front_cap));
  FS_CUDA(c, s->fix_list.reserve(s->front_cap));
  FS_CUDA(c, c->frontiers.reserve(s->front_cap));
  return FITSLAM_OK;
}

// Enqueue one search on the context stream; counters land in *host_counters (pinned) after the stream drains.
static fitslam_status enqueue_search(fitslam_ctx *c, FrontierState *s, uint32_t robot_cell, double rx, double ry,
                                     Counters *host_counters, const EarlyHook *hook, int phase) {
  SearchParams S;
  memset(&S, 0, sizeof(S));
  S.map = map_view(c);
  S.wpr = (c->W + 31u) / 32u;
  S.lethal = (uint32_t)c->prm.lethal_threshold;
  S.robot_cell = robot_cell;
  S.robot_x = rx;
  S.robot_y = ry;
  const double mfd = s->search_distance >= 0.0 ? s->search_distance : c->prm.max_frontier_distance;
  S.dist_limit = mfd + (c->prm.max_frontier_cluster_size * c->res * 1.414);  // FrontierSearch.cpp:74
  {
    const double L = S.dist_limit / c->res;
    S.near_cells_sq = L > 4.0 ? (int64_t)std::min(L * L * (1.0 - 1e-6), 4e18) - 2 : 0;
    S.far_cells_sq = (int64_t)std::min(L * L * (1.0 + 1e-6), 4e18) + 2;
  }
  if (s->force_global < 0) {
    const char *e = getenv("FITSLAM_WAVEFRONT");
    s->force_global = (e && !strcmp(e, "global")) ? 1 : 0;
  }
  {
    const char *e = getenv("FITSLAM_OW_QCAP");
    const long v = e ? atol(e) : (long)OW_QCAP;
    S.ow_qcap = (uint32_t)std::min<long>(std::max<long>(v, 1), (long)OW_QCAP);
  }
  // a batch is about 1/12 of the frontiers; the launches are sized for a quarter of the last search's count
  S.early_cap = std::min<uint32_t>(std::max<uint32_t>(EARLY_SLICE_CAP, s->last.n_frontiers / 4u), 8192u);
  if (const char *e = getenv("FITSLAM_EARLY_CAP")) S.early_cap = (uint32_t)std::min<long>(std::max<long>(atol(e), 1), 8192);
  S.early_limit = hook ? hook->limit : 0xFFFFFFFFu;
  S.ow_part_levels = (uint32_t)std::min<double>(std::max<double>(S.dist_limit / c->res / OW_PARTS, 64.0), 1e6);
  {
    // When the parts end.  First search on a map: every ow_part_levels levels (the search radius in cells / OW_PARTS).  After
    // that the number of levels is known from the last search, and the boundaries crowd towards its end: what is found after
    // the last boundary waits for the search to finish before anything of it can be scored, and on wide maps a fifth of the
    // frontiers used to be found there.  (A search shorter than the last one just never reaches the late boundaries.)
    const uint32_t known = s->last.bfs_levels;
    double late[4] = {0.85, 0.92, 0.965, 0.99};
    if (const char *e = getenv("FITSLAM_OW_LATE")) sscanf(e, "%lf,%lf,%lf,%lf", &late[0], &late[1], &late[2], &late[3]);  // (tuning)
    uint32_t prev = 0;
    for (int k = 0; k < OW_PARTS; ++k) {
      uint32_t e;
      const int n_uniform = OW_PARTS - 1 - 4;  // boundaries 0 .. n_uniform - 1 share the first 3/4 of the levels
      if (!known || OW_PARTS < 8) e = (uint32_t)(k + 1) * S.ow_part_levels;
      else if (k < n_uniform) e = (uint32_t)((k + 1) * 0.75 * known / n_uniform);
      else if (k < OW_PARTS - 1) e = (uint32_t)(late[k - n_uniform] * known);
      else e = 0xFFFFFFFFu;
      e = std::max(e, prev + 16u);
      S.ow_part_end[k] = e;
      prev = e;
    }
  }
  S.ow_wide = getenv("FITSLAM_OW_NOWIDE") ? 0u : 1u;
  S.fix_all = getenv("FITSLAM_FIX_ALL") ? 1u : 0u;
  S.use_smem = (s->force_global || c->W > 16384u || c->H > 16384u) ? 0u : 1u;  // queue entries pack x | y << 14 | mask << 28
  if (!s->attr_set) {
    FS_CUDA(c, cudaFuncSetAttribute(ow_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)OW_SMEM));
    FS_CUDA(c, cudaFuncSetAttribute(ow_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)OW_SMEM));
    FS_CUDA(c, cudaFuncSetAttribute(replay_warp_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)iw_smem(IW_CELLS, IW_BITS_LARGE)));
    // the cluster wavefront: 16 CTAs (non-portable size) where the device takes it, else 8, else the single-CTA kernel
    s->ow_cluster = 0;
    if (const char *e = getenv("FITSLAM_OW_CLUSTER_MIN")) s->ow_cluster_min = (uint32_t)atol(e);
    const char *ev = getenv("FITSLAM_OW_CLUSTER");  // 0: single-CTA kernel; 8 / 16: that cluster size
    const int want = ev ? atoi(ev) : 16;
    if (want > 1 && cudaFuncSetAttribute(ow_cluster_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)OWC_SMEM) == cudaSuccess) {
      for (int cs = want >= 16 ? 16 : 8; cs >= 8 && !s->ow_cluster; cs -= 8) {
        if (cs > 8 && cudaFuncSetAttribute(ow_cluster_kernel, cudaFuncAttributeNonPortableClusterSizeAllowed, 1) != cudaSuccess) continue;
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(cs);
        cfg.blockDim = dim3(OWC_THREADS);
        cfg.dynamicSmemBytes = OWC_SMEM;
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeClusterDimension;
        at[0].val.clusterDim.x = cs; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
        cfg.attrs = at;
        cfg.numAttrs = 1;
        int nc = 0;
        if (cudaOccupancyMaxActiveClusters(&nc, ow_cluster_kernel, &cfg) == cudaSuccess && nc > 0) s->ow_cluster = cs;
      }
      cudaGetLastError();
    }
    s->attr_set = true;
  }
  S.min_size = c->prm.min_frontier_cluster_size;
  S.max_size = c->prm.max_frontier_cluster_size;
  S.mask_cap = s->mask_cap;
  S.queue_cap = (uint32_t)std::min<size_t>(s->queue.cap, 0xFFFFFFF0u);
  S.comp_cap = s->comp_cap;
  S.cell_cap = s->cell_cap;
  S.front_cap = s->front_cap;
  Counters *cnt = s->counters.p;
  // phase 0: the search.  phase 1 (only after the host has seen ow_fallback == 1 in the counters of phase 0: a level too
  // wide for the shared-memory queue on a map too large for the spill instantiation): the global-memory wavefront redoes
  // the search, then a collection of what it found.  The fallback kernels used to follow the shared-memory kernel
  // unconditionally and return at once; but each is one CTA of 1024 threads that needs a whole free SM, and waited up to
  // 0.5 ms for one while the scoring launches filled them all.
  const uint64_t n_words = (uint64_t)S.wpr * c->H;
  const unsigned wide = 148 * 8;
  if (phase == 0) {
    FS_CUDA(c, cudaMemsetAsync(cnt, 0, sizeof(Counters), c->stream));
    FS_CUDA(c, cudaMemsetAsync(s->ftag.p, 0, s->front_cap, c->stream));
    const uint64_t mask_threads = (uint64_t)S.wpr * ((c->H + MASK_ROWS - 1) / MASK_ROWS);
    mask_kernel<<<(unsigned)((mask_threads + 255) / 256), 256, 0, c->stream>>>(S, s->bits.p, s->trav.p, s->ebits.p, s->tm.p, s->slot.p, s->mcell.p, cnt);
    c->launches += 1;
  }
  if (!s->have_ev) {
    for (auto &e : s->ev_bfs) FS_CUDA(c, cudaEventCreate(&e));
    for (auto &e : s->ev_part) FS_CUDA(c, cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    for (auto &e : s->ev_coll) FS_CUDA(c, cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    FS_CUDA(c, cudaEventCreateWithFlags(&s->ev_join, cudaEventDisableTiming));
    for (auto &e : s->ev_join_large) FS_CUDA(c, cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    for (auto &row : s->ev_chunk)
      for (auto &e : row) FS_CUDA(c, cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    for (auto &e : s->ev_score) FS_CUDA(c, cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    // the collections are short kernels on the critical path of their components; the scoring launches they feed are
    // long and fill every SM, so they run at the lowest priority and the collections take the slots that free up first
    int prio_lo = 0, prio_hi = 0;
    FS_CUDA(c, cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi));
    FS_CUDA(c, cudaStreamCreateWithPriority(&s->stream2, cudaStreamNonBlocking, prio_hi));
    for (auto &st : s->stream_large) FS_CUDA(c, cudaStreamCreateWithPriority(&st, cudaStreamNonBlocking, prio_hi));
    for (auto &st : s->score_stream) FS_CUDA(c, cudaStreamCreateWithPriority(&st, cudaStreamNonBlocking, prio_lo));
    if (const char *e = getenv("FITSLAM_OW_PARTS")) s->ow_parts = (uint32_t)std::min<long>(std::max<long>(atol(e), 1), OW_PARTS);
    s->have_ev = true;
  }
  // the union-find feeds the collections, not the outer wavefront: when collections run on the side stream, so does it
  const bool side = S.use_smem && s->ow_parts > 1 && phase == 0;
  cudaStream_t uf_stream = side ? s->stream2 : c->stream;
  if (side) {  // the side stream starts behind the counters reset and the mask
    FS_CUDA(c, cudaEventRecord(s->ev_part[0], c->stream));
    FS_CUDA(c, cudaStreamWaitEvent(s->stream2, s->ev_part[0], 0));
  }
  if (phase == 0) {
    uf_init_kernel<<<wide, 256, 0, uf_stream>>>(s->parent.p, s->rkey.p, s->rsize.p, s->rbox.p, cnt, S.mask_cap);
    uf_merge_kernel<<<wide, 256, 0, uf_stream>>>(S, s->bits.p, s->slot.p, s->mcell.p, s->parent.p, cnt);
    uf_compress_kernel<<<wide, 256, 0, uf_stream>>>(s->parent.p, s->rsize.p, s->rbox.p, s->mcell.p, c->W, cnt, S.mask_cap);
    c->launches += 3;
  }
  FS_CUDA(c, cudaEventRecord(s->ev_bfs[0], c->stream));
  // one collection: seeds of the components whose keys are final, their discovery order and offsets on stream `st`, their
  // inner wavefronts behind them; the large size class (the long poles: one warp walks a thin component level by level
  // for milliseconds) goes to `st_large`, so that the collections queued behind it are not held up
  auto collect = [&](cudaStream_t st, cudaStream_t st_large, int part) -> cudaError_t {
    const int coll = part < 0 ? OW_PARTS - 1 : part;
    touch_reduce_kernel<<<wide, 256, 0, st>>>(S, s->mcell.p, s->parent.p, s->claim.p, s->rkey.p, cnt, part);
    seed_collect_kernel<<<wide, 256, 0, st>>>(S, s->mcell.p, s->parent.p, s->claim.p, s->rkey.p, s->rsize.p, s->rbox.p, s->ckey.p,
                                              s->cseed.p, s->csize.p, s->cbox.p, cnt, part);
    comp_order_kernel<<<148 * 2, 256, 0, st>>>(S, s->ckey.p, s->cseed.p, s->csize.p, s->cbox.p, s->oseed.p, s->osize.p, s->obox.p, cnt);
    comp_scan_kernel<<<1, 1024, 0, st>>>(S, s->osize.p, s->ocell0.p, s->ofront0.p, cnt, coll);
    c->launches += 4;
    if (S.use_smem) {
      if (st_large != st) {
        cudaError_t e = cudaEventRecord(s->ev_coll[coll], st);
        if (e == cudaSuccess) e = cudaStreamWaitEvent(st_large, s->ev_coll[coll], 0);
        if (e != cudaSuccess) return e;
      }
      replay_line_kernel<<<148 * 4, 32, 0, st_large>>>(S, s->bits.p, s->slot.p, s->oseed.p, s->osize.p, s->ocell0.p, s->cells.p, cnt,
                                                       IW_SMALL + 1, 0xFFFFFFFFu, coll);
      replay_warp_kernel<<<148 * 16, 32, iw_smem(IW_SMALL, IW_BITS_SMALL), st>>>(
          S, s->bits.p, s->tm.p, s->oseed.p, s->osize.p, s->ocell0.p, s->obox.p, s->cells.p, cnt, 0, IW_SMALL, IW_BITS_SMALL, coll);
      c->launches += 2;
      if (part >= 0 && hook) {
        // the frontiers of the components just replayed are final: build them and hand them to the scorer, per size class
        // on the stream that replayed it (frontier.cu knows nothing about scoring: cycle.cu's hook launches it)
        const uint32_t *range = &cnt->coll_fr[coll][0];
        cudaStream_t sts[2] = {st, st_large};
        const uint32_t lo[2] = {0u, IW_SMALL + 1u}, hi[2] = {IW_SMALL, 0xFFFFFFFFu};
        for (int cls = 0; cls < 2; ++cls) {
          const uint32_t tag = 1u + 2u * (uint32_t)coll + (uint32_t)cls;
          chunk_kernel<<<148 * 2, 64, 0, sts[cls]>>>(S, s->osize.p, s->ocell0.p, s->ofront0.p, s->cells.p, c->frontiers.p,
                                                     s->centroids.p, s->ftag.p, s->fix_list.p, cnt, coll, lo[cls], hi[cls], tag);
          ++c->launches;
          ScoreSlice sl;
          sl.range = range;
          sl.cap = S.early_cap;
          sl.limit = S.early_limit;
          sl.tags = s->ftag.p;
          sl.tag = tag;
          cudaError_t e = cudaEventRecord(s->ev_chunk[coll][cls], sts[cls]);
          cudaStream_t sc = s->score_stream[cls == 0 ? 0 : 1 + coll % OW_LARGE_STREAMS];
          if (e == cudaSuccess) e = cudaStreamWaitEvent(sc, s->ev_chunk[coll][cls], 0);
          if (e != cudaSuccess) return e;
          if (hook->fn(c, sc, sl, hook->user) != FITSLAM_OK) return cudaErrorUnknown;
        }
      }
    }
    return cudaSuccess;
  };
  if (S.use_smem && phase == 0) {
    // bounding box of the search disc (+ a halo) clipped to the map, in 16 x 8-cell tiles of the padded arrays
    const OwLayout L = ow_layout(c->W, c->H);
    const uint32_t rcx = robot_cell % c->W, rcy = robot_cell / c->W;
    const uint32_t rad = (uint32_t)std::min<double>(S.dist_limit / c->res + 8.0, 1e9);
    const uint32_t bx0 = rcx > rad ? rcx - rad : 0u, bx1 = (uint32_t)std::min<uint64_t>((uint64_t)rcx + rad, c->W - 1);
    const uint32_t by0 = rcy > rad ? rcy - rad : 0u, by1 = (uint32_t)std::min<uint64_t>((uint64_t)rcy + rad, c->H - 1);
    // (a halo of OW_PAD cells around the box, so the padded coordinates of its corners are bx0 .. bx1 + 2 * OW_PAD)
    const uint32_t tx0 = bx0 >> 4, tx1 = std::min((bx1 + 2u * OW_PAD) >> 4, L.Wp / 16u - 1u);
    const uint32_t ty0 = by0 >> 3, ty1 = std::min((by1 + 2u * OW_PAD) >> 3, L.Hp / 8u - 1u);
    const uint32_t n_parts = s->ow_parts;
    // narrow levels (a few hundred cells) are bound by the latency of a level, which one CTA keeps lowest; wide ones by the
    // lookups per level, which the cluster spreads over its SMs.  Decided by the widest level of the last search (first
    // search on a map: by the size of the search disc)
    const uint32_t guess = s->last.bfs_levels ? s->last.widest_level : (uint32_t)std::min<double>(S.dist_limit / c->res, 1e9);
    const bool cluster = s->ow_cluster && guess >= s->ow_cluster_min;
    ow_prep_kernel<<<wide, 256, 0, c->stream>>>(S, L, s->ebits.p, s->bits.p, s->rk.p, tx0, ty0, tx1 - tx0 + 1, ty1 - ty0 + 1);
    ++c->launches;
    if (cluster) {
      cudaLaunchConfig_t cfg = {};
      cfg.gridDim = dim3(s->ow_cluster);
      cfg.blockDim = dim3(OWC_THREADS);
      cfg.dynamicSmemBytes = OWC_SMEM;
      cfg.stream = c->stream;
      cudaLaunchAttribute at[1];
      at[0].id = cudaLaunchAttributeClusterDimension;
      at[0].val.clusterDim.x = s->ow_cluster; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
      cfg.attrs = at;
      cfg.numAttrs = 1;
      FS_CUDA(c, cudaLaunchKernelEx(&cfg, ow_cluster_kernel, S, L, s->rk.p, s->claim.p, s->queue.p, cnt, n_parts));
    }
    s->last_ow_ctas = cluster ? (uint32_t)s->ow_cluster : 1u;
    // the search on the context stream; behind each published part but the last, a collection on the second stream
    if (cluster) {
    } else if (s->prefer_wide && S.ow_wide)
      ow_kernel<true><<<1, OW_THREADS, OW_SMEM, c->stream>>>(S, L, s->rk.p, s->claim.p, s->queue.p, cnt, n_parts);
    else
      ow_kernel<false><<<1, OW_THREADS, OW_SMEM, c->stream>>>(S, L, s->rk.p, s->claim.p, s->queue.p, cnt, n_parts);
    ++c->launches;
    FS_CUDA(c, cudaGetLastError());  // (the kernels below poll flags only this one sets)
    const auto host_t0 = std::chrono::steady_clock::now();
    for (uint32_t part = 0; part + 1 < n_parts; ++part) {
      wait_part_kernel<<<1, 1, 0, s->stream2>>>(cnt, part);
      ++c->launches;
      FS_CUDA(c, collect(s->stream2, s->stream_large[part % OW_LARGE_STREAMS], (int)part));
    }
    s->host_enqueue_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - host_t0).count();
    FS_CUDA(c, cudaEventRecord(s->ev_bfs[1], c->stream));
    if (n_parts > 1) {  // the kernels below rewrite or rescan what the collections' table kernels read
      FS_CUDA(c, cudaEventRecord(s->ev_join, s->stream2));
      FS_CUDA(c, cudaStreamWaitEvent(c->stream, s->ev_join, 0));
    }
    FS_CUDA(c, cudaEventRecord(s->ev_bfs[3], c->stream));
  } else {
    FS_CUDA(c, cudaEventRecord(s->ev_bfs[1], c->stream));
    FS_CUDA(c, cudaEventRecord(s->ev_bfs[3], c->stream));
    // the global-memory wavefront: the whole search when the shared-memory path is off, else (phase 1) for ow_fallback == 1
    rank_bfs_kernel<<<1, BFS_THREADS, 0, c->stream>>>(S, s->bits.p, s->trav.p, s->claim.p, s->queue.p, cnt);
    ++c->launches;
  }
  FS_CUDA(c, cudaEventRecord(s->ev_bfs[4], c->stream));
  FS_CUDA(c, collect(c->stream, c->stream, -1));
  if (!S.use_smem)  // (the global-memory path replays every component here)
    replay_kernel<<<148 * 8, REPLAY_THREADS, 0, c->stream>>>(S, s->bits.p, s->claim.p, s->oseed.p, s->osize.p, s->ocell0.p, s->cells.p, cnt);
  if (side)  // the large inner wavefronts still running on their streams: the frontiers not built yet may be theirs
    for (int i = 0; i < OW_LARGE_STREAMS; ++i) {
      FS_CUDA(c, cudaEventRecord(s->ev_join_large[i], s->stream_large[i]));
      FS_CUDA(c, cudaStreamWaitEvent(c->stream, s->ev_join_large[i], 0));
    }
  FS_CUDA(c, cudaEventRecord(s->ev_bfs[2], c->stream));
  chunk_kernel<<<148 * 4, 64, 0, c->stream>>>(S, s->osize.p, s->ocell0.p, s->ofront0.p, s->cells.p, c->frontiers.p,
                                               s->centroids.p, s->ftag.p, s->fix_list.p, cnt, -1, 0, 0xFFFFFFFFu, TAG_FINAL);
  cleanup_kernel<<<wide, 256, 0, c->stream>>>(S, s->queue.p, s->mcell.p, s->claim.p, cnt);
  c->launches += S.use_smem ? 2 : 3;  // (legacy replay,) chunk, cleanup
  FS_CUDA(c, cudaEventRecord(s->ev_bfs[5], c->stream));
  if (side && hook)
    for (int i = 0; i < 1 + OW_LARGE_STREAMS; ++i) {  // scores of the early batches: complete when the search is
      FS_CUDA(c, cudaEventRecord(s->ev_score[i], s->score_stream[i]));
      FS_CUDA(c, cudaStreamWaitEvent(c->stream, s->ev_score[i], 0));
    }
  FS_CUDA(c, cudaGetLastError());
  FS_CUDA(c, cudaEventRecord(s->ev_bfs[6], c->stream));
  FS_CUDA(c, cudaMemcpyAsync(host_counters, cnt, sizeof(Counters), cudaMemcpyDeviceToHost, c->stream));
  return FITSLAM_OK;
}

// widest level the shared-memory queue holds (FITSLAM_OW_QCAP lowers it in tests)
static uint32_t hc_qcap(const FrontierState *) {
  const char *e = getenv("FITSLAM_OW_QCAP");
  const long v = e ? atol(e) : (long)OW_QCAP;
  return (uint32_t)std::min<long>(std::max<long>(v, 1), (long)OW_QCAP);
}

// The median sort of the chunks the device would not decide (chunk_kernel), on the host and with the host's libm, exactly as
// the reference does it: getCentroidOfCells + SortByMedianFunctor + std::sort + the middle element
// (DEP/include/frontier_exploration/FrontierSearch.hpp:85-126, 156-184; DEP/src/FrontierSearch.cpp:161-165, 194-199).
static uint32_t host_median_cell(const fitslam_ctx *c, const uint32_t *cells, uint32_t n) {
  struct P { double first, second; uint32_t cell; };
  std::vector<P> v(n);
  const double res = c->res;
  double sumX = 0, sumY = 0;
  for (uint32_t k = 0; k < n; ++k) {
    const uint32_t my = cells[k] / c->W, mx = cells[k] - my * c->W;
    v[k] = P{c->ox + (mx + 0.5) * res, c->oy + (my + 0.5) * res, cells[k]};  // Costmap2D::mapToWorld
    sumX += v[k].first;
    sumY += v[k].second;
  }
  double centerX = static_cast<double>(sumX) / v.size(), centerY = static_cast<double>(sumY) / v.size();
  bool offset_centroid = false;
  double varX = 0, varY = 0;
  for (const P &p : v) {
    if (sqrt(pow(p.first - centerX, 2) + pow(p.second - centerY, 2)) < res * 3) offset_centroid = true;
    varX += std::abs(p.first - centerX);
    varY += std::abs(p.second - centerY);
  }
  const double distance_to_offset = res * 1.414 * 2;
  if (varX > varY && offset_centroid) centerY -= distance_to_offset;
  if (varX < varY && offset_centroid) centerX -= distance_to_offset;
  std::sort(v.begin(), v.end(), [centerX, centerY](const P &a, const P &b) {
    auto angle_a = atan2(a.second - centerY, a.first - centerX);
    if (angle_a < 0) angle_a = angle_a + (2 * M_PI);
    auto angle_b = atan2(b.second - centerY, b.first - centerX);
    if (angle_b < 0) angle_b = angle_b + (2 * M_PI);
    if (0 <= angle_a && angle_a <= M_PI / 2 && 3 * M_PI / 2 <= angle_b && angle_b <= 2 * M_PI) return false;
    if (0 <= angle_b && angle_b <= M_PI / 2 && 3 * M_PI / 2 <= angle_a && angle_a <= 2 * M_PI) return true;
    return angle_a < angle_b;
  });
  return v[static_cast<int>(v.size() / 2)].cell;
}

// after the search has drained: the flagged chunks come to the host, get their goal cell, and go back tagged TAG_FINAL
static fitslam_status fix_flagged_frontiers(fitslam_ctx *c, FrontierState *s, uint32_t n_flagged) {
  const uint32_t n = std::min(n_flagged, s->front_cap);
  if (n == 0) return FITSLAM_OK;
  FS_CUDA(c, s->fix_stage.reserve((size_t)n * FIX_WORDS));
  s->fix_host.resize((size_t)n * FIX_WORDS);
  fix_gather_kernel<<<std::min<uint32_t>((n * FIX_WORDS + 255) / 256, 148 * 8), 256, 0, c->stream>>>(s->fix_list.p, n, c->frontiers.p, s->cells.p,
                                                                                                  s->fix_stage.p);
  FS_CUDA(c, cudaMemcpyAsync(s->fix_host.data(), s->fix_stage.p, (size_t)n * FIX_WORDS * 4, cudaMemcpyDeviceToHost, c->stream));
  FS_CUDA(c, cudaStreamSynchronize(c->stream));
  std::vector<uint32_t> back(2 * (size_t)n);
  for (uint32_t e = 0; e < n; ++e) {
    const uint32_t *rec = s->fix_host.data() + (size_t)e * FIX_WORDS;
    back[2 * e] = rec[0];
    back[2 * e + 1] = host_median_cell(c, rec + 2, std::min<uint32_t>(rec[1], MAX_CHUNK));
  }
  FS_CUDA(c, cudaMemcpyAsync(s->fix_stage.p, back.data(), back.size() * 4, cudaMemcpyHostToDevice, c->stream));
  fix_apply_kernel<<<(n + 255) / 256, 256, 0, c->stream>>>(s->fix_stage.p, n, map_view(c), c->frontiers.p, s->ftag.p);
  FS_CUDA(c, cudaStreamSynchronize(c->stream));  // (`back` is pageable and goes out of scope)
  c->launches += 2;
  return FITSLAM_OK;
}

// FrontierSearch::searchFrom on the device; the frontier list stays resident in c->frontiers (c->n_frontiers entries).
uint32_t frontier_capacity(fitslam_ctx *c) {
  if (!c->have_map) return 0;
  FrontierState *s = fr_state(c);
  return ensure_frontier_buffers(c, s) == FITSLAM_OK ? s->front_cap : 0u;
}
const uint8_t *frontier_tags_device(fitslam_ctx *c) { return fr_state(c)->ftag.p; }

fitslam_status frontier_search_device(fitslam_ctx *c, double rx, double ry, const EarlyHook *hook) {
  if (!c->have_map) return fail(c, FITSLAM_E_NO_COSTMAP, "no costmap set");
  if (c->prm.max_frontier_cluster_size + 1 > MAX_CHUNK || c->prm.max_frontier_cluster_size + 1 <= c->prm.min_frontier_cluster_size)
    return fail(c, FITSLAM_E_INVALID, "search_frontiers: need min_frontier_cluster_size < max_frontier_cluster_size + 1 <= 64");
  FrontierState *s = fr_state(c);
  // Costmap2D::worldToMap (FrontierSearch.cpp:28-32)
  c->n_frontiers = 0;
  if (rx < c->ox || ry < c->oy) return fail(c, FITSLAM_E_ROBOT_OUT_OF_BOUNDS, "Robot out of costmap bounds, cannot search for frontiers");
  const unsigned int mx = static_cast<unsigned int>((rx - c->ox) / c->res), my = static_cast<unsigned int>((ry - c->oy) / c->res);
  if (!(mx < c->W && my < c->H)) return fail(c, FITSLAM_E_ROBOT_OUT_OF_BOUNDS, "Robot out of costmap bounds, cannot search for frontiers");
  fitslam_status st = ensure_pinned(c, sizeof(Counters));
  if (st != FITSLAM_OK) return st;
  bool grown = false;
  for (int attempt = 0; attempt < 10; ++attempt) {
    st = ensure_frontier_buffers(c, s);
    if (st != FITSLAM_OK) return st;
    Counters *hc = reinterpret_cast<Counters *>(c->pinned);
    // (once the lists have grown, the caller's buffers, sized for the first attempt, may be too small: no early scoring then)
    st = enqueue_search(c, s, my * c->W + mx, rx, ry, hc, grown ? nullptr : hook, 0);
    if (st != FITSLAM_OK) return st;
    FS_CUDA(c, cudaStreamSynchronize(c->stream));
    if (hc->ow_fallback == 2u && !hc->overflow && !s->prefer_wide) {
      s->prefer_wide = true;  // a level too wide for the shared-memory queue: again, with the instantiation that spills
      continue;
    }
    if (hc->ow_fallback == 1u && !hc->overflow) {  // ... on a map too large for it: the global-memory wavefront, then the rest
      st = enqueue_search(c, s, my * c->W + mx, rx, ry, hc, nullptr, 1);
      if (st != FITSLAM_OK) return st;
      FS_CUDA(c, cudaStreamSynchronize(c->stream));
    }
    if (s->prefer_wide && !hc->overflow && hc->widest_level <= hc_qcap(s)) s->prefer_wide = false;  // narrow levels again
    s->last = *hc;
    if (getenv("FITSLAM_PROF"))
      fprintf(stderr, "outer wavefront: levels %u expanded %u widest %u passes %llu cycles %llu\n", hc->bfs_levels, hc->expanded,
              hc->widest_level, hc->prof[4], hc->prof[5]);
    if (getenv("FITSLAM_PROF")) {
      float t[6] = {0, 0, 0, 0, 0, 0};
      const int seq[7] = {0, 1, 3, 4, 2, 5, 6};
      for (int i = 0; i < 6; ++i) cudaEventElapsedTime(&t[i], s->ev_bfs[seq[i]], s->ev_bfs[seq[i + 1]]);
      fprintf(stderr, "context stream, ms: outer wavefront %.3f | wait for the collections %.3f | fallback kernels %.3f | last collection + "
                      "large inner wavefronts %.3f | frontiers of the rest, cleanup %.3f | wait for the early scoring %.3f\n",
              t[0], t[1], t[2], t[3], t[4], t[5]);
      fprintf(stderr, "host: %.3f ms to enqueue the work behind the parts of the outer wavefront\n", s->host_enqueue_ms);
      fprintf(stderr, "collections (cells dequeued | frontiers):");
      for (int k = 0; k < OW_PARTS; ++k) fprintf(stderr, " %u|%u", hc->part_end[k], hc->coll_fr[k][1] - hc->coll_fr[k][0]);
      fprintf(stderr, "\n");
    }
    {  // the wavefront kernels alone: the parts, and the fallback kernels behind the join with the second stream
      float a = 0, b = 0;
      cudaEventElapsedTime(&a, s->ev_bfs[0], s->ev_bfs[1]);
      cudaEventElapsedTime(&b, s->ev_bfs[3], s->ev_bfs[4]);
      s->bfs_ms = a + b;
    }
    if (!hc->overflow) {
      c->n_frontiers = hc->n_frontiers;
      return fix_flagged_frontiers(c, s, s->last.pad0[0]);
    }
    // a list was too small: grow to what this run counted (plus slack) and search again
    grown = true;
    if (hc->overflow & 1u) s->mask_cap = hc->n_mask + hc->n_mask / 8 + 1024;
    if (hc->overflow & 4u) s->comp_cap = hc->n_comp + hc->n_comp / 8 + 1024;
    if (hc->overflow & 8u) s->cell_cap = hc->n_cells + hc->n_cells / 8 + 1024;
    if (hc->overflow & 16u) s->front_cap = hc->n_frontiers + hc->n_frontiers / 8 + 1024;
    s->cell_cap = std::max(s->cell_cap, s->mask_cap);
    if (hc->overflow & 2u) return fail(c, FITSLAM_E_INVALID, "search_frontiers: wavefront queue overflow");
    if (hc->overflow & 32u) return fail(c, FITSLAM_E_CUDA, "search_frontiers: the outer wavefront never published its progress");
  }
  return fail(c, FITSLAM_E_INVALID, "search_frontiers: buffers still too small after repeated attempts");
}

const uint32_t *frontier_cells_device(fitslam_ctx *c, uint32_t *n_cells, uint32_t *n_comp, uint32_t *levels, uint32_t *ambiguous) {
  FrontierState *s = fr_state(c);
  if (n_cells) *n_cells = s->last.n_cells;
  if (n_comp) *n_comp = s->last.n_comp;
  if (levels) *levels = s->last.bfs_levels;
  if (ambiguous) *ambiguous = s->last.ambiguous;
  c->last_bfs_ms = s->bfs_ms;
  return s->cells.p;
}

void frontier_wavefront_device(fitslam_ctx *c, uint32_t *widest_level, uint32_t *ctas) {
  FrontierState *s = fr_state(c);
  *widest_level = s->last.widest_level;
  *ctas = s->last_ow_ctas;
}

}  // namespace fitslam

using namespace fitslam;

extern "C" {

fitslam_status fitslam_set_search_distance(fitslam_ctx *c, double max_frontier_distance) {
  if (!c || !(max_frontier_distance >= 0.0)) return fail(c, FITSLAM_E_INVALID, "set_search_distance: bad argument");
  fr_state(c)->search_distance = max_frontier_distance;  // incrementSearchDistance / resetSearchDistance (FrontierSearch.hpp:39-52)
  return FITSLAM_OK;
}

fitslam_status fitslam_search_frontiers(fitslam_ctx *c, double robot_x, double robot_y, fitslam_frontier *out,
                                        uint32_t *n_frontiers, uint32_t *every_cells, uint32_t *n_cells) {
  if (!c || !n_frontiers) return fail(c, FITSLAM_E_INVALID, "search_frontiers: null argument");
  cudaSetDevice(c->device);
  const uint32_t cap = *n_frontiers;
  const uint32_t ccap = n_cells ? *n_cells : 0;
  *n_frontiers = 0;
  if (n_cells) *n_cells = 0;
  fitslam_status st = frontier_search_device(c, robot_x, robot_y);
  if (st != FITSLAM_OK) return st;
  FrontierState *s = c->fs;
  const uint32_t F = s->last.n_frontiers, NC = s->last.n_cells;
  *n_frontiers = F;
  if (n_cells) *n_cells = NC;
  st = FITSLAM_OK;
  if (out) {
    if (F > cap) st = fail(c, FITSLAM_E_CAPACITY, "search_frontiers: frontier buffer too small");
    const uint32_t m = std::min(F, cap);
    if (m) FS_CUDA(c, cudaMemcpyAsync(out, c->frontiers.p, (size_t)m * sizeof(fitslam_frontier), cudaMemcpyDefault, c->stream));
  }
  if (every_cells) {
    if (NC > ccap) st = fail(c, FITSLAM_E_CAPACITY, "search_frontiers: cell buffer too small");
    const uint32_t m = std::min(NC, ccap);
    if (m) FS_CUDA(c, cudaMemcpyAsync(every_cells, s->cells.p, (size_t)m * 4, cudaMemcpyDefault, c->stream));
  }
  FS_CUDA(c, cudaStreamSynchronize(c->stream));
  return st;
}

}  // extern "C"
