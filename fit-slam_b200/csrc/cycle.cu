// cycle.cu -- one whole exploration cycle, enqueued back to back on the context stream.
//
// The order is the behaviour tree's (FS2/fit_slam2/xml/active_slam_exploration.xml:5-47, DEP/src/ExplorationBT.cpp:166-637):
//   SearchForFrontiers  -> FrontierSearch::searchFrom                      (frontier.cu)
//   ProcessFrontierCosts -> setArrivalInformationForFrontier per frontier  (raycast.cu)
//                         -> setPlanForFrontierEuclidean                   (select.cu)
//                         -> U1 utility block                              (select.cu)
//   OptimizeFullPath     -> getFilteredFrontiersN                          (select.cu)
//   EvaluateFisherInformation -> isPoseSafe, here for every candidate pose (fisher.cu)
// fitslam_cycle_begin runs everything that shards across GPUs (extraction is replicated, arrival and Fisher are
// scored for this rank's block of frontiers) and leaves 16-byte score records laid out for an in-place all-gather;
// fitslam_cycle_end unpacks the merged records and runs the global steps (plan, normalise, select).
#include "ctx.h"

#include <algorithm>
#include <cmath>
#include <cstring>
#include <vector>

namespace fitslam {

struct CycleState {
  DevBuf<fitslam_pose> poses;
  DevBuf<float> fisher;        // per pose
  DevBuf<float> fisher_best;   // per frontier: Fisher information of its best-yaw pose (cycle_fetch)
  DevBuf<uint8_t> fisher_safe;
  DevBuf<uint32_t> fisher_nvis;
  DevBuf<double> yaw_q;        // per yaw window: qz, qw of orientationAroundZAxis(window * delta_theta + fov / 2)
  // atan2(goal - robot) per frontier for the path heading, evaluated by the host's libm while the scoring kernels run
  DevBuf<double> goal_yaw;
  fitslam_frontier *h_fr = nullptr;  // pinned
  double *h_yaw = nullptr;           // pinned
  size_t h_cap = 0;
  cudaEvent_t ev_fr = nullptr;
  // Work lists of the Fisher launches (ScoreSlice::work_*): WORK_LAUNCHES counters, then one list of WORK_LIST frontier indices
  // per early launch (index = the launch's tag); work_final: counter + list of the launch after the search.  hist[] = how many
  // frontiers each early launch had in the last cycle: the next cycle sizes its grids from it (the host cannot know the
  // counts of a search that is still running; what a grid does not cover is scored after the search).
  static constexpr uint32_t WORK_LAUNCHES = 128, WORK_LIST = 8192;
  DevBuf<uint32_t> work, work_final;
  uint32_t *h_work = nullptr;  // pinned: WORK_LAUNCHES + 1
  uint32_t hist[WORK_LAUNCHES] = {};
  bool have_hist = false, hist_pending = false;
  uint32_t yaw_windows = 0;
  double yaw_dt = 0, yaw_fov = 0;
  cudaEvent_t ev[8] = {};
  bool have_events = false;
  // state carried from cycle_begin to cycle_end
  bool open = false;
  double rx = 0, ry = 0, ryaw = 0;
  int fisher_mode = 0;
  uint32_t n_poses = 0;
  fitslam_cycle_result res{};
};

static CycleState *cstate(fitslam_ctx *c) {
  if (!c->cy) c->cy = new CycleState();
  return c->cy;
}

void destroy_cycle_state(fitslam_ctx *c) {
  if (!c->cy) return;
  CycleState *s = c->cy;
  s->poses.release(); s->fisher.release(); s->fisher_best.release(); s->fisher_safe.release(); s->fisher_nvis.release(); s->yaw_q.release();
  s->goal_yaw.release();
  s->work.release(); s->work_final.release();
  if (s->h_work) cudaFreeHost(s->h_work);
  if (s->h_fr) cudaFreeHost(s->h_fr);
  if (s->h_yaw) cudaFreeHost(s->h_yaw);
  if (s->have_events) {
    for (auto &e : s->ev) cudaEventDestroy(e);
    cudaEventDestroy(s->ev_fr);
  }
  delete s;
  c->cy = nullptr;
}

// Candidate poses at the frontier goals: position = goal point, z = candidate_pose_z, yaw = window *
// delta_theta + fov / 2 for every window (ppf = n_windows) or for the best window only (ppf = 1), i.e. what
// setGoalOrientation produces (DEP/src/Frontier.cpp:50-56).  A frontier whose ray fan failed has orientation 0.
__global__ void make_poses_kernel(const fitslam_frontier *__restrict__ fr, const fitslam_arrival *__restrict__ arr,
                                  const double *__restrict__ yaw_q, const ScoreSlice slice, uint32_t ppf, double z,
                                  fitslam_pose *poses) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  const uint32_t slot = i / ppf, w = i - slot * ppf;
  uint32_t f = 0;
  if (!slice_frontier(slice, slot, f)) return;
  if (slice.work_count && w == 0) {
    const uint32_t k = atomicAdd(slice.work_count, 1u);
    if (k < slice.work_cap) slice.work_list[k] = f;
    else if (slice.retag) slice.retag[f] = (uint8_t)TAG_FINAL;  // more frontiers than the launch was sized for: scored after the search
  }
  const uint32_t win = ppf == 1 ? (uint32_t)arr[f].max_index : w;
  fitslam_pose p;
  p.x = fr[f].goal_x;
  p.y = fr[f].goal_y;
  p.z = z;
  p.qx = 0.0;
  p.qy = 0.0;
  if (ppf == 1 && arr[f].ray_failed) {
    p.qz = 0.0;
    p.qw = 1.0;
  } else {
    p.qz = yaw_q[2 * win];
    p.qw = yaw_q[2 * win + 1];
  }
  poses[(size_t)f * ppf + w] = p;
}

// Arrival information, candidate poses and Fisher information for one slice of the resident frontier list, on `stream`.
// The cycle calls it for the rank's block of frontiers; the frontier search calls it (through the hook) for every batch
// of frontiers that is final while the outer wavefront is still running.
struct ScoreCfg {
  CycleState *s;
  double poly[4];
  uint32_t ppf;  // 0: no Fisher stage
};
static fitslam_status score_slice(fitslam_ctx *c, cudaStream_t stream, const ScoreSlice &sl_in, void *user) {
  const ScoreCfg *cfg = static_cast<const ScoreCfg *>(user);
  CycleState *s = cfg->s;
  ScoreSlice sl = sl_in;  // this rank's share of the slice
  sl.mod = (uint32_t)c->nranks;
  sl.rem = (uint32_t)c->rank;
  fitslam_status st = raycast_score_slice(c, stream, c->frontiers.p, cfg->poly, c->arrival.p, c->records.p, sl);
  if (st != FITSLAM_OK || cfg->ppf == 0) return st;
  const uint64_t slots = sl.range ? sl.cap : sl.count;
  if (slots == 0) return FITSLAM_OK;
  if (sl.range && sl.tags && sl.tag < CycleState::WORK_LAUNCHES && s->work.p) {
    // an early launch: its Fisher grid is sized from what this launch found in the last cycle (+ 25 %), not from the cap
    sl.work_count = s->work.p + sl.tag;
    sl.work_list = s->work.p + CycleState::WORK_LAUNCHES + (size_t)sl.tag * CycleState::WORK_LIST;
    sl.work_cap = std::min<uint32_t>(sl.cap, CycleState::WORK_LIST);
    if (s->have_hist) sl.work_cap = std::min<uint32_t>(sl.work_cap, s->hist[sl.tag] + s->hist[sl.tag] / 4u + 8u);
    sl.retag = const_cast<uint8_t *>(sl.tags);
  }
  const uint64_t np = slots * cfg->ppf;
  make_poses_kernel<<<(unsigned)((np + 255) / 256), 256, 0, stream>>>(c->frontiers.p, c->arrival.p, s->yaw_q.p, sl, cfg->ppf,
                                                                       c->prm.candidate_pose_z, s->poses.p);
  ++c->launches;
  return fisher_score_slice(c, stream, s->poses.p, cfg->ppf, sl, s->fisher.p, s->fisher_safe.p, s->fisher_nvis.p, c->records.p,
                            c->arrival.p);
}

// merged 16-byte records -> fitslam_arrival (the part of the record that other ranks scored)
__global__ void unpack_records_kernel(const ScoreRecord *__restrict__ rec, fitslam_arrival *arr, uint32_t n, double dt,
                                      double half_fov) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const ScoreRecord r = rec[i];
  fitslam_arrival a;
  memset(&a, 0, sizeof(a));
  a.information = r.information;
  a.max_index = r.max_index;
  a.achievable = r.flags & 1u;
  a.footprint_in_lethal = (r.flags >> 1) & 1u;
  a.ray_failed = (r.flags >> 2) & 1u;
  a.theta_s_star = a.ray_failed ? 0.0 : ((double)r.max_index * dt) + half_fov;
  arr[i] = a;
}

__global__ void gather_fisher_kernel(const ScoreRecord *__restrict__ rec, float *out, uint32_t n) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = rec[i].fisher;
}

static fitslam_status ensure_yaw_table(fitslam_ctx *c, CycleState *s) {
  const uint32_t nw = c->shape.n_windows;
  if (s->yaw_windows == nw && s->yaw_dt == c->prm.delta_theta && s->yaw_fov == c->prm.camera_fov) return FITSLAM_OK;
  std::vector<double> q(2 * (size_t)nw);
  for (uint32_t w = 0; w < nw; ++w) {
    const double theta = (w * c->prm.delta_theta) + (c->prm.camera_fov / 2);  // DEP/src/CostCalculator.cpp:119
    // tf2::Quaternion::setRPY(0, 0, theta) (nav2_util::geometry_utils::orientationAroundZAxis)
    const double halfYaw = theta * 0.5, cosYaw = cos(halfYaw), sinYaw = sin(halfYaw);
    const double cosPitch = cos(0.0), sinPitch = sin(0.0), cosRoll = cos(0.0), sinRoll = sin(0.0);
    q[2 * w] = cosRoll * cosPitch * sinYaw - sinRoll * sinPitch * cosYaw;
    q[2 * w + 1] = cosRoll * cosPitch * cosYaw + sinRoll * sinPitch * sinYaw;
  }
  FS_CUDA(c, s->yaw_q.reserve(q.size()));
  FS_CUDA(c, cudaMemcpyAsync(s->yaw_q.p, q.data(), q.size() * sizeof(double), cudaMemcpyHostToDevice, c->stream));
  FS_CUDA(c, cudaStreamSynchronize(c->stream));
  s->yaw_windows = nw;
  s->yaw_dt = c->prm.delta_theta;
  s->yaw_fov = c->prm.camera_fov;
  return FITSLAM_OK;
}

}  // namespace fitslam

using namespace fitslam;

extern "C" {

fitslam_status fitslam_cycle_begin(fitslam_ctx *c, double rx, double ry, double ryaw, const double poly[4], int fisher_mode,
                                   uint32_t *n_frontiers) {
  if (!c) return FITSLAM_E_INVALID;
  if (n_frontiers) *n_frontiers = 0;
  if (!c->have_map) return fail(c, FITSLAM_E_NO_COSTMAP, "no costmap set");
  if (!poly) return fail(c, FITSLAM_E_NO_POLYGON, "Frontier cannot be selected, no polygon.");
  if (fisher_mode < FITSLAM_FISHER_OFF || fisher_mode > FITSLAM_FISHER_ALL_YAWS) return fail(c, FITSLAM_E_INVALID, "cycle: bad fisher_mode");
  cudaSetDevice(c->device);
  CycleState *s = cstate(c);
  s->open = false;
  if (!s->have_events) {
    for (auto &e : s->ev) FS_CUDA(c, cudaEventCreate(&e));
    FS_CUDA(c, cudaEventCreateWithFlags(&s->ev_fr, cudaEventDisableTiming));
    s->have_events = true;
  }
  if (fisher_mode != FITSLAM_FISHER_OFF && !fisher_ready(c))
    return fail(c, c->fi ? FITSLAM_E_NO_MAP_POINTS : FITSLAM_E_NO_LOOKUP_TABLE, "cycle: Fisher stage needs a lookup table and map points");
  memset(&s->res, 0, sizeof(s->res));
  s->res.selection.closest_global = -1;
  fitslam_status st = raycast_limits(c);  // once per calculator (DEP/src/CostCalculator.cpp:125-129)
  if (st != FITSLAM_OK) return st;
  const uint32_t ppf = fisher_mode == FITSLAM_FISHER_OFF ? 0u : (fisher_mode == FITSLAM_FISHER_ALL_YAWS ? c->shape.n_windows : 1u);
  if (ppf) {
    st = ensure_yaw_table(c, s);
    if (st != FITSLAM_OK) return st;
  }
  ScoreCfg cfg;
  cfg.s = s;
  for (int i = 0; i < 4; ++i) cfg.poly[i] = poly[i];
  cfg.ppf = ppf;
  // The frontiers are scored as the search produces them: buffers for as many frontiers as the search can return are laid
  // out first, and the search calls score_slice on its side streams.  With several ranks each scores the frontiers f with
  // f % nranks == rank into zeroed buffers (the caller's all-reduce merges the records).
  EarlyHook hook{score_slice, &cfg, 0};
  const bool early = !getenv("FITSLAM_NO_EARLY_SCORING");
  uint32_t cap = 0;
  if (early) {
    // (not for as many as the list can hold, N / 64 frontiers: 4 x what the last search found, at least 64 k; frontiers
    // beyond that are left to the launches after the search)
    cap = frontier_capacity(c);
    if (cap == 0) return fail(c, FITSLAM_E_CUDA, "cycle: frontier buffers");
    cap = std::min<uint32_t>(cap, std::max<uint32_t>(65536u, 4u * c->n_frontiers));
    hook.limit = cap;
    FS_CUDA(c, c->arrival.reserve(cap));
    FS_CUDA(c, c->records.reserve(cap));
    if (ppf) {
      FS_CUDA(c, s->poses.reserve((size_t)cap * ppf));
      FS_CUDA(c, s->fisher.reserve((size_t)cap * ppf));
      FS_CUDA(c, s->fisher_safe.reserve((size_t)cap * ppf));
      FS_CUDA(c, s->fisher_nvis.reserve((size_t)cap * ppf));
    }
    if (c->nranks > 1) {  // what the other ranks score stays zero
      FS_CUDA(c, cudaMemsetAsync(c->arrival.p, 0, (size_t)cap * sizeof(fitslam_arrival), c->stream));
      FS_CUDA(c, cudaMemsetAsync(c->records.p, 0, (size_t)cap * sizeof(ScoreRecord), c->stream));
      if (ppf) FS_CUDA(c, cudaMemsetAsync(s->fisher.p, 0, (size_t)cap * ppf * sizeof(float), c->stream));
    }
  }
  if (early && ppf) {
    FS_CUDA(c, s->work.reserve(CycleState::WORK_LAUNCHES * (size_t)(1 + CycleState::WORK_LIST)));
    FS_CUDA(c, cudaMemsetAsync(s->work.p, 0, CycleState::WORK_LAUNCHES * sizeof(uint32_t), c->stream));
    if (!s->h_work) FS_CUDA(c, cudaMallocHost((void **)&s->h_work, (CycleState::WORK_LAUNCHES + 1) * sizeof(uint32_t)));
  }
  s->hist_pending = false;
  FS_CUDA(c, cudaEventRecord(s->ev[0], c->stream));
  st = frontier_search_device(c, rx, ry, early ? &hook : nullptr);  // one host sync inside: the list sizes decide the next launches
  if (st != FITSLAM_OK) return st;
  if (early && ppf) {  // (the search has waited for its scoring streams: the counts are final; read at the end of the cycle)
    FS_CUDA(c, cudaMemcpyAsync(s->h_work, s->work.p, CycleState::WORK_LAUNCHES * sizeof(uint32_t), cudaMemcpyDeviceToHost, c->stream));
    s->hist_pending = true;
  }
  FS_CUDA(c, cudaEventRecord(s->ev[1], c->stream));
  const uint32_t F = c->n_frontiers;
  uint32_t n_cells = 0, n_comp = 0, levels = 0, ambiguous = 0;
  frontier_cells_device(c, &n_cells, &n_comp, &levels, &ambiguous);
  s->res.n_frontiers = F;
  s->res.n_frontier_cells = n_cells;
  s->res.n_components = n_comp;
  s->res.bfs_levels = levels;
  s->res.ambiguous_sorts = ambiguous;
  frontier_wavefront_device(c, &s->res.widest_level, &s->res.wavefront_ctas);
  if (n_frontiers) *n_frontiers = F;
  if (F == 0) return fail(c, FITSLAM_E_NO_FRONTIERS, "No frontiers found from frontier search.");
  // the goal points come to the host now (70 KB at C2), ahead of the scoring launches below
  if (s->h_cap < F) {
    if (s->h_fr) cudaFreeHost(s->h_fr);
    if (s->h_yaw) cudaFreeHost(s->h_yaw);
    s->h_fr = nullptr; s->h_yaw = nullptr; s->h_cap = 0;
    const size_t want = (size_t)F + F / 4 + 64;
    FS_CUDA(c, cudaMallocHost((void **)&s->h_fr, want * sizeof(fitslam_frontier)));
    FS_CUDA(c, cudaMallocHost((void **)&s->h_yaw, want * sizeof(double)));
    s->h_cap = want;
  }
  FS_CUDA(c, s->goal_yaw.reserve(F));
  FS_CUDA(c, cudaMemcpyAsync(s->h_fr, c->frontiers.p, (size_t)F * sizeof(fitslam_frontier), cudaMemcpyDeviceToHost, c->stream));
  FS_CUDA(c, cudaEventRecord(s->ev_fr, c->stream));
  // ---- arrival and Fisher information for this rank's share of the frontiers: the ones not scored during the search
  const uint64_t n_poses = (uint64_t)F * ppf;
  if (n_poses >= (1ull << 31)) return fail(c, FITSLAM_E_INVALID, "cycle: too many candidate poses");
  const bool scored_early = early && F <= cap;  // (a search that had to grow its lists was repeated without the hook)
  if (!scored_early) {
    FS_CUDA(c, c->arrival.reserve(F));
    FS_CUDA(c, c->records.reserve(F));
    if (ppf) {
      FS_CUDA(c, s->poses.reserve(n_poses));
      FS_CUDA(c, s->fisher.reserve(n_poses));
      FS_CUDA(c, s->fisher_safe.reserve(n_poses));
      FS_CUDA(c, s->fisher_nvis.reserve(n_poses));
    }
    if (c->nranks > 1) {
      FS_CUDA(c, cudaMemsetAsync(c->arrival.p, 0, (size_t)F * sizeof(fitslam_arrival), c->stream));
      FS_CUDA(c, cudaMemsetAsync(c->records.p, 0, (size_t)F * sizeof(ScoreRecord), c->stream));
      if (ppf) FS_CUDA(c, cudaMemsetAsync(s->fisher.p, 0, n_poses * sizeof(float), c->stream));
    }
  }
  c->records_rows = F;
  ScoreSlice rest;
  rest.first = 0;
  rest.count = F;
  rest.mod = (uint32_t)c->nranks;
  rest.rem = (uint32_t)c->rank;
  // (more frontiers than the early buffers hold: they were grown above, so every frontier is scored now, whatever its tag says)
  rest.tags = scored_early ? frontier_tags_device(c) : nullptr;
  rest.tag = TAG_FINAL;
  // two launches so that the stage times keep their meaning: arrival, then poses + Fisher
  ScoreCfg arrival_only = cfg;
  arrival_only.ppf = 0;
  st = score_slice(c, c->stream, rest, &arrival_only);
  if (st != FITSLAM_OK) return st;
  FS_CUDA(c, cudaEventRecord(s->ev[2], c->stream));
  s->res.n_candidates = F * c->shape.n_windows;
  s->n_poses = 0;
  if (ppf && rest.count) {
    // the frontiers left for this launch (this rank's, not scored during the search) go into a list, and the host reads
    // their number: a grid over all F frontiers would start up to F * ppf CTAs that find nothing to do
    FS_CUDA(c, s->work_final.reserve((size_t)F + 1));
    if (!s->h_work) FS_CUDA(c, cudaMallocHost((void **)&s->h_work, (CycleState::WORK_LAUNCHES + 1) * sizeof(uint32_t)));
    FS_CUDA(c, cudaMemsetAsync(s->work_final.p, 0, sizeof(uint32_t), c->stream));
    rest.work_count = s->work_final.p;
    rest.work_list = s->work_final.p + 1;
    rest.work_cap = F;
    make_poses_kernel<<<(unsigned)(((uint64_t)rest.count * ppf + 255) / 256), 256, 0, c->stream>>>(
        c->frontiers.p, c->arrival.p, s->yaw_q.p, rest, ppf, c->prm.candidate_pose_z, s->poses.p);
    ++c->launches;
    uint32_t *h_n = s->h_work + CycleState::WORK_LAUNCHES;
    FS_CUDA(c, cudaMemcpyAsync(h_n, s->work_final.p, sizeof(uint32_t), cudaMemcpyDeviceToHost, c->stream));
    FS_CUDA(c, cudaStreamSynchronize(c->stream));
    if (s->hist_pending) {  // (the same synchronisation brought the early launches' counts)
      memcpy(s->hist, s->h_work, sizeof(s->hist));
      s->have_hist = true;
      s->hist_pending = false;
    }
    rest.work_cap = std::min<uint32_t>(*h_n, F);
    st = fisher_score_slice(c, c->stream, s->poses.p, ppf, rest, s->fisher.p, s->fisher_safe.p, s->fisher_nvis.p, c->records.p,
                            c->arrival.p);
    if (st != FITSLAM_OK) return st;
  }
  if (ppf) {
    s->n_poses = (uint32_t)n_poses;
    s->res.n_fisher_poses = (uint32_t)n_poses;
  }
  FS_CUDA(c, cudaEventRecord(s->ev[3], c->stream));
  FS_CUDA(c, cudaGetLastError());
  // ... and while those kernels run: goal_yaw = atan2(goal - robot) with the reference's own libm (CostCalculator.cpp:469)
  FS_CUDA(c, cudaEventSynchronize(s->ev_fr));
  for (uint32_t i = 0; i < F; ++i) s->h_yaw[i] = atan2(s->h_fr[i].goal_y - ry, s->h_fr[i].goal_x - rx);
  FS_CUDA(c, cudaMemcpyAsync(s->goal_yaw.p, s->h_yaw, (size_t)F * sizeof(double), cudaMemcpyHostToDevice, c->stream));
  s->rx = rx; s->ry = ry; s->ryaw = ryaw;
  s->fisher_mode = fisher_mode;
  s->open = true;
  return FITSLAM_OK;
}

fitslam_status fitslam_cycle_end(fitslam_ctx *c, fitslam_cycle_result *result) {
  if (!c) return FITSLAM_E_INVALID;
  CycleState *s = cstate(c);
  if (!s->open) return fail(c, FITSLAM_E_INVALID, "cycle_end without a successful cycle_begin");
  s->open = false;
  cudaSetDevice(c->device);
  const uint32_t F = c->n_frontiers;
  FS_CUDA(c, cudaEventRecord(s->ev[4], c->stream));
  if (c->nranks > 1) {  // the caller has all-gathered the records: rebuild every frontier's arrival fields from them
    unpack_records_kernel<<<(F + 255) / 256, 256, 0, c->stream>>>(c->records.p, c->arrival.p, F, c->prm.delta_theta,
                                                                  c->prm.camera_fov / 2);
    ++c->launches;
  }
  fitslam_status st = plan_euclidean_device(c, F, c->frontiers.p, c->arrival.p, s->goal_yaw.p, s->rx, s->ry, s->ryaw);
  if (st != FITSLAM_OK) return st;
  FS_CUDA(c, cudaEventRecord(s->ev[5], c->stream));
  SelectState *ss = sstate(c);
  st = select_device(c, F, c->arrival.p, ss->ach.p, nullptr, ss->len.p, ss->head.p);
  if (st != FITSLAM_OK) return st;
  FS_CUDA(c, cudaEventRecord(s->ev[6], c->stream));
  st = ensure_pinned(c, sizeof(SelectMeta));
  if (st != FITSLAM_OK) return st;
  SelectMeta *m = reinterpret_cast<SelectMeta *>(c->pinned);
  FS_CUDA(c, cudaMemcpyAsync(m, ss->meta.p, sizeof(SelectMeta), cudaMemcpyDeviceToHost, c->stream));
  FS_CUDA(c, cudaStreamSynchronize(c->stream));
  s->res.selection = m->sel;
  float ms = 0;
  const int pairs[7][2] = {{0, 1}, {1, 2}, {4, 5}, {2, 3}, {5, 6}, {0, 3}, {4, 6}};
  for (int i = 0; i < 7; ++i) {
    FS_CUDA(c, cudaEventElapsedTime(&ms, s->ev[pairs[i][0]], s->ev[pairs[i][1]]));
    s->res.stage_ms[i] = ms;  // extract, arrival, distance, fisher, normalise+select, begin total, end total
  }
  s->res.stage_ms[7] = c->last_bfs_ms;  // the outer wavefront kernel alone (inside stage 0)
  if (result) *result = s->res;
  if (m->err_key != 0xffffffffu) {
    const bool arrival = (m->err_key & 1u) == 0;
    return fail(c, arrival ? FITSLAM_E_ARRIVAL_UTILITY : FITSLAM_E_COST_OUT_OF_BOUNDS,
                std::string(arrival ? "ARRIVAL UTILITY ERROR! MORE THAN 1" : "Cost out of bounds") + " (frontier " +
                    std::to_string(m->err_key >> 1) + ")");
  }
  return FITSLAM_OK;
}

fitslam_status fitslam_cycle(fitslam_ctx *c, double rx, double ry, double ryaw, const double poly[4], int fisher_mode,
                             fitslam_cycle_result *result) {
  uint32_t F = 0;
  fitslam_status st = fitslam_cycle_begin(c, rx, ry, ryaw, poly, fisher_mode, &F);
  if (st != FITSLAM_OK) {
    if (result && c && c->cy) *result = c->cy->res;
    return st;
  }
  st = fitslam_cycle_merge(c);  // (nothing on one GPU or without a communicator)
  if (st != FITSLAM_OK) return st;
  return fitslam_cycle_end(c, result);
}

fitslam_status fitslam_cycle_fetch(fitslam_ctx *c, fitslam_frontier *frontiers, fitslam_arrival *arrival, double *path_length_m,
                                   double *path_heading, double *weighted_cost, float *fisher_info, uint32_t *sorted,
                                   uint32_t capacity) {
  if (!c || !c->cy) return fail(c, FITSLAM_E_INVALID, "cycle_fetch before fitslam_cycle");
  cudaSetDevice(c->device);
  CycleState *s = c->cy;
  SelectState *ss = sstate(c);
  const uint32_t F = c->n_frontiers;
  if (capacity < F) return fail(c, FITSLAM_E_CAPACITY, "cycle_fetch: buffers too small");
  if (F == 0) return FITSLAM_OK;
  if (frontiers) FS_CUDA(c, cudaMemcpyAsync(frontiers, c->frontiers.p, (size_t)F * sizeof(fitslam_frontier), cudaMemcpyDefault, c->stream));
  if (arrival) FS_CUDA(c, cudaMemcpyAsync(arrival, c->arrival.p, (size_t)F * sizeof(fitslam_arrival), cudaMemcpyDefault, c->stream));
  if (path_length_m) FS_CUDA(c, cudaMemcpyAsync(path_length_m, ss->len.p, (size_t)F * 8, cudaMemcpyDefault, c->stream));
  if (path_heading) FS_CUDA(c, cudaMemcpyAsync(path_heading, ss->head.p, (size_t)F * 8, cudaMemcpyDefault, c->stream));
  if (weighted_cost) FS_CUDA(c, cudaMemcpyAsync(weighted_cost, ss->wc.p, (size_t)F * 8, cudaMemcpyDefault, c->stream));
  if (fisher_info) {  // Fisher information of each frontier's best-yaw pose, from the (merged) records; 0 when the stage was off
    FS_CUDA(c, s->fisher_best.reserve(F));
    gather_fisher_kernel<<<(F + 255) / 256, 256, 0, c->stream>>>(c->records.p, s->fisher_best.p, F);
    ++c->launches;
    FS_CUDA(c, cudaMemcpyAsync(fisher_info, s->fisher_best.p, (size_t)F * 4, cudaMemcpyDefault, c->stream));
  }
  if (sorted && s->res.selection.n_sorted)
    FS_CUDA(c, cudaMemcpyAsync(sorted, ss->sorted.p, (size_t)s->res.selection.n_sorted * 4, cudaMemcpyDefault, c->stream));
  FS_CUDA(c, cudaStreamSynchronize(c->stream));
  return FITSLAM_OK;
}

fitslam_status fitslam_cycle_fetch_candidates(fitslam_ctx *c, float *information, uint8_t *safe, uint32_t *n_visible, uint32_t capacity,
                                              uint32_t *n_poses) {
  if (!c || !c->cy) return fail(c, FITSLAM_E_INVALID, "cycle_fetch_candidates before fitslam_cycle");
  cudaSetDevice(c->device);
  CycleState *s = c->cy;
  const uint32_t n = s->n_poses;
  if (n_poses) *n_poses = n;
  if (s->fisher_mode == FITSLAM_FISHER_OFF) return fail(c, FITSLAM_E_INVALID, "cycle_fetch_candidates: the last cycle ran without the Fisher stage");
  if (capacity < n) return fail(c, FITSLAM_E_CAPACITY, "cycle_fetch_candidates: buffers too small");
  if (n == 0) return FITSLAM_OK;
  if (information) FS_CUDA(c, cudaMemcpyAsync(information, s->fisher.p, (size_t)n * 4, cudaMemcpyDefault, c->stream));
  if (safe) FS_CUDA(c, cudaMemcpyAsync(safe, s->fisher_safe.p, n, cudaMemcpyDefault, c->stream));
  if (n_visible) FS_CUDA(c, cudaMemcpyAsync(n_visible, s->fisher_nvis.p, (size_t)n * 4, cudaMemcpyDefault, c->stream));
  FS_CUDA(c, cudaStreamSynchronize(c->stream));
  return FITSLAM_OK;
}

}  // extern "C"
