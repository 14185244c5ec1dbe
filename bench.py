#!/usr/bin/env python
"""bench.py -- one exploration cycle of the FIT-SLAM hot path on B200, timed against its roofline.

Contract (driver): `python bench.py --gpus N --steps K --warmup W [--impl reference]`; for N > 1 it is launched
under torch.distributed.run, one rank per GPU.  Rank 0 prints ONE JSON line.

A "step" is one exploration cycle over the BASELINE.json workload `configs[1]` (C2: 4096^2 costmap at 5 cm,
~2k frontiers x 16 yaw windows x 64 rays, 200k map points), synthetic and seeded (fit-slam_b200/synth.py):
frontier extraction -> arrival information (raycast fan + footprint test + window argmax) -> Fisher information of
every (frontier, yaw window) candidate pose -> euclidean plan -> U1 normalisation -> top-k selection
(fitslam_cycle).  With N > 1 ranks every context carries an NCCL communicator (fitslam_comm_init): rank 0 uploads the
map, the library broadcasts it, every rank scores the frontiers f with f % N == rank, and fitslam_cycle merges the 16-byte
score records with an all-reduce on the context stream.

  value   candidates (frontier x yaw window) scored per second, inputs resident in HBM, CUDA-event timed on the
          context stream, L2 flushed between steps (a 256 MiB memset, outside the event pair).
  e2e     the same metric through the C ABI with HOST buffers: the costmap and the map points are copied from pinned
          host memory every step (rank 0; broadcast to the other ranks) and the frontier list, the per-frontier scores
          and the per-candidate Fisher results are read back, all inside the timed region.
  c4      a few cycles of configs[3] (16384^2 map, 1M map points, ~22k frontiers x 32 yaw windows) on the same ranks.
  --impl reference   the CPU oracle (oracle/liboracle.so: a restatement of the reference's single-threaded C++, equal bit for
          bit to the reference's own FrontierSearch / Helpers sources compiled in oracle/_ref; the rest of the reference
          needs ROS 2 / Nav2 / Eigen and cannot be built here) on the same inputs, all host threads, each step a bounded
          sample of the workload: what is timed is what is counted.
"""
import argparse
import ctypes as C
import importlib
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
import __graft_entry__ as graft  # noqa: E402

METRIC = "frontier-yaw candidates scored/sec at 4096^2 grid; ms per exploration cycle"
UNIT = "candidates/s"


def load():
    pkg = graft._load_package()
    return pkg, pkg.capi, importlib.import_module("fitslam_b200.synth")


# --------------------------------------------------------------------------------------------- clocks
class ClockSampler:
    """Samples SM clock and throttle reasons through NVML while the timed region runs."""

    def __init__(self, index=0, period=0.02):
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._thr = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = int(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
        except Exception:  # no NVML: clocks stay empty and the JSON says so
            self.nv = None
        self.period = period

    def _run(self):
        nv = self.nv
        names = {"hw_slowdown": 0x8, "sw_power_cap": 0x4, "sw_thermal_slowdown": 0x20, "hw_thermal_slowdown": 0x40,
                 "hw_power_brake_slowdown": 0x80}
        while not self._stop.is_set():
            try:
                self.samples.append(int(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)))
                r = int(nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)) if hasattr(
                    nv, "nvmlDeviceGetCurrentClocksEventReasons") else int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h))
                for k, bit in names.items():
                    if r & bit:
                        self.reasons.add(k)
            except Exception:
                pass
            time.sleep(self.period)

    def __enter__(self):
        if self.nv:
            self._thr = threading.Thread(target=self._run, daemon=True)
            self._thr.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        if self._thr:
            self._thr.join()

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": 0}
        return {"sm_mhz": float(np.median(self.samples)), "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(self.samples)}


# --------------------------------------------------------------------------------------------- workload
class Workload:
    def __init__(self, name, capi, synth):
        self.name = name
        self.a, self.meta, self.pts, self.ov = synth.make_config(name)
        self.capi, self.synth = capi, synth
        self.idx = int(name[1:])
        self.ryaw = 0.4
        self.fisher_mode = capi.FISHER_ALL_YAWS

    def params(self, base):
        return self.synth.apply_overrides(base, self.ov)

    def describe(self, F, n_rays, n_win, max_len):
        H, W = self.a.shape
        return (f"{self.name}: {W}x{H} costmap @ {self.meta['res']} m, {F} frontiers x {n_win} yaw windows x {n_rays} rays x "
                f"{max_len + 1} cells, {len(self.pts)} map points, Fisher information on all {F * n_win} candidate poses")


def algorithmic_bytes(W, H, F, n_rays, max_len, P):
    """bytes one cycle has to touch, independent of implementation passes."""
    return {"extract": W * H + F * 32, "arrival": F * (n_rays * (max_len + 1) + 625 + 16), "fisher": 12 * P + 8 * F,
            "select": F * 17 + 40}


# --------------------------------------------------------------------------------------------- GPU arm
STAGE_NAMES = ["extract", "arrival", "distance", "fisher", "normalise_select", "begin_total", "end_total", "outer_wavefront_kernel"]


class GpuJob:
    """One context on this rank's GPU with the workload's map resident.  With more than one rank the context carries an NCCL
    communicator (fitslam_comm_init): rank 0 alone uploads the costmap and the map points, the library broadcasts them over
    NVLink, every rank runs the (deterministic) extraction, scores the frontiers f with f % nranks == rank, and fitslam_cycle
    merges the 16-byte score records with an all-reduce on the context stream."""

    def __init__(self, wl, capi, torch, local, rank, world, uid):
        self.wl, self.capi, self.torch, self.rank, self.world = wl, capi, torch, rank, world
        self.ctx = capi.Context(wl.params(capi.default_params()), device=local)
        if world > 1:
            self.ctx.comm_init(uid, rank, world)
        a, pts = wl.a, wl.pts
        self.shape, self.n_pts = a.shape, len(pts)
        # the map lives on rank 0's host only (pinned: the copies inside the timed e2e region are DMA from pinned memory)
        self.a_pin = torch.from_numpy(a).pin_memory().numpy() if rank == 0 else None
        self.pts_pin = torch.from_numpy(pts).pin_memory().numpy() if rank == 0 else None
        self.upload()
        self.ctx.fisher_generate_lookup()
        self.stream = torch.cuda.ExternalStream(self.ctx.stream, device=torch.device("cuda", local))
        self.last = {}

    def upload(self):
        m = self.wl.meta
        self.ctx.set_costmap(self.a_pin, m["res"], m["ox"], m["oy"], shape=self.shape)
        self.ctx.set_map_points(self.pts_pin, n=self.n_pts)

    def cycle(self):
        m = self.wl.meta
        st, res = self.ctx.cycle(m["robot_x"], m["robot_y"], self.wl.ryaw, m["poly"], self.wl.fisher_mode)
        assert st == 0, st
        self.last["res"] = res
        return res

    def _result_buffers(self, n_frontiers, n_poses):
        """Pinned host buffers for the results, kept between steps as a caller of the C ABI would (grown when a cycle finds
        more frontiers than they hold)."""
        torch = self.torch
        b = getattr(self, "_buf", None)
        if b is None or len(b["frontiers"]) < n_frontiers or len(b["candidate_fisher"]) < n_poses:
            def pinned(n, dtype):
                t = torch.empty(max(n, 1) * np.dtype(dtype).itemsize, dtype=torch.uint8).pin_memory()
                return t.numpy().view(dtype)
            self._buf = b = self.capi.Context.fetch_buffers(2 * n_frontiers, 2 * max(n_poses, 1), alloc=pinned)
        return b

    def e2e_step(self):
        self.upload()
        res = self.cycle()
        buf = self._result_buffers(res.n_frontiers, res.n_fisher_poses)
        out = self.ctx.cycle_fetch(res.n_frontiers, out=buf)
        info, safe, nvis = self.ctx.cycle_fetch_candidates(res.n_fisher_poses, out=buf)  # every (frontier, yaw) candidate's Fisher result
        out.update(candidate_fisher=info, candidate_safe=safe, candidate_landmarks=nvis)
        self.last["out"] = out
        return res


def run_gpu(args):
    import torch
    import torch.distributed as dist
    pkg, capi, synth = load()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the hot path has no CPU fallback")
    torch.cuda.set_device(local)
    uids = [None, None]
    saved_stdout = None
    if world > 1:
        # NCCL prints its version banner on stdout when the first communicator is made: stdout carries the JSON line only
        sys.stdout.flush()
        saved_stdout = os.dup(1)
        os.dup2(2, 1)
    if world > 1:
        # torch.distributed is the launcher's plumbing only (barriers, the max over ranks of the timings, carrying the two
        # 128-byte NCCL ids); the data path's collectives are the library's own (csrc/comm.cu)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        if rank == 0:
            uids = [capi.Context.comm_unique_id(), capi.Context.comm_unique_id()]
        dist.broadcast_object_list(uids, src=0)
        dist.barrier()
        torch.cuda.synchronize()
        sys.stdout.flush()
        os.dup2(saved_stdout, 1)
        os.close(saved_stdout)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(job, step, steps, warmup, device_events):
        stream, ctx = job.stream, job.ctx
        for _ in range(warmup):
            with torch.cuda.stream(stream):
                flush.zero_()
            step()
        barrier()
        l0 = ctx.launch_count
        t_dev, wall_in = 0.0, 0.0
        stages = np.zeros(8)
        for _ in range(steps):
            with torch.cuda.stream(stream):
                flush.zero_()
            stream.synchronize()
            if world > 1:
                dist.barrier()
                torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            w0 = time.perf_counter()
            e0.record(stream)
            res = step()
            e1.record(stream)
            e1.synchronize()
            wall_in += time.perf_counter() - w0
            t_dev += e0.elapsed_time(e1)
            stages += np.array(list(res.stage_ms))
        barrier()
        ms = (t_dev if device_events else wall_in * 1e3) / steps
        if world > 1:
            t = torch.tensor([ms], device="cuda", dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms, ctx.launch_count - l0, {n: float(v) for n, v in zip(STAGE_NAMES, stages / steps)}

    def serial_stages(job):
        """One more cycle with the overlap of scoring and search switched off, so that each stage's own time shows."""
        os.environ["FITSLAM_NO_EARLY_SCORING"] = "1"
        try:
            _, _, st = timed(job, job.cycle, 1, 1, True)
        finally:
            del os.environ["FITSLAM_NO_EARLY_SCORING"]
        return st

    wl = Workload(args.config, capi, synth)
    job = GpuJob(wl, capi, torch, local, rank, world, uids[0])
    ctx, meta = job.ctx, wl.meta
    H, W = wl.a.shape
    n_rays, k, n_win, max_len = ctx.ray_shape()
    with ClockSampler(local) as clk:
        ms_dev, launches, stage_ms = timed(job, job.cycle, args.steps, args.warmup, True)
        ms_e2e, _, _ = timed(job, job.e2e_step, max(3, args.steps // 4), max(3, args.warmup // 2), False)
    clocks = clk.summary()
    stage_serial = serial_stages(job)
    res = job.last["res"]
    F = res.n_frontiers
    candidates = F * n_win
    assert res.n_fisher_poses == candidates and int(job.last["out"]["arrival"]["information"].max()) > 0

    # ---- roofline of the dominant kernel of the step (SURVEY 8d algorithmic bytes; measured HBM peak)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    ab = algorithmic_bytes(W, H, F, n_rays, max_len, len(wl.pts))
    # the outer wavefront kernel dominates the cycle; its algorithmic traffic is the searched map.  `traffic`: dram bytes read +
    # written by that kernel in the ncu --set full capture of this round's build (profiles/r2_traffic.json names the commit)
    dom_ms = stage_ms["outer_wavefront_kernel"]
    dom_name = f"ow_cluster_kernel ({res.wavefront_ctas} CTAs)" if res.wavefront_ctas > 1 else "ow_kernel"
    traffic, traffic_source = None, None
    try:
        tj = json.load(open(os.path.join(ROOT, "profiles", "r2_traffic.json")))
        traffic, traffic_source = tj.get(dom_name.split(" ")[0]), tj.get("source")
    except Exception:
        pass
    roof = {"bound": "hbm", "kernel": dom_name + " (inside stage extract)", "achieved": ab["extract"] / (dom_ms * 1e-3) / 1e9 if dom_ms > 0 else None,
            "peak": peak, "unit": "GB/s", "peak_source": "measured" if peaks else "fallback", "traffic": traffic,
            "traffic_source": traffic_source, "algorithmic_bytes_per_launch": ab["extract"], "kernel_ms": dom_ms,
            "note": "latency-bound level-synchronous wavefront (dependent levels, exact FIFO order): see DESIGN.md; per-stage figures below. "
                    "The arrival + Fisher scoring of finished frontier batches runs on side streams while this kernel runs, so "
                    "stage_ms arrival / fisher only hold the batch scored after the search; stage_ms_serial is one cycle with that overlap off",
            "stages": {n: {"ms": stage_serial[n], "algorithmic_bytes": ab[k2], "gbs": ab[k2] / (stage_serial[n] * 1e-3) / 1e9 if stage_serial[n] > 0 else None}
                       for n, k2 in (("extract", "extract"), ("arrival", "arrival"), ("fisher", "fisher"), ("normalise_select", "select"))}}
    roof["frac"] = roof["achieved"] / peak if roof["achieved"] else None
    h2d = int(wl.a.nbytes + wl.pts.nbytes)
    d2h = int(sum(v.nbytes for v in job.last["out"].values()))
    out = {
        "metric": METRIC, "value": candidates / (ms_dev * 1e-3), "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_dev, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "u8/int32 (extraction, arrival), f32 (Fisher), f64 (geometry, costs)", "data": "synthetic",
        "config": {"workload": wl.describe(F, n_rays, n_win, max_len),
                   "stages": ["extract", "arrival", "fisher(all yaws)", "plan", "normalise", "select"],
                   "l2": "flushed between steps (256 MiB memset)",
                   "sharding": (f"frontier f scored by rank f % {world}; extraction replicated; costmap + map points uploaded by rank 0 and "
                                "broadcast by the library (ncclBroadcast, inside the e2e region); 16 B score records merged by the library "
                                "(ncclAllReduce, int32 sum into zeroed rows) inside fitslam_cycle" if world > 1 else "none"),
                   "components": int(res.n_components), "frontier_cells": int(res.n_frontier_cells),
                   "wavefront_levels": int(res.bfs_levels), "widest_level": int(res.widest_level), "wavefront_kernel": dom_name},
        "stage_ms": stage_ms, "stage_ms_serial": stage_serial,
        "e2e": {"value": candidates / (ms_e2e * 1e-3), "unit": UNIT, "ms_per_step": ms_e2e,
                "h2d_bytes_per_step": h2d if rank == 0 else 0, "d2h_bytes_per_step": d2h,
                "note": "rank 0: pinned host costmap + map points -> device (+ broadcast to the other ranks), cycle, frontier list + "
                        "per-frontier scores + per-candidate Fisher results -> host"},
        "gpu_launches": int(launches), "clocks": clocks, "roofline": roof,
    }
    job.ctx.close()
    del job

    # ---- the 16384^2 configuration (C4), a few cycles on every line: the config the multi-GPU target is written for
    if args.config != "C4" and not args.no_c4:
        wl4 = Workload("C4", capi, synth)
        job4 = GpuJob(wl4, capi, torch, local, rank, world, uids[1])
        ms4, launches4, st4 = timed(job4, job4.cycle, 3, 2, True)
        ms4_e2e, _, _ = timed(job4, job4.e2e_step, 2, 1, False)
        ser4 = serial_stages(job4)
        r4 = job4.last["res"]
        out["c4"] = {"workload": wl4.describe(r4.n_frontiers, *[job4.ctx.ray_shape()[i] for i in (0, 2, 3)]), "steps": 3, "warmup": 2,
                     "ms_per_step": ms4, "e2e_ms_per_step": ms4_e2e, "value": r4.n_fisher_poses / (ms4 * 1e-3), "unit": UNIT,
                     "outer_wavefront_ms": st4["outer_wavefront_kernel"], "fisher_ms": ser4["fisher"], "arrival_ms": ser4["arrival"],
                     "frontiers": int(r4.n_frontiers), "candidates": int(r4.n_fisher_poses), "wavefront_levels": int(r4.bfs_levels),
                     "widest_level": int(r4.widest_level),
                     "wavefront_kernel": f"ow_cluster_kernel ({r4.wavefront_ctas} CTAs)" if r4.wavefront_ctas > 1 else "ow_kernel",
                     "gpu_launches": int(launches4), "stage_ms": st4, "stage_ms_serial": ser4,
                     "note": "fisher_ms / arrival_ms: the stage alone, from one cycle with the search/scoring overlap off"}
        job4.ctx.close()
        del job4
    if rank == 0 and world == 1 and not args.no_cpu:
        out["cpu_baseline"] = cpu_baseline(wl, capi, threads=1, budget_s=20.0)
    if rank == 0:
        print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


# --------------------------------------------------------------------------------------------- CPU arms
class CpuCycle:
    """The same cycle on the CPU oracle (oracle/liboracle.so), on a bounded SAMPLE of the workload: the extraction runs in full
    (one thread, as the reference does); then every k-th frontier is scored -- arrival information, and Fisher information on
    all of its yaw-window poses, brute force over all map points per pose like the reference's exhaustive landmark query --
    and the sampled frontiers go through plan, U1 and selection.  Arrival and Fisher scoring are per-frontier / per-pose
    independent, so with threads > 1 they are split over host threads (ctypes releases the GIL).  What is timed is what is
    counted: value = the sampled candidates / the measured time of the step."""

    def __init__(self, wl, capi, threads, sample_frontiers):
        from oracle.binding import Oracle
        self.wl, self.capi, self.threads = wl, capi, max(1, threads)
        self.orcs = [Oracle(capi) for _ in range(self.threads)]
        for o in self.orcs:
            wl.params(o.params)
            o.fisher_generate()
        self.sample_frontiers = sample_frontiers
        self.n_win = self.orcs[0].ray_shape(wl.meta["res"])[2]

    def _parallel(self, fn, chunks):
        if self.threads == 1:
            return [fn(0, chunks[0])]
        out = [None] * len(chunks)

        def run(i):
            out[i] = fn(i, chunks[i])
        th = [threading.Thread(target=run, args=(i,)) for i in range(len(chunks))]
        for t in th:
            t.start()
        for t in th:
            t.join()
        return out

    def step(self, reference_workload=False):
        """reference_workload: the cycle as the reference itself runs it -- every frontier's arrival information, then Fisher
        information on ONE pose (the selected frontier's best yaw), all on one thread."""
        wl, o = self.wl, self.orcs[0]
        a, meta = wl.a, wl.meta
        t = [time.perf_counter()]
        st, fr, cells, seeds, cent = o.search_frontiers(a, meta, meta["robot_x"], meta["robot_y"], capacity=1 << 18,
                                                        cell_capacity=1 << 23)
        t.append(time.perf_counter())
        F = len(fr)
        K = F if reference_workload else min(self.sample_frontiers, F)
        pick = np.unique(np.linspace(0, F - 1, K).astype(np.int64))
        K = len(pick)
        frs = fr[pick]
        parts = np.array_split(np.arange(K), 1 if reference_workload else self.threads)
        res = self._parallel(lambda i, idx: self.orcs[i].score_arrival(a, meta, frs[idx], meta["poly"])[0], parts)
        arr = np.concatenate(res)
        t.append(time.perf_counter())
        ach, pl, ph = o.plan_euclidean(meta["robot_x"], meta["robot_y"], wl.ryaw, frs, arr["achievable"])
        max_gt = o.arrival_limits(a, meta)[0]
        o.normalise(arr, ach, pl, ph, None, max_gt)
        o.select(ach, pl)
        t.append(time.perf_counter())
        if reference_workload:
            theta = np.array([0.5 * o.params.camera_fov])
            poses = wl.synth.frontier_poses(frs[:1], theta, o.params.candidate_pose_z)
            pparts = [np.arange(1)]
        else:
            f_idx = np.repeat(np.arange(K), self.n_win)
            theta = np.tile(np.arange(self.n_win), K) * o.params.delta_theta + o.params.camera_fov / 2
            poses = wl.synth.frontier_poses(frs[f_idx], theta, o.params.candidate_pose_z)
            pparts = np.array_split(np.arange(len(poses)), self.threads)
        self._parallel(lambda i, idx: self.orcs[i].fisher_poses(wl.pts, poses[idx]), pparts)
        t.append(time.perf_counter())
        d = np.diff(t)
        return dict(F=F, sampled_frontiers=K, candidates=len(poses), s_total=float(t[-1] - t[0]),
                    ms={"extract": d[0] * 1e3, "arrival": d[1] * 1e3, "plan_normalise_select": d[2] * 1e3, "fisher": d[3] * 1e3})


def sample_text(r, wl, threads):
    return (f"each step: extraction of the whole map in full (1 thread, {r['F']} frontiers found), then {r['sampled_frontiers']} of the "
            f"frontiers (every {r['F'] // max(1, r['sampled_frontiers'])}th) scored on {threads} host thread(s): arrival information, "
            f"Fisher information on all {r['candidates']} of their yaw-window poses (brute force over all {len(wl.pts)} map points, as "
            "the reference's exhaustive landmark query), plan, U1, selection; value = those candidates / the measured step time")


def cpu_baseline(wl, capi, threads, budget_s):
    cyc = CpuCycle(wl, capi, threads, sample_frontiers=8 * threads)
    cyc.step()
    t0, runs = time.perf_counter(), []
    while time.perf_counter() - t0 < budget_s and len(runs) < 20:
        runs.append(cyc.step())
    s_total = float(np.median([x["s_total"] for x in runs]))
    r = runs[len(runs) // 2]
    refw = [cyc.step(reference_workload=True) for _ in range(3)][1:]
    rw = refw[-1]
    return {"value": r["candidates"] / s_total, "unit": UNIT, "cores": threads, "kind": "port", "ms_per_step": s_total * 1e3,
            "stage_ms": r["ms"], "sample": f"{len(runs)} steps; " + sample_text(r, wl, threads),
            "reference_workload": {
                "ms_per_cycle": float(np.mean([x["s_total"] for x in refw])) * 1e3, "stage_ms": rw["ms"], "cores": 1,
                "what": f"the cycle as the reference runs it: extraction, arrival information of all {rw['F']} frontiers, plan, U1, "
                        "selection, and Fisher information on ONE pose (isPoseSafe of the chosen goal), single thread; not the "
                        "metric's workload (the metric scores every candidate pose), quoted for scale"}}


def host_threads(limit=32):
    """Host threads this process may really use: the affinity mask, cut by a cgroup CPU quota where one is set."""
    n = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    try:
        quota, period = open("/sys/fs/cgroup/cpu.max").read().split()
        if quota != "max":
            n = min(n, max(1, int(int(quota) / int(period))))
    except Exception:
        pass
    return max(1, min(n, limit))


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    pkg, capi, synth = load()
    wl = Workload(args.config, capi, synth)
    threads = host_threads()
    # about a second of Fisher work per step whatever the core count: 8 frontiers x n_win poses per thread
    cyc = CpuCycle(wl, capi, threads, sample_frontiers=8 * threads)
    for _ in range(args.warmup):
        cyc.step()
    t0 = time.perf_counter()
    runs = [cyc.step() for _ in range(args.steps)]
    wall = time.perf_counter() - t0
    r = runs[-1]
    ms = wall / len(runs) * 1e3
    val = r["candidates"] / (ms * 1e-3)
    n_rays, k, n_win, max_len = cyc.orcs[0].ray_shape(wl.meta["res"])
    sample = sample_text(r, wl, threads)
    out = {"impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": int(os.environ.get("WORLD_SIZE", "1")),
           "steps": len(runs), "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
           "scaling": "strong", "vs_baseline": None, "dtype": "u8/int32, f32, f64", "data": "synthetic",
           "config": {"workload": wl.describe(r["F"], n_rays, n_win, max_len),
                      "stages": ["extract", "arrival", "fisher(all yaws)", "plan", "normalise", "select"]},
           "stage_ms": {n: float(np.mean([x["ms"][n] for x in runs])) for n in r["ms"]},
           "candidates_per_step": r["candidates"], "timed_region_s": wall,
           "cpu_baseline": {"value": val, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
           "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(out))


def run_c5(args):
    """BASELINE.json configs[4] (rapid mode): `--steps` successive incremental costmap updates on the C2 map (a random walk of
    0.2 m steps revealing a 2 m disc each: SURVEY 8d), every one followed by a full re-extraction and rescoring cycle.  An update
    uploads only the dirty rectangle from pinned host memory.  Timed per update + cycle on the host clock (the call returns
    after its own sync), L2 not flushed (a rapid-mode loop would not either); the reference ticks at 5 Hz
    (fit_slam2/xml/rapid_exploration.xml:4)."""
    import torch
    pkg, capi, synth = load()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the hot path has no CPU fallback")
    wl = Workload("C2", capi, synth)
    a, meta = wl.a, wl.meta
    H, W = a.shape
    ctx = capi.Context(wl.params(capi.default_params()))
    a_pin = torch.from_numpy(a.copy()).pin_memory().numpy()
    ctx.set_costmap(a_pin, meta["res"], meta["ox"], meta["oy"])
    ctx.fisher_generate_lookup()
    ctx.set_map_points(wl.pts)
    rng = np.random.default_rng(55)
    cx = int((meta["robot_x"] - meta["ox"]) / meta["res"])
    cy = int((meta["robot_y"] - meta["oy"]) / meta["res"])
    n_win = ctx.ray_shape()[2]
    ms, cand, dirty, fr = [], [], [], []
    l0 = None
    with ClockSampler(0) as clk:
        for step in range(args.warmup + args.steps):
            cx = int(np.clip(cx + rng.integers(-4, 5), 100, W - 100))
            cy = int(np.clip(cy + rng.integers(-4, 5), 100, H - 100))
            rect = synth.reveal_step(a_pin, meta, cx, cy, rng)  # (host work of the producer, outside the timed region)
            rx, ry = meta["ox"] + (cx + 0.5) * meta["res"], meta["oy"] + (cy + 0.5) * meta["res"]
            if step == args.warmup:
                l0 = ctx.launch_count
            t0 = time.perf_counter()
            ctx.set_costmap(a_pin, meta["res"], meta["ox"], meta["oy"], dirty=rect)
            st, res = ctx.cycle(rx, ry, wl.ryaw, meta["poly"], wl.fisher_mode)
            t1 = time.perf_counter()
            assert st == 0, st
            if step >= args.warmup:
                ms.append((t1 - t0) * 1e3)
                cand.append(res.n_frontiers * n_win)
                dirty.append(rect[2] * rect[3])
                fr.append(res.n_frontiers)
    ms = np.array(ms)
    out = {"metric": METRIC, "value": float(np.sum(cand) / (ms.sum() * 1e-3)), "unit": UNIT, "n_gpus": 1, "steps": args.steps, "warmup": args.warmup,
           "ms_per_step": float(ms.mean()), "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
           "dtype": "u8/int32 (extraction, arrival), f32 (Fisher), f64 (geometry, costs)", "data": "synthetic",
           "config": {"workload": f"C5: {args.steps} successive incremental costmap updates on the C2 map ({W}x{H} @ {meta['res']} m, {len(wl.pts)} map "
                                  f"points), re-extraction + rescoring of all (frontier, yaw) candidates per update; dirty rectangle uploaded from pinned host memory",
                      "l2": "not flushed (successive cycles of one loop)"},
           "c5": {"updates": int(args.steps), "ms_median": float(np.median(ms)), "ms_p99": float(np.percentile(ms, 99)), "ms_max": float(ms.max()),
                  "ms_min": float(ms.min()), "dirty_bytes_mean": float(np.mean(dirty)), "frontiers_mean": float(np.mean(fr)),
                  "candidates_mean": float(np.mean(cand)), "tick_budget_ms": 200.0,
                  "note": "per update: dirty-rectangle H2D + whole cycle through the C ABI, host clock around both calls"},
           "e2e": {"value": float(np.sum(cand) / (ms.sum() * 1e-3)), "unit": UNIT, "ms_per_step": float(ms.mean()),
                   "h2d_bytes_per_step": int(np.mean(dirty)), "d2h_bytes_per_step": 0},
           "gpu_launches": int(ctx.launch_count - l0), "clocks": clk.summary()}
    ctx.close()
    print(json.dumps(out))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--config", default="C2")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-c4", action="store_true", help="skip the c4 block (three cycles of the 16384^2 configuration)")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    elif args.config == "C5":
        run_c5(args)
    else:
        run_gpu(args)


if __name__ == "__main__":
    main()
