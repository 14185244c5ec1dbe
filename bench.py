#!/usr/bin/env python
"""bench.py -- one exploration cycle of the FIT-SLAM hot path on B200, timed against its roofline.

Contract (driver): `python bench.py --gpus N --steps K --warmup W [--impl reference]`; for N > 1 it is launched
under torch.distributed.run, one rank per GPU.  Rank 0 prints ONE JSON line.

A "step" is one exploration cycle over the BASELINE.json workload `configs[1]` (C2: 4096^2 costmap at 5 cm,
~2k frontiers x 16 yaw windows x 64 rays, 200k map points), synthetic and seeded (fit-slam_b200/synth.py):
frontier extraction -> arrival information (raycast fan + footprint test + window argmax) -> Fisher information of
every (frontier, yaw window) candidate pose -> euclidean plan -> U1 normalisation -> top-k selection
(fitslam_cycle; with N > 1 ranks: fitslam_cycle_begin on this rank's block of frontiers, NCCL all-gather of the
16-byte score records, fitslam_cycle_end on every rank).

  value   candidates (frontier x yaw window) scored per second, inputs resident in HBM, CUDA-event timed on the
          context stream, L2 flushed between steps (a 256 MiB memset, outside the event pair).
  e2e     the same metric through the C ABI with HOST buffers: the costmap (and map points) are copied from pinned
          host memory every step and the per-frontier records are read back, all inside the timed region.
  --impl reference   the CPU oracle (oracle/liboracle.so, a restatement of the reference's single-threaded C++;
          the reference itself needs ROS 2/Nav2/Eigen and cannot be built here) on the same inputs.
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
def run_gpu(args):
    import torch
    import torch.distributed as dist
    pkg, capi, synth = load()
    shard = importlib.import_module("fitslam_b200.shard")
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the hot path has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    wl = Workload(args.config, capi, synth)
    ctx = capi.Context(wl.params(capi.default_params()), device=local)
    a, meta, pts = wl.a, wl.meta, wl.pts
    H, W = a.shape
    a_pin = torch.from_numpy(a).pin_memory().numpy()
    pts_pin = torch.from_numpy(pts).pin_memory().numpy()
    # map update (once): costmap + map points to the device, lookup table; with N ranks every rank holds a replica
    ctx.set_costmap(a_pin, meta["res"], meta["ox"], meta["oy"])
    ctx.fisher_generate_lookup()
    ctx.set_map_points(pts_pin)
    if world > 1:
        ctx.set_shard(rank, world)
    stream = torch.cuda.ExternalStream(ctx.stream, device=torch.device("cuda", local))
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    rx, ry, ryaw, poly, mode = meta["robot_x"], meta["robot_y"], wl.ryaw, meta["poly"], wl.fisher_mode
    n_rays, k, n_win, max_len = ctx.ray_shape()
    last = {}

    def cycle():
        if world == 1:
            st, res = ctx.cycle(rx, ry, ryaw, poly, mode)
            assert st == 0, st
        else:
            st, F = ctx.cycle_begin(rx, ry, ryaw, poly, mode)
            assert st == 0, st
            rec = ctx.records_tensor()
            with torch.cuda.stream(stream):  # NCCL orders itself after the scoring kernels of this stream
                shard.allreduce_records_(rec, world)
            res = ctx.cycle_end()
        last["res"] = res
        return res

    def resident_step():
        return cycle()

    def e2e_step():
        ctx.set_costmap(a_pin, meta["res"], meta["ox"], meta["oy"])
        ctx.set_map_points(pts_pin)
        res = cycle()
        last["out"] = ctx.cycle_fetch(res.n_frontiers)
        return res

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(step, steps, warmup, device_events):
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
        return ms, ctx.launch_count - l0, stages / steps

    with ClockSampler(local) as clk:
        ms_dev, launches, stages = timed(resident_step, args.steps, args.warmup, True)
        ms_e2e, _, _ = timed(e2e_step, max(3, args.steps // 4), max(3, args.warmup // 2), False)
    clocks = clk.summary()
    res = last["res"]
    F = res.n_frontiers
    candidates = F * n_win
    assert res.n_fisher_poses == candidates and int(last["out"]["arrival"]["information"].max()) > 0

    # ---- roofline of the dominant kernel of the step (SURVEY 8d algorithmic bytes; measured HBM peak)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    ab = algorithmic_bytes(W, H, F, n_rays, max_len, len(pts))
    stage_names = ["extract", "arrival", "distance", "fisher", "normalise_select", "begin_total", "end_total", "outer_wavefront_kernel"]
    stage_ms = {n: float(v) for n, v in zip(stage_names, stages)}
    # the outer wavefront kernel (rank_bfs_smem_kernel) dominates the cycle; its algorithmic traffic is the searched map.
    # `traffic`: dram__bytes_read.sum + dram__bytes_write.sum of that kernel from the committed ncu --set full capture
    dom_ms = stage_ms["outer_wavefront_kernel"]
    traffic = None
    try:
        traffic = json.load(open(os.path.join(ROOT, "profiles", "r1_traffic.json"))).get("rank_bfs_smem_kernel")
    except Exception:
        pass
    roof = {"bound": "hbm", "kernel": "rank_bfs_smem_kernel (inside stage extract)", "achieved": ab["extract"] / (dom_ms * 1e-3) / 1e9 if dom_ms > 0 else None,
            "peak": peak, "unit": "GB/s", "peak_source": "measured" if peaks else "fallback", "traffic": traffic,
            "algorithmic_bytes_per_launch": ab["extract"], "kernel_ms": dom_ms,
            "note": "latency-bound level-synchronous wavefront (one CTA, 2205 dependent levels): see DESIGN.md; per-stage fractions below. "
                    "On one GPU the arrival + Fisher scoring of finished frontier batches runs on side streams while this kernel runs, so "
                    "stages arrival / fisher only hold the batch scored after the search (FITSLAM_NO_EARLY_SCORING=1 separates them)",
            "stages": {n: {"ms": stage_ms[n], "algorithmic_bytes": ab[k2], "gbs": ab[k2] / (stage_ms[n] * 1e-3) / 1e9 if stage_ms[n] > 0 else None}
                       for n, k2 in (("extract", "extract"), ("arrival", "arrival"), ("fisher", "fisher"), ("normalise_select", "select"))}}
    roof["frac"] = roof["achieved"] / peak if roof["achieved"] else None
    h2d = int(a.nbytes + pts.nbytes)
    d2h = int(sum(v.nbytes for v in last["out"].values()))
    out = {
        "metric": METRIC, "value": candidates / (ms_dev * 1e-3), "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_dev, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "u8/int32 (extraction, arrival), f32 (Fisher), f64 (geometry, costs)", "data": "synthetic",
        "config": {"workload": wl.describe(F, n_rays, n_win, max_len),
                   "stages": ["extract", "arrival", "fisher(all yaws)", "plan", "normalise", "select"],
                   "l2": "flushed between steps (256 MiB memset)",
                   "sharding": (f"frontier f scored by rank f % {world}, extraction replicated, NCCL all-reduce (integer sum into zeroed rows) of the 16 B records"
                                if world > 1 else "none"),
                   "components": int(res.n_components), "frontier_cells": int(res.n_frontier_cells),
                   "wavefront_levels": int(res.bfs_levels)},
        "stage_ms": stage_ms,
        "e2e": {"value": candidates / (ms_e2e * 1e-3), "unit": UNIT, "ms_per_step": ms_e2e, "h2d_bytes_per_step": h2d,
                "d2h_bytes_per_step": d2h},
        "gpu_launches": int(launches), "clocks": clocks, "roofline": roof,
    }
    if rank == 0 and world == 1 and not args.no_cpu:
        out["cpu_baseline"] = cpu_baseline(wl, capi, threads=1, budget_s=20.0)
    ctx.close()
    if rank == 0:
        print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


# --------------------------------------------------------------------------------------------- CPU arms
class CpuCycle:
    """The same cycle on the CPU oracle (oracle/liboracle.so).  Extraction, plan, U1 and selection run as the
    reference does (one thread).  Arrival scoring and Fisher scoring are per-frontier / per-pose independent, so with
    threads > 1 they are split over host threads (ctypes releases the GIL).  Fisher information is brute force over all
    map points per pose, like the reference's exhaustive landmark query; it is timed on a SAMPLE of the candidate
    poses and scaled to all of them (the full set would take minutes per cycle)."""

    def __init__(self, wl, capi, threads, fisher_sample):
        from oracle.binding import Oracle
        self.wl, self.capi, self.threads = wl, capi, max(1, threads)
        self.orcs = [Oracle(capi) for _ in range(self.threads)]
        for o in self.orcs:
            wl.params(o.params)
            o.fisher_generate()
        self.fisher_sample = fisher_sample
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

    def step(self):
        wl, o = self.wl, self.orcs[0]
        a, meta = wl.a, wl.meta
        t = [time.perf_counter()]
        st, fr, cells, seeds, cent = o.search_frontiers(a, meta, meta["robot_x"], meta["robot_y"], capacity=1 << 18,
                                                        cell_capacity=1 << 23)
        t.append(time.perf_counter())
        parts = np.array_split(np.arange(len(fr)), self.threads)
        res = self._parallel(lambda i, idx: self.orcs[i].score_arrival(a, meta, fr[idx], meta["poly"])[0], parts)
        arr = np.concatenate(res)
        t.append(time.perf_counter())
        ach, pl, ph = o.plan_euclidean(meta["robot_x"], meta["robot_y"], wl.ryaw, fr, arr["achievable"])
        max_gt = o.arrival_limits(a, meta)[0]
        o.normalise(arr, ach, pl, ph, None, max_gt)
        o.select(ach, pl)
        t.append(time.perf_counter())
        # Fisher: sample of the F x n_win candidate poses (every k-th), yaw = window * delta_theta + fov / 2
        n_poses = len(fr) * self.n_win
        S = min(self.fisher_sample, n_poses)
        pick = np.linspace(0, n_poses - 1, S).astype(np.int64)
        f_idx, w_idx = pick // self.n_win, pick % self.n_win
        theta = w_idx * o.params.delta_theta + o.params.camera_fov / 2
        poses = wl.synth.frontier_poses(fr[f_idx], theta, o.params.candidate_pose_z)
        pparts = np.array_split(np.arange(S), self.threads)
        t.append(time.perf_counter())
        self._parallel(lambda i, idx: self.orcs[i].fisher_poses(wl.pts, poses[idx]), pparts)
        t.append(time.perf_counter())
        d = np.diff(t)
        fisher_full = d[4] * n_poses / S
        total = d[0] + d[1] + d[2] + fisher_full
        return dict(F=len(fr), candidates=n_poses, s_total=total, s_measured=d[0] + d[1] + d[2] + d[4],
                    ms={"extract": d[0] * 1e3, "arrival": d[1] * 1e3, "plan_normalise_select": d[2] * 1e3,
                        "fisher_sample": d[4] * 1e3, "fisher_scaled": fisher_full * 1e3}, fisher_sample=S)


def cpu_baseline(wl, capi, threads, budget_s):
    cyc = CpuCycle(wl, capi, threads, fisher_sample=256)
    cyc.step()
    t0, runs = time.perf_counter(), []
    while time.perf_counter() - t0 < budget_s and len(runs) < 20:
        runs.append(cyc.step())
    r = runs[len(runs) // 2]
    s_total = float(np.median([x["s_total"] for x in runs]))
    return {"value": r["candidates"] / s_total, "unit": UNIT, "cores": threads, "kind": "port", "ms_per_step": s_total * 1e3,
            "stage_ms": r["ms"],
            "sample": f"{len(runs)} cycles of the same workload; extraction, arrival ({r['F']} frontiers), plan, U1 and selection in "
                      f"full; Fisher information timed on {r['fisher_sample']} of the {r['candidates']} candidate poses (brute force "
                      f"over all {len(wl.pts)} map points, as the reference's exhaustive landmark query) and scaled to all of them"}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    pkg, capi, synth = load()
    wl = Workload(args.config, capi, synth)
    threads = min(os.cpu_count() or 1, 32)
    cyc = CpuCycle(wl, capi, threads, fisher_sample=max(256, 32 * threads))
    for _ in range(min(args.warmup, 2)):
        cyc.step()
    runs = [cyc.step() for _ in range(max(1, min(args.steps, 10)))]  # each step is a bounded sample (see `sample`)
    s_total = float(np.mean([x["s_total"] for x in runs]))
    r = runs[-1]
    val = r["candidates"] / s_total
    n_rays, k, n_win, max_len = cyc.orcs[0].ray_shape(wl.meta["res"])
    sample = (f"{len(runs)} cycles; extraction, plan, U1, selection single-threaded as in the reference; arrival scoring split over "
              f"{threads} host threads; Fisher information timed on {r['fisher_sample']} of {r['candidates']} candidate poses over "
              f"{threads} threads and scaled to all of them")
    out = {"impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": int(os.environ.get("WORLD_SIZE", "1")),
           "steps": len(runs), "warmup": min(args.warmup, 2), "ms_per_step": s_total * 1e3, "higher_is_better": True,
           "scaling": "strong", "vs_baseline": None, "dtype": "u8/int32, f32, f64", "data": "synthetic",
           "config": {"workload": wl.describe(r["F"], n_rays, n_win, max_len),
                      "stages": ["extract", "arrival", "fisher(all yaws)", "plan", "normalise", "select"]},
           "stage_ms": r["ms"],
           "cpu_baseline": {"value": val, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
           "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(out))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--config", default="C2")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_gpu(args)


if __name__ == "__main__":
    main()
