"""The C++ host layer (fit-slam_b200/cpp/frontier_exploration_cuda.hpp: FrontierSearch, Frontier, FrontierCostCalculator,
FisherInformationManager with the reference's method names) built with g++ against libfitslam_cuda.so and driven by
tests/cpp/shim_harness.cpp in the behaviour tree's call order.
  * not-gpu: it compiles, links, and its host-only checks pass (no compute call);
  * gpu:     its per-frontier output equals the oracle's (goal point bits, size, UID, arrival information, orientation,
             achievability; Fisher information within 1e-4 relative)."""
import os
import struct
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "fit-slam_b200")


@pytest.fixture(scope="module")
def harness(tmp_path_factory):
    out = str(tmp_path_factory.mktemp("shim") / "shim_harness")
    assert os.path.exists(os.path.join(PKG, "libfitslam_cuda.so")), "build the library first (__graft_entry__.build())"
    subprocess.check_call(["g++", "-std=c++17", "-O1", "-Wall", "-I", os.path.join(ROOT, "include"), "-I", os.path.join(PKG, "cpp"),
                           os.path.join(ROOT, "tests", "cpp", "shim_harness.cpp"), "-o", out, "-L", PKG, "-lfitslam_cuda",
                           "-Wl,-rpath," + PKG, "-ldl", "-lpthread"])
    return out


def test_shim_compiles_and_host_checks_pass(harness):
    r = subprocess.run([harness, "--no-gpu"], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "0 failed" in r.stdout



@pytest.mark.gpu
def test_shim_matches_oracle(harness, capi, oracle, synth, tmp_path):
    a, meta = synth.make_costmap(360, 300, seed=31, n_disc=7, r_lo=20, r_hi=55, n_holes=8)
    H, W = a.shape
    rng = np.random.default_rng(31)
    pts = np.stack([rng.uniform(meta["ox"], meta["ox"] + W * meta["res"], 2500), rng.uniform(meta["oy"], meta["oy"] + H * meta["res"], 2500),
                    rng.uniform(0.0, 3.0, 2500)], axis=1).astype(np.float32)
    path = tmp_path / "in.bin"
    with open(path, "wb") as f:
        f.write(struct.pack("<3I5d", W, H, len(pts), meta["res"], meta["ox"], meta["oy"], meta["robot_x"], meta["robot_y"]))
        f.write(np.ascontiguousarray(a).tobytes())
        f.write(pts.tobytes())
    r = subprocess.run([harness, str(path)], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    lines = r.stdout.strip().splitlines()
    # the oracle on the same inputs
    st, fr, cells, seeds, cent = oracle.search_frontiers(a, meta, meta["robot_x"], meta["robot_y"])
    poly = [meta["ox"], meta["oy"], meta["ox"] + W * meta["res"], meta["oy"] + H * meta["res"]]
    arr, _, _ = oracle.score_arrival(a, meta, fr, poly)
    oracle.fisher_generate()
    poses = synth.frontier_poses(fr, arr["theta_s_star"], 0.3)
    q = np.array([oracle.orientation_around_z(t) for t in arr["theta_s_star"]])
    poses["qx"], poses["qy"], poses["qz"], poses["qw"] = q.T
    stf, finfo, *_ = oracle.fisher_poses(pts, poses)
    assert float.fromhex(lines[0].split()[1]) == oracle.arrival_limits(a, meta)[0]
    assert lines[1].split() == ["frontiers", str(len(fr)), "cells", str(len(cells))]
    rows = [ln.split() for ln in lines if ln.startswith("F ")]
    assert len(rows) == len(fr) > 0
    import ctypes as C
    for i, row in enumerate(rows):
        gx, gy = float.fromhex(row[1]), float.fromhex(row[2])
        assert gx == fr["goal_x"][i] and gy == fr["goal_y"][i]
        assert int(row[3]) == fr["size"][i]
        assert int(row[4]) == int(oracle.lib.orc_generate_uid(C.c_double(gx), C.c_double(gy)))
        assert float.fromhex(row[5]) == float(arr["information"][i])
        assert float.fromhex(row[6]) == q[i][2]
        assert int(row[7]) == int(arr["achievable"][i])
        fi = float.fromhex(row[8])
        assert abs(fi - finfo[i]) <= 1e-4 * abs(finfo[i]) + 1e-6  # tolerance: 1e-4 relative (north star)
    cx = meta["ox"] + ((cells % W) + 0.5) * meta["res"]
    cy = meta["oy"] + ((cells // W) + 0.5) * meta["res"]
    sx, sy = 0.0, 0.0
    for vx, vy in zip(cx.tolist(), cy.tolist()):
        sx += vx
        sy += vy
    cs = [ln.split() for ln in lines if ln.startswith("cellsum")][0]
    assert float.fromhex(cs[1]) == sx and float.fromhex(cs[2]) == sy
    assert lines[-1].split() == ["offmap", "0"]
    # roadmap edges of the first frontier (FrontierRoadMap::isConnectable, radius 2.6 m)
    seg = np.stack([np.full(len(fr) - 1, fr["goal_x"][0]), np.full(len(fr) - 1, fr["goal_y"][0]), fr["goal_x"][1:], fr["goal_y"][1:]], axis=1)
    wok, _ = oracle.roadmap_connectable(a, meta, seg, 2.6)
    edges = [ln.split() for ln in lines if ln.startswith("edges")][0]
    assert [int(v) for v in edges[1:]] == list(wok)
    assert [ln.split() for ln in lines if ln.startswith("edge1")][0][1] == str(int(wok[0]))
    # SensorSimulator::markRays and LethalMarker::updateCosts through the shim classes
    latest = np.where(a == 254, 100, np.where(a == 255, -1, 0)).astype(np.int8)
    want, n_rays = oracle.sensor_mark_rays(latest, np.full_like(latest, -1), meta["res"], meta["ox"], meta["oy"], meta["robot_x"], meta["robot_y"],
                                           0.4, capi.default_sensor_params())
    idx = np.flatnonzero(want.ravel() != -1)
    wsum = int((idx.astype(np.uint64) * (want.ravel()[idx].astype(np.int64) + 2).astype(np.uint64)).sum())
    assert [ln.split() for ln in lines if ln.startswith("sensor")][0][1:] == [str(n_rays), str(len(idx)), str(wsum)]
    areas = [(meta["robot_x"] + 1.0, meta["robot_y"], 3.5, 0.7, 1), (meta["robot_x"] - 2.0, meta["robot_y"] + 1.0, 1.7, 0.0, 0),
             (meta["ox"] - 5.0, meta["oy"], 1.7, 0.0, 0)]
    stamped, marked = oracle.mark_lethal(a, meta, areas)
    i253 = np.flatnonzero(stamped.ravel() == 253)
    assert marked == 2 and len(i253) > 50
    assert [ln.split() for ln in lines if ln.startswith("lethal")][0][1:] == [str(marked), str(len(i253)), str(int(i253.sum())), str(int((stamped != a).sum()))]
    # FisherInformationManager::getLandmarksInView for the first frontier's pose
    want = oracle.landmarks_in_view(pts, poses[:1]).astype(np.float64)
    lm = [ln.split() for ln in lines if ln.startswith("landmarks")][0]
    assert int(lm[1]) == len(want)
    sums = [0.0, 0.0, 0.0]
    for row in want.tolist():
        for k in range(3):
            sums[k] += row[k]
    assert [float.fromhex(v) for v in lm[2:5]] == sums
    # ... and over a keyframe atlas (setKeyframes + the four-argument getLandmarksInView / isPoseSafe): every point once, and
    # the maxLandmarks cutoff
    P = len(pts)
    lists = [np.arange(0, P, 2), np.arange(1, P, 2), np.arange(min(10, P))]
    kfs = np.zeros(3, dtype=capi.KEYFRAME_DTYPE)
    kfs["Rwc"] = np.eye(3, dtype=np.float32).ravel()
    kfs["count"] = [len(l) for l in lists]
    kfs["first"] = np.concatenate([[0], np.cumsum(kfs["count"])[:-1]])
    idx = np.concatenate(lists).astype(np.uint32)
    w_all, _ = oracle.landmarks_in_view_keyframes(pts, kfs, idx, poses[:1], 1 << 30, 5.0, True)
    w_cut, _ = oracle.landmarks_in_view_keyframes(pts, kfs, idx, poses[:1], 20, 5.0, True)
    sto, winfo, winfo64, wsafe, wn, wused = oracle.fisher_pose_keyframes(pts, kfs, idx, poses[:1], 20, 5.0, True)
    lk = [ln.split() for ln in lines if ln.startswith("landmarks_kf")][0]
    fsum = lambda rows: float(sum(float(v) for v in rows[:, 0].astype(np.float64).tolist())) if len(rows) else 0.0
    assert int(lk[1]) == len(w_all) == len(want) and float.fromhex(lk[2]) == fsum(w_all)
    assert int(lk[3]) == len(w_cut) <= len(w_all) and float.fromhex(lk[4]) == fsum(w_cut)
    assert abs(float.fromhex(lk[5]) - winfo) <= 1e-4 * abs(winfo) + 1e-6 and (abs(winfo - 550.0) < 1.0 or int(lk[6]) == wsafe)

    # ---- cost assignment and goal selection (CostAssigner::getFrontierCosts, FullPathOptimizer::getNextGoal), BT order
    f32 = lambda v: float(np.float32(v))
    poly32 = [f32(meta["ox"]), f32(meta["oy"]), f32(meta["ox"] + W * meta["res"]), f32(meta["oy"] + H * meta["res"])]
    arr2, _, _ = oracle.score_arrival(a, meta, fr, poly32)
    bl = np.zeros(len(fr), np.uint8)
    if len(fr) > 4:
        bl[3] = 1
    ach_in = arr2["achievable"].copy()
    ach, pl, ph = oracle.plan_euclidean(meta["robot_x"], meta["robot_y"], 0.0, fr, ach_in)
    stn, wc, ua, ud, err = oracle.normalise(arr2, ach, pl, ph, bl, oracle.arrival_limits(a, meta)[0])
    assert stn == 0
    head = [ln.split() for ln in lines if ln.startswith("costs")][0]
    assert head[1:] == ["1", str(len(fr))] and not any("NOT THE SAME POINTER" in ln for ln in lines)
    crow = [ln.split() for ln in lines if ln.startswith("C ")]
    assert len(crow) == len(fr)
    DBL_MAX = np.finfo(np.float64).max
    for i, row in enumerate(crow):
        cost, dist, info = (float.fromhex(v) for v in row[1:4])
        if bl[i]:
            assert cost == DBL_MAX and dist == DBL_MAX and info == 0.0
            continue
        assert cost == wc[i], (i, cost, wc[i])                     # weighted cost: bit-exact
        assert dist == (pl[i] if ach[i] else DBL_MAX)
        assert info == float(arr2["information"][i])
        assert int(row[4]) == int(ach[i])
    # getNextGoal: the shortest-path local frontier (no travelling-salesman step in this build), or the closest global one
    elig = np.flatnonzero(ach.astype(bool) & (bl == 0))
    order = elig[np.argsort(pl[elig], kind="stable")]
    nxt = [ln.split() for ln in lines if ln.startswith("next")][0]
    local = [i for i in order[:5] if pl[i] <= 12.0]
    if len(order) <= 5 and len(local) == len(order) and local:
        local = local[:-1]  # need_to_pop
    want_next = local[0] if local else (order[0] if len(order) else None)
    if want_next is not None:
        assert float.fromhex(nxt[2]) == fr["goal_x"][want_next] and float.fromhex(nxt[3]) == fr["goal_y"][want_next]
    rp = synth.frontier_poses(fr[:1], 0.0, 0.3)
    rp["x"], rp["y"] = meta["robot_x"], meta["robot_y"]
    rp["qx"], rp["qy"], rp["qz"], rp["qw"] = 0.0, 0.0, 0.0, 1.0
    str_, rinfo, rsafe, *_ = oracle.fisher_poses(pts, rp)
    assert abs(float.fromhex(nxt[5]) - rinfo[0]) <= 1e-4 * abs(rinfo[0]) + 1e-6
    assert int(nxt[1]) == (0 if rsafe[0] else 1) and int(nxt[4]) == int(rsafe[0])   # SAFE / UNSAFE = isRobotPoseSafe
    assert [ln for ln in lines if ln.startswith("duplicates")][0] == "duplicates Duplicate frontiers found."
    lut = [ln.split() for ln in lines if ln.startswith("lut")][0]
    assert int(lut[1]) > 500 and int(lut[2]) == 0   # the helper free functions reproduce the device's table bit for bit
