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
