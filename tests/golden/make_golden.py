"""Regenerates tests/golden/cycle_small.npz: the outputs of the CPU oracle (oracle/oracle.cpp) for one small seeded
exploration cycle.  The reference ships no golden vectors for this path and cannot be built here, so
these vectors pin the ORACLE (a change in its arithmetic shows up as a diff against history) and give the CUDA path a
fixture that travels to the GPU box; they are not outputs of the reference itself ("parity unpinned").

    python tests/golden/make_golden.py
"""
import importlib
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def inputs(synth):
    a, meta = synth.make_costmap(320, 288, seed=77, n_disc=7, r_lo=20, r_hi=55, n_holes=8, inflation=3)
    rng = np.random.default_rng(77)
    H, W = a.shape
    pts = np.stack([rng.uniform(meta["ox"], meta["ox"] + W * meta["res"], 3000),
                    rng.uniform(meta["oy"], meta["oy"] + H * meta["res"], 3000),
                    rng.uniform(0.0, 3.0, 3000)], axis=1).astype(np.float32)
    poly = [meta["ox"], meta["oy"], meta["ox"] + W * meta["res"], meta["oy"] + H * meta["res"]]
    return a, meta, pts, poly


def run_oracle(capi, synth):
    from oracle.binding import Oracle
    a, meta, pts, poly = inputs(synth)
    orc = Oracle(capi)
    st, fr, cells, seeds, cent = orc.search_frontiers(a, meta, meta["robot_x"], meta["robot_y"])
    assert st == 0
    arr, rays, wins = orc.score_arrival(a, meta, fr, poly)
    ach, pl, ph = orc.plan_euclidean(meta["robot_x"], meta["robot_y"], 0.3, fr, arr["achievable"])
    stn, wc, ua, ud, err = orc.normalise(arr, ach, pl, ph, None, orc.arrival_limits(a, meta)[0])
    srt, sel = orc.select(ach, pl)
    orc.fisher_generate()
    poses = synth.frontier_poses(fr, arr["theta_s_star"], orc.params.candidate_pose_z)
    q = np.array([orc.orientation_around_z(t) for t in arr["theta_s_star"]])
    poses["qx"], poses["qy"], poses["qz"], poses["qw"] = q.T
    stf, finfo, *_ = orc.fisher_poses(pts, poses)
    return dict(goal_mx=fr["goal_mx"], goal_my=fr["goal_my"], size=fr["size"], component=fr["component"], chunk=fr["chunk"],
                first_cell=fr["first_cell"], cells=cells, seeds=seeds, information=arr["information"],
                max_index=arr["max_index"], achievable=arr["achievable"], footprint_in_lethal=arr["footprint_in_lethal"],
                info_along_ray=rays, window_sums=wins, path_length_m=pl, path_heading=ph, weighted_cost=wc, status_normalise=np.int32(stn),
                sorted=np.asarray(srt, np.uint32), sel=np.array([sel.n_sorted, sel.n_local, sel.global_begin, sel.n_global, sel.closest_global], np.int64),
                fisher=np.asarray(finfo, np.float32))


if __name__ == "__main__":
    from conftest import load_package
    pkg = load_package()
    synth = importlib.import_module("fitslam_b200.synth")
    out = run_oracle(pkg.capi, synth)
    np.savez_compressed(os.path.join(HERE, "cycle_small.npz"), **out)
    print({k: (v.shape, str(v.dtype)) for k, v in out.items()})
