"""Committed fixture tests/golden/cycle_small.npz (made by tests/golden/make_golden.py from the CPU oracle):
  * not-gpu: the oracle still reproduces it bit for bit (pins the oracle against drift);
  * gpu:     the CUDA path, through the C ABI, reproduces it (integer fields bit-exact, weighted cost bit-exact
             relative, Fisher information within the north star's 1e-4 relative)."""
import importlib
import os
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "golden"))
import make_golden  # noqa: E402

GOLD = np.load(os.path.join(HERE, "golden", "cycle_small.npz"))
INT_FIELDS = ("goal_mx", "goal_my", "size", "component", "chunk", "first_cell", "cells", "seeds", "information", "max_index",
              "achievable", "footprint_in_lethal", "info_along_ray", "window_sums", "sorted", "sel")


def test_oracle_reproduces_golden(capi, synth):
    out = make_golden.run_oracle(capi, synth)
    for k in INT_FIELDS:
        np.testing.assert_array_equal(out[k], GOLD[k], err_msg=k)
    for k in ("path_length_m", "path_heading", "weighted_cost"):
        np.testing.assert_array_equal(out[k].view(np.uint64), GOLD[k].view(np.uint64), err_msg=k)
    np.testing.assert_array_equal(out["fisher"].view(np.uint32), GOLD["fisher"].view(np.uint32))


@pytest.mark.gpu
def test_cuda_cycle_reproduces_golden(capi, synth):
    a, meta, pts, poly = make_golden.inputs(synth)
    ctx = capi.Context()
    try:
        ctx.set_costmap(a, meta["res"], meta["ox"], meta["oy"])
        ctx.fisher_generate_lookup()
        ctx.set_map_points(pts)
        st, res = ctx.cycle(meta["robot_x"], meta["robot_y"], 0.3, poly, capi.FISHER_BEST_YAW)
        assert st == 0
        got = ctx.cycle_fetch(res.n_frontiers)
        assert res.n_frontiers == len(GOLD["size"]) and res.n_frontier_cells == len(GOLD["cells"])
        for k in ("goal_mx", "goal_my", "size", "component", "chunk", "first_cell"):
            np.testing.assert_array_equal(got["frontiers"][k], GOLD[k], err_msg=k)
        for k in ("information", "max_index", "achievable", "footprint_in_lethal"):
            np.testing.assert_array_equal(got["arrival"][k], GOLD[k], err_msg=k)
        np.testing.assert_array_equal(got["path_length_m"].view(np.uint64), GOLD["path_length_m"].view(np.uint64))
        np.testing.assert_array_equal(got["weighted_cost"].view(np.uint64), GOLD["weighted_cost"].view(np.uint64))
        sel = res.selection
        np.testing.assert_array_equal(got["sorted"][:sel.n_sorted], GOLD["sorted"])
        assert [sel.n_sorted, sel.n_local, sel.global_begin, sel.n_global, sel.closest_global] == GOLD["sel"].tolist()
        f = GOLD["fisher"].astype(np.float64)
        assert np.all(np.abs(got["fisher"] - f) <= 1e-4 * np.abs(f) + 1e-6)  # tolerance: 1e-4 relative (north star)
        # the stage-wise arrays through the stand-alone entry points
        st, fr, cells = ctx.search_frontiers(meta["robot_x"], meta["robot_y"])
        np.testing.assert_array_equal(cells, GOLD["cells"])
        arr, rays, wins = ctx.score_arrival(fr, poly, want_rays=True, want_windows=True)
        np.testing.assert_array_equal(rays, GOLD["info_along_ray"])
        np.testing.assert_array_equal(wins, GOLD["window_sums"])
    finally:
        ctx.close()
