"""BASELINE.json's big and incremental configurations on the GPU, checked through size-independent properties where the
oracle would take minutes:
  * C4 (16384^2 map, 1M map points, ~20k frontiers x 32 yaw windows): a second cycle on the same context gives the same
    bits (scratch state is clean between cycles), the union of 4 candidate shards equals the unsharded records, every goal
    cell is an unknown cell next to known space, sizes obey the cluster bounds, the every_frontier list has no duplicates;
    the whole frontier list and every_frontier list equal the oracle's bit for bit; samples of the frontiers / candidate poses
    are scored by the oracle (arrival information bit-exact, Fisher information 1e-4);
  * C5 (rapid mode): a random walk reveals discs on a C1-size map; each step uploads only the dirty rectangle, and the
    cycle must equal a cycle on a fresh context with the full map, and the oracle every 8th step."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

REC_FIELDS = ("information", "max_index", "achievable", "footprint_in_lethal", "ray_failed")


def _fetch(ctx, capi, meta, mode):
    st, res = ctx.cycle(meta["robot_x"], meta["robot_y"], 0.3, meta["poly"], mode)
    assert st == 0
    return res, ctx.cycle_fetch(res.n_frontiers)


def test_c4_full_size_properties(capi, oracle, synth):
    a, meta, pts, ov = synth.make_config("C4")
    params = synth.apply_overrides(capi.default_params(), ov)
    synth.apply_overrides(oracle.params, ov)
    H, W = a.shape
    ctx = capi.Context(params)
    try:
        ctx.set_costmap(a, meta["res"], meta["ox"], meta["oy"])
        ctx.fisher_generate_lookup()
        ctx.set_map_points(pts)
        res1, out1 = _fetch(ctx, capi, meta, capi.FISHER_ALL_YAWS)
        res2, out2 = _fetch(ctx, capi, meta, capi.FISHER_ALL_YAWS)
        info, safe, nvis = ctx.cycle_fetch_candidates(res2.n_fisher_poses)
        F = res1.n_frontiers
        n_rays, k, n_win, max_len = ctx.ray_shape()
        assert F > 5000 and n_win == 32 and res1.n_candidates == F * n_win == res1.n_fisher_poses
        print(f"C4: {F} frontiers, {res1.n_components} components, {res1.bfs_levels} levels, stage_ms first cycle {[round(v, 3) for v in res1.stage_ms]}, "
              f"second cycle {[round(v, 3) for v in res2.stage_ms]}")
        # idempotence: the second cycle is bit-identical
        assert res2.n_frontiers == F
        for key in out1:
            assert out1[key].tobytes() == out2[key].tobytes(), key
        fr = out1["frontiers"]
        # every goal is an unknown cell with a known, non-lethal 4-neighbour (isNewFrontierCell), sizes within bounds
        gx, gy = fr["goal_mx"].astype(np.int64), fr["goal_my"].astype(np.int64)
        assert np.all(a[gy, gx] == 255)
        nb = np.stack([a[gy, gx - 1], a[gy, gx + 1], a[gy - 1, gx], a[gy + 1, gx]])
        assert np.all((nb < params.lethal_threshold).any(axis=0))
        assert fr["size"].min() > params.min_frontier_cluster_size and fr["size"].max() <= params.max_frontier_cluster_size + 1
        st, fr2, cells = ctx.search_frontiers(meta["robot_x"], meta["robot_y"], capacity=1 << 17, cell_capacity=1 << 23)
        assert len(cells) == res1.n_frontier_cells and len(np.unique(cells)) == len(cells)
        assert fr2.tobytes() == fr.tobytes()
        # shards: the union of the 4 shares equals the unsharded arrival records and Fisher information
        merged_arr = np.zeros(F, dtype=out1["arrival"].dtype)
        merged_f = np.zeros(F, np.float32)
        for r in range(4):
            ctx.set_shard(r, 4)
            st, Fr = ctx.cycle_begin(meta["robot_x"], meta["robot_y"], 0.3, meta["poly"], capi.FISHER_ALL_YAWS)
            assert st == 0 and Fr == F
            import torch
            torch.cuda.synchronize()  # cycle_begin leaves the scoring kernels in flight on the context stream
            rec = ctx.records_tensor().cpu().numpy()
            assert rec.shape[0] == F
            own = np.arange(F) % 4 == r  # the cycle shards by frontier index modulo the rank count
            assert not rec[~own].any()
            merged_arr["information"][own] = rec[own, 0]
            merged_arr["max_index"][own] = rec[own, 1]
            merged_f[own] = rec[own, 2].view(np.float32)
            ctx.cycle_end()
        ctx.set_shard(0, 1)
        np.testing.assert_array_equal(merged_arr["information"], out1["arrival"]["information"])
        np.testing.assert_array_equal(merged_arr["max_index"], out1["arrival"]["max_index"])
        np.testing.assert_array_equal(merged_f.view(np.uint32), out1["fisher"].view(np.uint32))
        # the whole extraction against the oracle (18 M expanded cells: the cluster wavefront and the 30 k-cell rim component
        # at their real sizes): frontier list and every_frontier list bit for bit
        sto, wfr, wcells, wseeds, wcent = oracle.search_frontiers(a, meta, meta["robot_x"], meta["robot_y"], capacity=1 << 17, cell_capacity=1 << 23)
        assert sto == 0 and len(wfr) == F and len(wseeds) == res1.n_components
        np.testing.assert_array_equal(cells, wcells)
        for key in ("goal_mx", "goal_my", "size", "component", "chunk", "first_cell"):
            np.testing.assert_array_equal(fr[key], wfr[key], err_msg=key)
        np.testing.assert_array_equal(fr["goal_x"].view(np.uint64), wfr["goal_x"].view(np.uint64))
        np.testing.assert_array_equal(fr["goal_y"].view(np.uint64), wfr["goal_y"].view(np.uint64))
        # a sample of the frontiers through the oracle: arrival information bit-exact
        pick = np.linspace(0, F - 1, 400).astype(np.int64)
        arr, _, _ = oracle.score_arrival(a, meta, fr[pick], meta["poly"])
        for key in REC_FIELDS:
            np.testing.assert_array_equal(out1["arrival"][key][pick], arr[key], err_msg=key)
        # path lengths, the sorted list and the selection of all 21 k frontiers (the ranks of a list this long are counted by
        # several CTAs per frontier): the oracle's plan + std::sort on the device's arrival records
        ach, pl, ph = oracle.plan_euclidean(meta["robot_x"], meta["robot_y"], 0.3, fr, out1["arrival"]["achievable"])
        np.testing.assert_array_equal(out1["path_length_m"].view(np.uint64), pl.view(np.uint64))
        wsrt, wsel = oracle.select(ach, pl)
        assert res1.selection.n_sorted == wsel.n_sorted == len(wsrt) > 0 and F > 4096
        np.testing.assert_array_equal(out1["sorted"][:wsel.n_sorted], wsrt)
        for f in ("n_local", "global_begin", "n_global", "closest_global"):
            assert getattr(res1.selection, f) == getattr(wsel, f), f
        # ... and Fisher information of 96 sampled candidates (3 yaw windows of 32 frontiers) over the 1 M map points: 1e-4
        oracle.fisher_generate()
        pf = np.repeat(np.linspace(0, F - 1, 32).astype(np.int64), 3)
        pw = np.tile(np.array([0, 11, 31]), 32)
        th = pw * params.delta_theta + params.camera_fov / 2
        poses = synth.frontier_poses(fr[pf], th, params.candidate_pose_z)
        q = np.array([oracle.orientation_around_z(t) for t in th])
        poses["qx"], poses["qy"], poses["qz"], poses["qw"] = q.T
        stf, winfo, wsafe, wvis, wused = oracle.fisher_poses(pts, poses)
        g = info[pf * n_win + pw]
        okp = out1["arrival"]["ray_failed"][pf] == 0
        assert np.all(np.abs(g - winfo)[okp] <= 1e-4 * np.abs(winfo[okp]) + 1e-6), np.max(np.abs(g - winfo)[okp] / np.abs(winfo[okp]))
        np.testing.assert_array_equal(nvis[pf * n_win + pw][okp], wused[okp])
    finally:
        ctx.close()


def test_c5_rapid_mode_dirty_rectangles(capi, oracle, synth):
    a, meta = synth.make_costmap(1000, 1000, seed=55, n_disc=10, r_lo=30, r_hi=80, n_holes=15)
    rng = np.random.default_rng(55)
    H, W = a.shape
    pts = synth.make_points(5000, synth.explored_bbox(a, meta), 55)
    ctx = capi.Context()
    try:
        ctx.set_costmap(a, meta["res"], meta["ox"], meta["oy"])
        ctx.fisher_generate_lookup()
        ctx.set_map_points(pts)
        cx = int((meta["robot_x"] - meta["ox"]) / meta["res"])
        cy = int((meta["robot_y"] - meta["oy"]) / meta["res"])
        for step in range(24):
            cx = int(np.clip(cx + rng.integers(-4, 5), 100, W - 100))  # 0.2 m steps
            cy = int(np.clip(cy + rng.integers(-4, 5), 100, H - 100))
            rect = synth.reveal_step(a, meta, cx, cy, rng)
            rx, ry = meta["ox"] + (cx + 0.5) * meta["res"], meta["oy"] + (cy + 0.5) * meta["res"]
            ctx.set_costmap(a, meta["res"], meta["ox"], meta["oy"], dirty=rect)
            st, res = ctx.cycle(rx, ry, 0.3, meta["poly"], capi.FISHER_BEST_YAW)
            assert st == 0
            got = ctx.cycle_fetch(res.n_frontiers)
            if step % 4 == 0:  # a fresh context with the whole map gives the same bits
                ref = capi.Context()
                ref.set_costmap(a, meta["res"], meta["ox"], meta["oy"])
                ref.fisher_generate_lookup()
                ref.set_map_points(pts)
                st2, res2 = ref.cycle(rx, ry, 0.3, meta["poly"], capi.FISHER_BEST_YAW)
                want = ref.cycle_fetch(res2.n_frontiers)
                ref.close()
                assert st2 == 0 and res2.n_frontiers == res.n_frontiers
                for key in got:
                    assert got[key].tobytes() == want[key].tobytes(), (step, key)
            if step % 8 == 0:  # and the oracle agrees
                sto, fr, cells, seeds, cent = oracle.search_frontiers(a, meta, rx, ry)
                assert len(fr) == res.n_frontiers
                for key in ("goal_mx", "goal_my", "size", "component", "chunk", "first_cell"):
                    np.testing.assert_array_equal(got["frontiers"][key], fr[key], err_msg=key)
                arr, _, _ = oracle.score_arrival(a, meta, fr, meta["poly"])
                for key in REC_FIELDS:
                    np.testing.assert_array_equal(got["arrival"][key], arr[key], err_msg=key)
    finally:
        ctx.close()


def test_c5_thousand_updates_on_c2(capi, oracle, synth):
    """configs[4] at its real size: 1 000 successive dirty-rectangle updates on the 4096^2 map, a full cycle after each; the
    oracle checks the extraction and the arrival information every 50th step, a fresh context the whole cycle every 250th."""
    a, meta, pts, ov = synth.make_config("C2")
    a = a.copy()
    params = synth.apply_overrides(capi.default_params(), ov)
    synth.apply_overrides(oracle.params, ov)
    rng = np.random.default_rng(55)
    H, W = a.shape
    ctx = capi.Context(params)
    try:
        ctx.set_costmap(a, meta["res"], meta["ox"], meta["oy"])
        ctx.fisher_generate_lookup()
        ctx.set_map_points(pts)
        cx = int((meta["robot_x"] - meta["ox"]) / meta["res"])
        cy = int((meta["robot_y"] - meta["oy"]) / meta["res"])
        for step in range(1000):
            cx = int(np.clip(cx + rng.integers(-4, 5), 100, W - 100))
            cy = int(np.clip(cy + rng.integers(-4, 5), 100, H - 100))
            rect = synth.reveal_step(a, meta, cx, cy, rng)
            rx, ry = meta["ox"] + (cx + 0.5) * meta["res"], meta["oy"] + (cy + 0.5) * meta["res"]
            ctx.set_costmap(a, meta["res"], meta["ox"], meta["oy"], dirty=rect)
            st, res = ctx.cycle(rx, ry, 0.3, meta["poly"], capi.FISHER_BEST_YAW)
            assert st == 0
            if step % 50 == 49:
                got = ctx.cycle_fetch(res.n_frontiers)
                sto, fr, cells, seeds, cent = oracle.search_frontiers(a, meta, rx, ry, capacity=1 << 17, cell_capacity=1 << 22)
                assert len(fr) == res.n_frontiers, step
                for key in ("goal_mx", "goal_my", "size", "component", "chunk", "first_cell"):
                    np.testing.assert_array_equal(got["frontiers"][key], fr[key], err_msg=f"{step} {key}")
                pick = np.linspace(0, len(fr) - 1, 200).astype(np.int64)
                arr, _, _ = oracle.score_arrival(a, meta, fr[pick], meta["poly"])
                for key in REC_FIELDS:
                    np.testing.assert_array_equal(got["arrival"][key][pick], arr[key], err_msg=f"{step} {key}")
            if step % 250 == 249:
                got = ctx.cycle_fetch(res.n_frontiers)
                ref = capi.Context(params)
                ref.set_costmap(a, meta["res"], meta["ox"], meta["oy"])
                ref.fisher_generate_lookup()
                ref.set_map_points(pts)
                st2, res2 = ref.cycle(rx, ry, 0.3, meta["poly"], capi.FISHER_BEST_YAW)
                want = ref.cycle_fetch(res2.n_frontiers)
                ref.close()
                assert st2 == 0 and res2.n_frontiers == res.n_frontiers
                for key in got:
                    assert got[key].tobytes() == want[key].tobytes(), (step, key)
    finally:
        ctx.close()
