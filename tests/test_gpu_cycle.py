"""Whole-cycle parity: fitslam_cycle (search -> arrival -> euclidean plan -> Fisher -> U1 -> selection) against the same
chain on the oracle, and the sharded begin / all-gather / end form against the single-GPU result."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _oracle_chain(oracle, synth, a, meta, pts, ryaw, fisher):
    st, fr, cells, seeds, cent = oracle.search_frontiers(a, meta, meta["robot_x"], meta["robot_y"], capacity=1 << 18,
                                                         cell_capacity=1 << 23)
    arr, _, _ = oracle.score_arrival(a, meta, fr, meta["poly"])
    ach, pl, ph = oracle.plan_euclidean(meta["robot_x"], meta["robot_y"], ryaw, fr, arr["achievable"])
    max_gt = oracle.arrival_limits(a, meta)[0]
    st, wc, ua, ud, err = oracle.normalise(arr, ach, pl, ph, None, max_gt)
    assert st == 0
    srt, sel = oracle.select(ach, pl)
    finfo = None
    if fisher:
        oracle.fisher_generate()
        theta = arr["theta_s_star"]
        poses = synth.frontier_poses(fr, theta, oracle.params.candidate_pose_z)
        q = np.array([oracle.orientation_around_z(t) for t in theta])
        poses["qx"], poses["qy"], poses["qz"], poses["qw"] = q.T
        st, finfo, fsafe, fvis, fused = oracle.fisher_poses(pts, poses)
    return dict(fr=fr, arr=arr, ach=ach, pl=pl, ph=ph, wc=wc, srt=srt, sel=sel, fisher=finfo, n_cells=len(cells),
                n_comp=len(seeds))


def _check_against(capi, got, res, want, fisher):
    F = len(want["fr"])
    assert res.n_frontiers == F and res.n_frontier_cells == want["n_cells"] and res.n_components == want["n_comp"]
    for f in ("goal_mx", "goal_my", "size", "component", "chunk", "first_cell"):
        np.testing.assert_array_equal(got["frontiers"][f], want["fr"][f], err_msg=f)
    for f in ("information", "max_index", "achievable", "footprint_in_lethal", "ray_failed"):
        np.testing.assert_array_equal(got["arrival"][f], want["arr"][f], err_msg=f)
    np.testing.assert_array_equal(got["arrival"]["theta_s_star"].view(np.uint64), want["arr"]["theta_s_star"].view(np.uint64))
    np.testing.assert_array_equal(got["path_length_m"].view(np.uint64), want["pl"].view(np.uint64))  # sqrt: exact
    # path_heading: atan2(goal - robot) is evaluated by the host's libm, as the reference's; everything after it is IEEE
    np.testing.assert_array_equal(got["path_heading"].view(np.uint64), want["ph"].view(np.uint64))
    np.testing.assert_array_equal(got["weighted_cost"].view(np.uint64), want["wc"].view(np.uint64))
    sel, wsel = res.selection, want["sel"]
    for f in ("n_sorted", "n_local", "global_begin", "n_global", "closest_global"):
        assert getattr(sel, f) == getattr(wsel, f), f
    np.testing.assert_array_equal(got["sorted"][:sel.n_sorted], want["srt"])
    if fisher:
        w = want["fisher"]
        assert np.all(np.abs(got["fisher"] - w) <= 1e-4 * np.abs(w) + 1e-6)
    # ambiguous_sorts counts the chunks whose median sort went to the host (an angle whose last bit glibc may round the other
    # way, next to another angle or a threshold of the comparator): a diagnostic, the goal cells above are what must match
    assert res.ambiguous_sorts <= res.n_frontiers


@pytest.mark.parametrize("name,fisher", [("C1", 0), ("C1", 1), ("C3", 0), ("C2", 1)])
def test_cycle_matches_oracle_chain(capi, oracle, synth, name, fisher):
    a, meta, pts, ov = synth.make_config(name, reference_rays=True)
    synth.apply_overrides(oracle.params, ov)
    ryaw = 0.7
    want = _oracle_chain(oracle, synth, a, meta, pts, ryaw, fisher)
    ctx = capi.Context(oracle.params)
    ctx.set_costmap(a, meta["res"], meta["ox"], meta["oy"])
    if fisher:
        ctx.fisher_generate_lookup()
        ctx.set_map_points(pts)
    for _ in range(2):
        st, res = ctx.cycle(meta["robot_x"], meta["robot_y"], ryaw, meta["poly"], fisher)
        assert st == 0
        got = ctx.cycle_fetch(res.n_frontiers)
        _check_against(capi, got, res, want, fisher)
    assert sum(res.stage_ms[:5]) > 0
    ctx.close()


def test_cycle_all_yaws_record_is_best_yaw(capi, oracle, synth):
    a, meta, pts, ov = synth.make_config("C1")  # BASELINE shape: 64 rays, 8 yaw windows
    synth.apply_overrides(oracle.params, ov)
    ctx = capi.Context(oracle.params)
    ctx.set_costmap(a, meta["res"], meta["ox"], meta["oy"])
    ctx.fisher_generate_lookup()
    ctx.set_map_points(pts)
    st, r1 = ctx.cycle(meta["robot_x"], meta["robot_y"], 0.0, meta["poly"], capi.FISHER_BEST_YAW)
    best = ctx.cycle_fetch(r1.n_frontiers)["fisher"].copy()
    st, r2 = ctx.cycle(meta["robot_x"], meta["robot_y"], 0.0, meta["poly"], capi.FISHER_ALL_YAWS)
    allyaw = ctx.cycle_fetch(r2.n_frontiers)["fisher"]
    assert r2.n_fisher_poses == r2.n_frontiers * ctx.ray_shape()[2] and r1.n_fisher_poses == r1.n_frontiers
    ok = ctx.cycle_fetch(r2.n_frontiers)["arrival"]["ray_failed"] == 0
    np.testing.assert_array_equal(best[ok], allyaw[ok])
    ctx.close()


@pytest.mark.parametrize("name,n_check", [("C1", 0), ("C2", 256)])
def test_cycle_all_yaws_candidates_match_oracle(capi, oracle, synth, name, n_check):
    """The benchmarked mode (FISHER_ALL_YAWS on the BASELINE ray shape): EVERY candidate's Fisher information, read through
    fitslam_cycle_fetch_candidates, against the oracle's isPoseSafe for the pose (frontier goal, yaw of the window):
    1e-4 relative, no allowance for the number of landmarks.  C2: the first 256 frontiers (4 096 poses x 200 k points)."""
    a, meta, pts, ov = synth.make_config(name)  # 64 rays; 8 (C1) / 16 (C2) yaw windows
    synth.apply_overrides(oracle.params, ov)
    ctx = capi.Context(oracle.params)
    ctx.set_costmap(a, meta["res"], meta["ox"], meta["oy"])
    ctx.fisher_generate_lookup()
    ctx.set_map_points(pts)
    st, res = ctx.cycle(meta["robot_x"], meta["robot_y"], 0.0, meta["poly"], capi.FISHER_ALL_YAWS)
    assert st == 0
    Y = ctx.ray_shape()[2]
    assert res.n_fisher_poses == res.n_frontiers * Y and Y == (8 if name == "C1" else 16)
    got = ctx.cycle_fetch(res.n_frontiers)
    info, safe, nvis = ctx.cycle_fetch_candidates(res.n_fisher_poses)
    assert len(info) == res.n_fisher_poses
    F = n_check or res.n_frontiers
    fr = got["frontiers"][:F]
    oracle.fisher_generate()
    th0 = np.arange(Y) * oracle.params.delta_theta + oracle.params.camera_fov / 2   # setGoalOrientation of window w
    th = np.tile(th0, F)
    poses = synth.frontier_poses(np.repeat(fr, Y), th, oracle.params.candidate_pose_z)
    q = np.array([oracle.orientation_around_z(t) for t in th])
    poses["qx"], poses["qy"], poses["qz"], poses["qw"] = q.T
    st, winfo, wsafe, wvis, wused = oracle.fisher_poses(pts, poses)
    ok = np.repeat(got["arrival"]["ray_failed"][:F] == 0, Y)
    g = info[:F * Y]
    assert np.all(np.abs(g - winfo)[ok] <= 1e-4 * np.abs(winfo[ok]) + 1e-6), np.max(np.abs(g - winfo)[ok] / np.maximum(np.abs(winfo[ok]), 1))
    np.testing.assert_array_equal(nvis[:F * Y][ok], wused[ok])
    thr = oracle.params.fisher_information_threshold
    clear = np.abs(winfo - thr) > 1e-3 * thr
    np.testing.assert_array_equal(safe[:F * Y][ok & clear], wsafe[ok & clear])
    # the record of a frontier carries the information of its best yaw
    best = info.reshape(-1, Y)[np.arange(res.n_frontiers), got["arrival"]["max_index"]]
    okf = got["arrival"]["ray_failed"] == 0
    np.testing.assert_array_equal(best[okf], got["fisher"][okf])
    ctx.close()


@pytest.mark.parametrize("nranks", [2, 3])
def test_sharded_cycle_equals_single(capi, oracle, synth, nranks):
    """N ranks emulated in one process on one GPU: each context scores its share (frontier f on rank f % N), the records
    are merged the way the NCCL all-reduce does (integer sum of the zero-padded buffers), then every rank finishes and
    must match the single-context cycle."""
    import torch
    a, meta, pts, ov = synth.make_config("C1")
    synth.apply_overrides(oracle.params, ov)
    args = (meta["robot_x"], meta["robot_y"], 0.3, meta["poly"], capi.FISHER_BEST_YAW)

    def make(rank, n):
        c = capi.Context(oracle.params)
        c.set_shard(rank, n)
        c.set_costmap(a, meta["res"], meta["ox"], meta["oy"])
        c.fisher_generate_lookup()
        c.set_map_points(pts)
        return c
    one = make(0, 1)
    st, res = one.cycle(*args)
    want = one.cycle_fetch(res.n_frontiers)
    ranks = [make(r, nranks) for r in range(nranks)]
    for c in ranks:
        st, F = c.cycle_begin(*args)
        assert st == 0 and F == res.n_frontiers
    torch.cuda.synchronize()
    recs = [c.records_tensor() for c in ranks]
    assert all(r.shape[0] == F for r in recs)
    f = torch.arange(F, device="cuda")
    for r in range(nranks):
        assert not recs[r][f % nranks != r].any()  # nothing scored outside the rank's share
    merged = torch.stack(recs).sum(0, dtype=torch.int32)
    for r in range(nranks):
        recs[r].copy_(merged)
    torch.cuda.synchronize()
    for c in ranks:
        r2 = c.cycle_end()
        got = c.cycle_fetch(r2.n_frontiers)
        for k in ("arrival", "path_length_m", "path_heading", "weighted_cost", "fisher", "sorted"):
            assert got[k].tobytes() == want[k].tobytes(), k
        assert bytes(r2.selection) == bytes(res.selection)
    for c in ranks + [one]:
        c.close()


def test_cycle_with_host_sorted_chunks(capi, oracle, synth, monkeypatch):
    # the frontiers whose median sort the host redoes are scored by the launches after the search (TAG_HOST -> TAG_FINAL):
    # with every chunk forced down that path the whole cycle still equals the oracle's chain
    monkeypatch.setenv("FITSLAM_FIX_ALL", "1")
    a, meta, pts, ov = synth.make_config("C1", reference_rays=True)
    synth.apply_overrides(oracle.params, ov)
    want = _oracle_chain(oracle, synth, a, meta, pts, 0.7, 1)
    ctx = capi.Context(oracle.params)
    ctx.set_costmap(a, meta["res"], meta["ox"], meta["oy"])
    ctx.fisher_generate_lookup()
    ctx.set_map_points(pts)
    st, res = ctx.cycle(meta["robot_x"], meta["robot_y"], 0.7, meta["poly"], 1)
    assert st == 0 and res.ambiguous_sorts == res.n_frontiers
    _check_against(capi, ctx.cycle_fetch(res.n_frontiers), res, want, 1)
    ctx.close()


@pytest.mark.parametrize("name", ["C1", "C3"])
def test_cycle_matches_the_compiled_reference(capi, oracle, synth, name):
    """The device cycle against the REFERENCE'S OWN code, no oracle in between: FrontierSearch::searchFrom, then
    FrontierCostsManager::assignCosts({"ArrivalInformation", "EuclideanDistance"}) and FullPathOptimizer::getFilteredFrontiersN,
    all compiled from the reference sources into oracle/_ref (oracle/Makefile).  Goal points, arrival information, achievability,
    path lengths, headings, weighted costs and the selection, bit for bit."""
    from oracle import ref_binding
    if not ref_binding.available():
        pytest.skip("oracle/_ref is built where /root/reference exists")
    ref = ref_binding.Reference()
    a, meta, pts, ov = synth.make_config(name, reference_rays=True)
    params = synth.apply_overrides(capi.default_params(), ov)
    ryaw = 0.7
    rx, ry = meta["robot_x"], meta["robot_y"]
    q = oracle.orientation_around_z(ryaw)
    _, _, ref_yaw = ref.plan_euclidean(rx, ry, q, np.zeros((0, 2)), np.zeros(0, np.uint8), np.zeros(0))
    n, rfr, rcells = ref.search_frontiers(a, meta, params, rx, ry, capacity=1 << 17, cell_capacity=1 << 22)
    assert n > 50
    st, rows, loc, glo, closest = ref.assign_costs(a, meta, params, rfr[:, :2], rfr[:, 2], meta["poly"], rx, ry, q, None)
    assert st == 0
    ctx = capi.Context(params)
    try:
        ctx.set_costmap(a, meta["res"], meta["ox"], meta["oy"])
        st, res = ctx.cycle(rx, ry, ref_yaw, meta["poly"], capi.FISHER_OFF)  # the yaw the reference extracts from the quaternion
        assert st == 0 and res.n_frontiers == n
        got = ctx.cycle_fetch(n)
        fr = got["frontiers"]
        np.testing.assert_array_equal(fr["goal_x"].view(np.uint64), rfr[:, 0].copy().view(np.uint64))
        np.testing.assert_array_equal(fr["goal_y"].view(np.uint64), rfr[:, 1].copy().view(np.uint64))
        np.testing.assert_array_equal(fr["size"].astype(np.float64), rfr[:, 2])
        np.testing.assert_array_equal(got["weighted_cost"].view(np.uint64), rows[:, 0].copy().view(np.uint64))
        np.testing.assert_array_equal(got["arrival"]["information"].astype(np.float64), rows[:, 4])
        ach = rows[:, 3].astype(bool)
        # (the reference's achievable flag after the plan step; the record keeps arrival and plan achievability apart)
        np.testing.assert_array_equal(np.isfinite(got["path_length_m"]) & (got["path_length_m"] < 1e300), ach)
        np.testing.assert_array_equal(got["path_length_m"].view(np.uint64), rows[:, 5].copy().view(np.uint64))
        np.testing.assert_array_equal(got["path_heading"].view(np.uint64), rows[:, 6].copy().view(np.uint64))
        sel = res.selection
        np.testing.assert_array_equal(got["sorted"][:sel.n_local].astype(np.int32), loc)
        np.testing.assert_array_equal(got["sorted"][sel.global_begin:sel.global_begin + sel.n_global].astype(np.int32), glo)
        assert sel.closest_global == closest
    finally:
        ctx.close()


def test_cycle_after_a_much_smaller_one(capi, oracle, synth):
    """The Fisher launches that run during the search are sized from what the same launch found in the previous cycle.  A cycle
    that finds far more frontiers than the one before leaves the excess to the launches after the search: every candidate is
    still scored once, and the results equal a fresh context's bit for bit."""
    a, meta, pts, ov = synth.make_config("C1")
    synth.apply_overrides(oracle.params, ov)
    H, W = a.shape
    walled = a.copy()  # a lethal box close around the robot: the search stays inside it and finds few frontiers
    ry, rx = int((meta["robot_y"] - meta["oy"]) / meta["res"]), int((meta["robot_x"] - meta["ox"]) / meta["res"])
    r = min(H, W) // 16
    walled[max(ry - r, 0):ry + r, max(rx - r, 0)] = 254
    walled[max(ry - r, 0):ry + r, min(rx + r, W - 1)] = 254
    walled[max(ry - r, 0), max(rx - r, 0):rx + r] = 254
    walled[min(ry + r, H - 1), max(rx - r, 0):rx + r + 1] = 254

    def run(ctx, grid):
        ctx.set_costmap(grid, meta["res"], meta["ox"], meta["oy"])
        st, res = ctx.cycle(meta["robot_x"], meta["robot_y"], 0.4, meta["poly"], capi.FISHER_ALL_YAWS)
        assert st in (0, capi.E_NO_FRONTIERS)
        if st != 0:
            return res, None, None
        return res, ctx.cycle_fetch(res.n_frontiers), ctx.cycle_fetch_candidates(res.n_fisher_poses)

    fresh = capi.Context(oracle.params)
    fresh.fisher_generate_lookup()
    fresh.set_map_points(pts)
    want_res, want, want_c = run(fresh, a)
    fresh.close()
    ctx = capi.Context(oracle.params)
    ctx.fisher_generate_lookup()
    ctx.set_map_points(pts)
    for _ in range(2):
        small_res, _, _ = run(ctx, walled)
    assert 0 < small_res.n_frontiers * 4 < want_res.n_frontiers
    for _ in range(2):  # the first one runs on the small cycle's sizes, the second on its own
        res, got, got_c = run(ctx, a)
        assert res.n_frontiers == want_res.n_frontiers
        for k in ("frontiers", "arrival", "fisher", "weighted_cost", "sorted"):
            np.testing.assert_array_equal(got[k], want[k], err_msg=k)
        for g, w in zip(got_c, want_c):
            np.testing.assert_array_equal(g, w)
    ctx.close()
