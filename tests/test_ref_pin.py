"""Pins the oracle to the REFERENCE ITSELF: oracle/_ref/libfitslam_ref.so is DEP/src/FrontierSearch.cpp, Frontier.cpp, the
costmap half of Helpers.cpp and the member functions of CostCalculator.cpp, FrontierCostsManager.cpp and FullPathOptimizer.cpp
that make up stages R and N, compiled from /root/reference where it lies (oracle/Makefile) against stand-in ROS headers.
Every comparison is exact: goal points, every_frontier coordinates, path lengths, headings and weighted costs bit for bit,
sizes, orders, ray counts, selections.
Rows pinned: S1-S6 (searchFrom, nearestFreeCell, isNewFrontierCell, buildNewFrontier, centroid + median sort
with glibc atan2, getAllFrontiers), R1-R6 (setArrivalInformationForFrontier, getTracedCells, bresenham2D,
isRobotFootprintInLethal, setMaxArrivalInformation, recomputeNormalizationFactors), T4 (RayTracedCells), N1-N3 (assignCosts with
the U1 block, setPlanForFrontierEuclidean, getFilteredFrontiersN), F3-F7 (the Fisher helpers, generateLookupTable / loadLookupTable,
getVoxelCoordinate, getInformationFromLookup with the crowding factor, isPoseSafe: FisherInfoManager.cpp compiled whole); next row
8f-3 (LethalMarker: the ray / circle / triangle functions, addNewMarkedArea, addNewMarkedAreaFOV, updateCosts; the threshold node's
callback) and next rows
8f-2 (FrontierRoadMap::isConnectable) and 8f-4 (SensorSimulator::markRay) against the roadmap-explorer package's own code in
oracle/_ref/libfitslam_ref_re.so; rows F1, F2 and the keyframe mode of 8f-1 in control flow (the ORB-SLAM3 wrapper's functions against
stand-in ORB_SLAM3 / Sophus types, oracle/_ref/libfitslam_ref_orbw.so; see the section at the end)."""
import math

import numpy as np
import pytest

from oracle import ref_binding

pytestmark = pytest.mark.skipif(not ref_binding.available(), reason="oracle/_ref is built where /root/reference exists")


@pytest.fixture(scope="module")
def ref():
    return ref_binding.Reference()


def _cells_world(cells, a, meta):
    W = a.shape[1]
    my, mx = cells // W, cells % W
    return np.stack([meta["ox"] + (mx + 0.5) * meta["res"], meta["oy"] + (my + 0.5) * meta["res"]], axis=1)


def _same_search(oracle, ref, a, meta, rx, ry, ov=None):
    for k, v in (ov or {}).items():
        setattr(oracle.params, k, v)
    st, fr, cells, seeds, cent = oracle.search_frontiers(a, meta, rx, ry, capacity=1 << 17, cell_capacity=1 << 22)
    n, rfr, rcells = ref.search_frontiers(a, meta, oracle.params, rx, ry, capacity=1 << 17, cell_capacity=1 << 22)
    if n < 0:
        assert st != 0 and len(fr) == 0
        return 0
    assert st == 0 and n == len(fr), (n, len(fr))
    np.testing.assert_array_equal(fr["goal_x"].view(np.uint64), rfr[:, 0].copy().view(np.uint64))
    np.testing.assert_array_equal(fr["goal_y"].view(np.uint64), rfr[:, 1].copy().view(np.uint64))
    np.testing.assert_array_equal(fr["size"].astype(np.float64), rfr[:, 2])
    assert len(cells) == len(rcells)
    np.testing.assert_array_equal(_cells_world(cells.astype(np.int64), a, meta).view(np.uint64), rcells.copy().view(np.uint64))
    return n


@pytest.mark.parametrize("name", ["C1", "C3"])
def test_reference_search_on_configs(oracle, ref, synth, name):
    a, meta, pts, ov = synth.make_config(name, reference_rays=True)
    n = _same_search(oracle, ref, a, meta, meta["robot_x"], meta["robot_y"], dict(max_frontier_distance=ov["max_frontier_distance"]))
    assert n > 100


@pytest.mark.parametrize("seed", [1, 2, 3, 4])
def test_reference_search_inflated_small_radius(oracle, ref, synth, seed):
    a, meta = synth.make_costmap(701, 577, seed=seed, n_disc=12, r_lo=25, r_hi=70, n_holes=30, inflation=5 if seed % 2 else 0)
    assert _same_search(oracle, ref, a, meta, meta["robot_x"], meta["robot_y"], dict(max_frontier_distance=6.0 + 3 * seed)) > 0


def test_reference_search_cluster_size_parameters(oracle, ref, synth):
    a, meta = synth.make_costmap(600, 600, seed=5, n_disc=10, r_lo=30, r_hi=80, n_holes=20)
    for mn, mx in ((1, 20), (3, 7), (0, 40), (5, 62)):
        assert _same_search(oracle, ref, a, meta, meta["robot_x"], meta["robot_y"],
                            dict(min_frontier_cluster_size=mn, max_frontier_cluster_size=mx)) > 0


def test_reference_search_random_noise(oracle, ref):
    rng = np.random.default_rng(7)
    for W, H in ((64, 48), (97, 131), (200, 200)):
        for k in range(6):
            a = rng.choice(np.array([0, 0, 0, 100, 200, 253, 254, 255, 255], np.uint8), size=(H, W))
            meta = dict(res=0.05 if k % 2 else 0.1, ox=-1.3 * k, oy=0.7 * k)
            rx = meta["ox"] + (W // 2 + 0.3) * meta["res"]
            ry = meta["oy"] + (H // 2 + 0.6) * meta["res"]
            _same_search(oracle, ref, a, meta, rx, ry, dict(max_frontier_distance=2.0 + k))


@pytest.mark.parametrize("res", [0.05, 0.1, 1.0 / 16])
def test_reference_straight_frontiers_median_sort(oracle, ref, res):
    # straight axis-aligned frontiers found from their middle: two-branch chunks of collinear cells, whose median sort
    # compares atan2 values that differ in the last place at 0.05 m (the case VERDICT r1 "What's weak" #2 names)
    for W, H, horizontal in ((400, 40, True), (40, 400, False), (3000, 12, True)):
        a = np.full((H, W), 255, np.uint8)
        if horizontal:
            a[H // 2:, :] = 0
            a[H // 2 + 4:, :] = 254
        else:
            a[:, W // 2:] = 0
            a[:, W // 2 + 4:] = 254
        meta = dict(res=res, ox=-W * res / 2, oy=-H * res / 2)
        cy, cx = (H // 2 + 1, W // 2) if horizontal else (H // 2, W // 2 + 1)
        rx, ry = meta["ox"] + (cx + 0.5) * res, meta["oy"] + (cy + 0.5) * res
        assert _same_search(oracle, ref, a, meta, rx, ry, dict(max_frontier_distance=500.0)) >= W * H // (40 * 21) - 2


def test_reference_robot_on_lethal_unknown_and_off_map(oracle, ref, synth):
    a, meta = synth.make_costmap(300, 300, seed=9, n_disc=6, r_lo=20, r_hi=50, n_holes=5)
    H, W = a.shape
    b = a.copy()
    b[H // 2 - 3:H // 2 + 4, W // 2 - 3:W // 2 + 4] = 254       # nearestFreeCell has to walk out of a lethal block
    _same_search(oracle, ref, b, meta, meta["robot_x"], meta["robot_y"])
    b = a.copy()
    b[H // 2 - 2:H // 2 + 3, W // 2 - 2:W // 2 + 3] = 255       # ... of an unknown block
    _same_search(oracle, ref, b, meta, meta["robot_x"], meta["robot_y"])
    _same_search(oracle, ref, a, meta, meta["ox"] - 1.0, meta["robot_y"])    # off the map: empty list
    _same_search(oracle, ref, a, meta, meta["ox"] + W * meta["res"] + 0.01, meta["robot_y"])


def test_reference_nearest_free_cell_and_neighbourhoods(oracle, ref):
    rng = np.random.default_rng(3)
    for _ in range(20):
        W, H = int(rng.integers(3, 40)), int(rng.integers(3, 40))
        a = rng.choice(np.array([0, 159, 160, 254, 255], np.uint8), size=(H, W), p=[0.1, 0.1, 0.3, 0.3, 0.2])
        start = int(rng.integers(0, W * H))
        assert oracle.nearest_free_cell(a, start, 160) == ref.nearest_free_cell(a, start, 160)
    # corner, edge and interior cells: the order of nhood4 / nhood8 (DEP/src/Helpers.cpp:185-262)
    W, H = 7, 5
    for idx in (0, 3, 6, 7, 17, 28, 31, 34):
        x, y = idx % W, idx // W
        want4 = [i for ok, i in ((x > 0, idx - 1), (x < W - 1, idx + 1), (y > 0, idx - W), (y < H - 1, idx + W)) if ok]
        assert ref.nhood(W, H, idx, False) == want4
        assert ref.nhood(W, H, idx, True)[:len(want4)] == want4


def _ray_ends(params, meta, a, gx, gy, poly):
    """End points of the ray fan of setArrivalInformationForFrontier (DEP/src/CostCalculator.cpp:36-48)."""
    H, W = a.shape
    size_x, size_y = (W - 1 + 0.5) * meta["res"], (H - 1 + 0.5) * meta["res"]
    ends = []
    theta = 0.0
    while theta <= 2 * math.pi:
        wx = gx + params.max_camera_depth * math.cos(theta)
        wy = gy + params.max_camera_depth * math.sin(theta)
        wx = max(poly[0], max(meta["ox"], min(poly[2], min(meta["ox"] + size_x, wx))))
        wy = max(poly[1], max(meta["oy"], min(poly[3], min(meta["oy"] + size_y, wy))))
        ends.append((wx, wy))
        theta += params.delta_theta
    return ends


def test_reference_traced_cells_and_footprint(oracle, ref, synth):
    """Per-ray unknown-cell counts of the oracle's arrival stage == the reference's getTracedCells + RayTracedCells on the
    same rays; its footprint flag == isRobotFootprintInLethal."""
    a, meta = synth.make_costmap(400, 400, seed=11, n_disc=8, r_lo=20, r_hi=60, n_holes=15, inflation=3)
    st, fr, cells, seeds, cent = oracle.search_frontiers(a, meta, meta["robot_x"], meta["robot_y"])
    assert st == 0 and len(fr) > 20
    poly = [meta["ox"], meta["oy"], meta["ox"] + a.shape[1] * meta["res"], meta["oy"] + a.shape[0] * meta["res"]]
    arr, rays, wins = oracle.score_arrival(a, meta, fr, poly)
    max_length = float(int(oracle.params.max_camera_depth / meta["res"]))
    radius_cells = math.ceil(oracle.params.robot_radius / meta["res"])
    for f in range(0, len(fr), max(1, len(fr) // 40)):
        gx, gy = float(fr["goal_x"][f]), float(fr["goal_y"][f])
        ends = _ray_ends(oracle.params, meta, a, gx, gy, poly)
        assert len(ends) == rays.shape[1]
        for i, (wx, wy) in enumerate(ends):
            ok, pushed, hit, unknown, allc = ref.traced_cells(a, meta, gx, gy, wx, wy, max_length)
            assert ok and pushed == int(rays[f, i]), (f, i, pushed, int(rays[f, i]))
        assert ref.footprint_in_lethal(a, meta, int(fr["goal_mx"][f]), int(fr["goal_my"][f]), radius_cells) == bool(arr["footprint_in_lethal"][f])


def test_reference_roadmap_rays(oracle, ref, synth):
    """FrontierRoadMap::isConnectable's ray (obstacle band 253..254, every cell traced) against the oracle's restatement."""
    a, meta = synth.make_costmap(300, 300, seed=13, n_disc=6, r_lo=20, r_hi=50, n_holes=10, inflation=2)
    rng = np.random.default_rng(5)
    H, W = a.shape
    seg = np.stack([meta["ox"] + rng.uniform(1, W - 1, 300) * meta["res"], meta["oy"] + rng.uniform(1, H - 1, 300) * meta["res"],
                    meta["ox"] + rng.uniform(1, W - 1, 300) * meta["res"], meta["oy"] + rng.uniform(1, H - 1, 300) * meta["res"]], axis=1)
    radius = 2.6
    ok_o, unk_o = oracle.roadmap_connectable(a, meta, seg, radius)
    for k, (sx, sy, ex, ey) in enumerate(seg):
        ok, pushed, hit, unknown, allc = ref.traced_cells(a, meta, sx, sy, ex, ey, float(int(1.5 * radius / meta["res"])) if False else 1.5 * radius / meta["res"],
                                                          obstacle=(253, 254), trace=(0, 255))
        assert ok
        want = (not hit) and unknown <= radius / meta["res"] * 0.3
        assert bool(ok_o[k]) == want, (k, hit, unknown)


# ---- stage R and the Euclidean plan against the reference's own CostCalculator.cpp member functions -----------------------------
def _arrival_vs_reference(oracle, ref, a, meta, fr, poly):
    arr, rays, wins = oracle.score_arrival(a, meta, fr, poly)
    rows, (ret, max_gt, min_gt) = ref.score_arrival(a, meta, oracle.params, np.stack([fr["goal_x"], fr["goal_y"]], axis=1), fr["size"], poly)
    omax, omin, oval = oracle.arrival_limits(a, meta)
    assert (ret, max_gt, min_gt) == (float(oval), omax, omin)  # setMaxArrivalInformation: return value, * 1.2, * 0.70 (bit for bit)
    np.testing.assert_array_equal(rows[:, 0], arr["information"].astype(np.float64))
    np.testing.assert_array_equal(rows[:, 1].astype(np.uint8), arr["achievable"])
    # setGoalOrientation((maxIndex * DELTA_THETA) + (CAMERA_FOV / 2)), or 0.0 after a failed ray
    theta = np.where(arr["ray_failed"] != 0, 0.0, arr["max_index"] * oracle.params.delta_theta + oracle.params.camera_fov / 2)
    want_q = np.array([oracle.orientation_around_z(t) for t in theta])
    np.testing.assert_array_equal(rows[:, 2:6].copy().view(np.uint64), want_q.view(np.uint64))
    return arr


@pytest.mark.parametrize("name,reference_rays", [("C1", True), ("C1", False), ("C3", True)])
def test_reference_arrival_information_on_configs(oracle, ref, synth, capi, name, reference_rays):
    a, meta, pts, ov = synth.make_config(name, reference_rays=reference_rays)
    synth.apply_overrides(oracle.params, ov)
    st, fr, cells, seeds, cent = oracle.search_frontiers(a, meta, meta["robot_x"], meta["robot_y"])
    assert st == 0 and len(fr) > 50
    arr = _arrival_vs_reference(oracle, ref, a, meta, fr, meta["poly"])
    assert arr["information"].max() > 0 and (arr["achievable"] == 0).any() and (arr["achievable"] == 1).any()
    # a boundary polygon that cuts the rays short, and one that lies off the frontiers
    box = synth.explored_bbox(a, meta)
    cx, cy = 0.5 * (box[0] + box[2]), 0.5 * (box[1] + box[3])
    _arrival_vs_reference(oracle, ref, a, meta, fr, [cx - 6.0, cy - 4.0, cx + 5.0, cy + 7.0])
    _arrival_vs_reference(oracle, ref, a, meta, fr[:64], [box[2] + 1.0, box[3] + 1.0, box[2] + 3.0, box[3] + 3.0])


def test_reference_arrival_information_near_lethal_and_borders(oracle, ref, synth, capi):
    # goals next to lethal cells (footprint penalty: small frontiers become unachievable), goals whose rays leave the map
    a, meta = synth.make_costmap(700, 600, seed=77, n_disc=14, r_lo=30, r_hi=90, n_holes=80)
    st, fr, cells, seeds, cent = oracle.search_frontiers(a, meta, meta["robot_x"], meta["robot_y"])
    assert st == 0 and len(fr) > 20
    H, W = a.shape
    # keep the footprint disc inside the array (the reference does not bounds-check it: DESIGN.md section 1)
    r = int(math.ceil(oracle.params.robot_radius / meta["res"])) + 1
    ok = (fr["goal_mx"] >= r) & (fr["goal_mx"] < W - r) & (fr["goal_my"] >= r) & (fr["goal_my"] < H - r)
    arr = _arrival_vs_reference(oracle, ref, a, meta, fr[ok], meta["poly"])
    assert arr["footprint_in_lethal"].any()


def test_reference_plan_euclidean_and_normalisation_factors(oracle, ref, synth, capi):
    a, meta, pts, ov = synth.make_config("C1")
    synth.apply_overrides(oracle.params, ov)
    st, fr, cells, seeds, cent = oracle.search_frontiers(a, meta, meta["robot_x"], meta["robot_y"])
    arr, _, _ = oracle.score_arrival(a, meta, fr, meta["poly"])
    rng = np.random.default_rng(8)
    goals = np.stack([fr["goal_x"], fr["goal_y"]], axis=1)
    for yaw_in in [0.0, 0.4, -2.9, 3.1, 1.5707963267948966, -1e-9] + list(rng.uniform(-np.pi, np.pi, 6)):
        q = oracle.orientation_around_z(yaw_in)
        # the robot at the map's robot pose, and right on top of a frontier (length < 0.5 m: not achievable)
        for rx, ry in ((meta["robot_x"], meta["robot_y"]), (float(fr["goal_x"][3]) + 0.1, float(fr["goal_y"][3]))):
            rows, norm, ref_yaw = ref.plan_euclidean(rx, ry, q, goals, arr["achievable"], arr["information"].astype(np.float64))
            # the oracle takes the yaw the reference extracts from the quaternion (tf2 getRPY)
            ach, pl, ph = oracle.plan_euclidean(rx, ry, ref_yaw, fr, arr["achievable"])
            np.testing.assert_array_equal(rows[:, 0].astype(np.uint8), ach)
            np.testing.assert_array_equal(rows[:, 1].view(np.uint64), pl.view(np.uint64))
            np.testing.assert_array_equal(rows[:, 2].view(np.uint64), ph.view(np.uint64))
            # recomputeNormalizationFactors over the list: min / max path length and min arrival information of the achievable ones
            m = ach != 0
            assert norm == (pl[m].min(), pl[m].max(), float(arr["information"][m].min()))
            assert abs(ref_yaw - math.atan2(math.sin(yaw_in), math.cos(yaw_in))) < 1e-12


# ---- stage N against the reference's own FrontierCostsManager::assignCosts and FullPathOptimizer::getFilteredFrontiersN -----------
def _oracle_chain(oracle, a, meta, fr, poly, rx, ry, yaw, blacklisted):
    arr, _, _ = oracle.score_arrival(a, meta, fr, poly)
    ach, pl, ph = oracle.plan_euclidean(rx, ry, yaw, fr, arr["achievable"])
    max_gt = oracle.arrival_limits(a, meta)[0]
    st, wc, ua, ud, err = oracle.normalise(arr, ach, pl, ph, blacklisted, max_gt)
    eligible = ach.astype(bool) if blacklisted is None else (ach.astype(bool) & ~np.asarray(blacklisted, bool))
    srt, sel = oracle.select(eligible, pl)
    return arr, ach, pl, ph, st, wc, ua, ud, srt, sel


@pytest.mark.parametrize("name,with_blacklist", [("C1", False), ("C1", True), ("C3", False)])
def test_reference_assign_costs_and_selection(oracle, ref, synth, capi, name, with_blacklist):
    a, meta, pts, ov = synth.make_config(name, reference_rays=True)
    synth.apply_overrides(oracle.params, ov)
    st, fr, cells, seeds, cent = oracle.search_frontiers(a, meta, meta["robot_x"], meta["robot_y"])
    assert st == 0 and len(fr) > 50
    rng = np.random.default_rng(3)
    bl = (rng.random(len(fr)) < 0.15).astype(np.uint8) if with_blacklist else None
    for yaw_in in (0.4, -2.2):
        q = oracle.orientation_around_z(yaw_in)
        rx, ry = meta["robot_x"], meta["robot_y"]
        _, _, ref_yaw = ref.plan_euclidean(rx, ry, q, np.zeros((0, 2)), np.zeros(0, np.uint8), np.zeros(0))
        arr, ach, pl, ph, sto, wc, ua, ud, srt, sel = _oracle_chain(oracle, a, meta, fr, meta["poly"], rx, ry, ref_yaw, bl)
        st, rows, loc, glo, closest = ref.assign_costs(a, meta, oracle.params, np.stack([fr["goal_x"], fr["goal_y"]], axis=1), fr["size"],
                                                       meta["poly"], rx, ry, q, bl)
        if with_blacklist:
            # The reference itself cannot finish this call: a blacklisted frontier skips the plan step (FrontierCostsManager.cpp:
            # 75-85), stays "achievable", and the U1 loop then reads its path heading, which was never set:
            # Frontier::getPathHeading throws "Path heading frontier property is null" (Frontier.cpp:194-199).  The oracle and the
            # device give such frontiers weighted_cost = DBL_MAX and carry on (DESIGN.md section 1 lists the deviation).
            assert st == 4 and sto == 0
            assert np.all(wc[bl.astype(bool)] == np.finfo(np.float64).max)
            continue
        assert st == 0 and sto == 0
        blm = np.zeros(len(fr), bool) if bl is None else bl.astype(bool)
        # weighted cost of every frontier, bit for bit (DBL_MAX for unachievable and blacklisted ones)
        np.testing.assert_array_equal(rows[:, 0].copy().view(np.uint64), wc.view(np.uint64))
        m = ~blm
        np.testing.assert_array_equal(rows[m, 1].copy().view(np.uint64), ua[m].view(np.uint64))
        np.testing.assert_array_equal(rows[m, 2].copy().view(np.uint64), ud[m].view(np.uint64))
        np.testing.assert_array_equal(rows[m, 3].astype(np.uint8), ach[m])
        np.testing.assert_array_equal(rows[m, 4], arr["information"][m].astype(np.float64))
        np.testing.assert_array_equal(rows[m, 5].copy().view(np.uint64), pl[m].view(np.uint64))
        np.testing.assert_array_equal(rows[m, 6].copy().view(np.uint64), ph[m].view(np.uint64))
        # getFilteredFrontiersN: local list, global list, closest global frontier
        np.testing.assert_array_equal(loc, srt[:sel.n_local].astype(np.int32))
        np.testing.assert_array_equal(glo, srt[sel.global_begin:sel.global_begin + sel.n_global].astype(np.int32))
        assert closest == sel.closest_global
        assert len(loc) + len(glo) > 0


def test_reference_assign_costs_few_frontiers_and_errors(oracle, ref, synth, capi):
    a, meta, pts, ov = synth.make_config("C1", reference_rays=True)
    synth.apply_overrides(oracle.params, ov)
    st, fr, cells, seeds, cent = oracle.search_frontiers(a, meta, meta["robot_x"], meta["robot_y"])
    q = oracle.orientation_around_z(0.0)
    rx, ry = meta["robot_x"], meta["robot_y"]
    _, _, ref_yaw = ref.plan_euclidean(rx, ry, q, np.zeros((0, 2)), np.zeros(0, np.uint8), np.zeros(0))
    # the five cases of getFilteredFrontiersN's comment: fewer / more local candidates than n, none global, none local
    d = np.hypot(fr["goal_x"] - rx, fr["goal_y"] - ry)
    near, far = np.argsort(d)[:12], np.argsort(d)[-6:]
    for pick in (near[:3], near[:5], near[:6], near, far, np.concatenate([near[:2], far]), np.concatenate([near, far])):
        sub = fr[np.sort(pick)]
        arr, ach, pl, ph, sto, wc, ua, ud, srt, sel = _oracle_chain(oracle, a, meta, sub, meta["poly"], rx, ry, ref_yaw, None)
        st, rows, loc, glo, closest = ref.assign_costs(a, meta, oracle.params, np.stack([sub["goal_x"], sub["goal_y"]], axis=1), sub["size"],
                                                       meta["poly"], rx, ry, q, None)
        assert st == 0 and sto == 0
        np.testing.assert_array_equal(rows[:, 0].copy().view(np.uint64), wc.view(np.uint64))
        np.testing.assert_array_equal(loc, srt[:sel.n_local].astype(np.int32))
        np.testing.assert_array_equal(glo, srt[sel.global_begin:sel.global_begin + sel.n_global].astype(np.int32))
        assert closest == sel.closest_global
    # no frontiers: assignCosts returns false (FrontierCostsManager.cpp:55-59)
    st, rows, loc, glo, closest = ref.assign_costs(a, meta, oracle.params, np.zeros((0, 2)), np.zeros(0), meta["poly"], rx, ry, q, None)
    assert st == 1


# ---- stage F against the reference's own FisherInformationManager (compiled whole) and its Fisher helpers ---------------------------
# (the fixed-size float algebra underneath is oracle/ref_shims/Eigen/Core's restatement of Eigen 3.4's evaluation order: Eigen
# itself is not in the image.  Everything above it -- the table's loops, keys, de-duplication and file format, getVoxelCoordinate,
# the crowding factor, the order and type of the accumulation in isPoseSafe -- is the reference's code.)
def test_reference_fisher_lookup_table_file_is_byte_identical(oracle, ref, tmp_path):
    p_ref, p_orc = str(tmp_path / "ref.dat"), str(tmp_path / "orc.dat")
    ref.fisher_generate(p_ref, oracle.params)
    assert oracle.fisher_generate(p_orc) == 0
    b_ref, b_orc = open(p_ref, "rb").read(), open(p_orc, "rb").read()
    assert len(b_ref) == len(b_orc) > 16 * 500_000  # 16-byte records: 3 float keys + the value
    assert b_ref == b_orc
    # loadLookupTable: both read the same number of distinct keys (the trailing (0, 0, 0) record overwrites the one generated)
    h = ref.fisher_manager(p_orc)
    assert h is not None
    ref.fisher_destroy(h)
    assert ref.fisher_manager(str(tmp_path / "missing.dat")) is None  # "Cannot load lookup table. Does it exist in the path?"


def test_reference_fisher_helpers_known_answers(oracle, ref):
    rng = np.random.default_rng(4)
    for x, y, z in [(0.3, 0.0, 0.0), (0.0, 0.3, 0.0), (2.1, -0.9, 0.6), (20.7, 14.4, -14.7)] + [tuple(v) for v in rng.uniform(-15, 21, (200, 3))]:
        a, b = ref.information_of_point(x, y, z), oracle.information_of_point_local(x, y, z)
        assert np.float32(a).view(np.uint32) == np.float32(b).view(np.uint32), (x, y, z, a, b)
    assert abs(ref.information_of_point(0.3, 0.0, 0.0) - (2 + 2 / 0.09)) < 1e-3  # trace = 2 + 2 / |p|^2
    for x, y, z in [tuple(v) for v in rng.uniform(-16, 22, (500, 3))] + [(0.15, -0.15, 0.45), (0.149999, 0.150001, -0.45)]:
        ok, v = ref.fisher_voxel(x, y, z)
        w = oracle.voxel_coordinate(x, y, z)
        assert ok and np.array_equal(np.array(v, np.float32).view(np.uint32), np.asarray(w, np.float32).view(np.uint32)), (x, y, z)
    for num in list(range(1, 40)) + [100, 511, 512]:
        assert np.float32(ref.crowd_factor(num, 0.8)).view(np.uint32) == np.float32(oracle.crowd_factor(num)).view(np.uint32)


def test_reference_is_pose_safe_equals_oracle(oracle, ref, synth, capi, tmp_path):
    path = str(tmp_path / "lut.dat")
    assert oracle.fisher_generate(path) == 0
    h = ref.fisher_manager(path)
    assert h is not None
    try:
        rng = np.random.default_rng(12)
        a, meta, pts, ov = synth.make_config("C1")
        box = synth.explored_bbox(a, meta)
        n_safe = 0
        for t in range(24):
            pose = np.zeros(1, dtype=capi.POSE_DTYPE)
            pose["x"], pose["y"], pose["z"] = rng.uniform(box[0], box[2]), rng.uniform(box[1], box[3]), 0.3
            yaw = rng.uniform(-np.pi, np.pi)
            pose["qz"], pose["qw"] = np.sin(yaw / 2), np.cos(yaw / 2)
            st, winfo, wsafe, wvis, wused = oracle.fisher_poses(pts, pose)
            lm = oracle.landmarks_in_view(pts, pose)  # the service's response for this pose (float32 widened on the wire)
            assert st == 0 and len(lm) == wvis[0]
            safe, info = ref.fisher_pose(h, lm.astype(np.float64), oracle.params.fisher_information_threshold)
            # the same float accumulation in the same order: bit for bit
            assert np.float32(info).view(np.uint32) == winfo[:1].view(np.uint32)[0], (t, info, winfo[0])
            assert safe == bool(wsafe[0])
            n_safe += safe
        # a dense clump: hundreds of landmarks per voxel (crowding factors), and the empty response
        clump = (rng.normal(0, 0.05, (3000, 3)) + np.array([3.0, 0.2, 0.1])).astype(np.float32)
        pose = np.zeros(1, dtype=capi.POSE_DTYPE)
        pose["qw"] = 1.0
        st, winfo, wsafe, wvis, wused = oracle.fisher_poses(clump, pose)
        lm = oracle.landmarks_in_view(clump, pose)
        safe, info = ref.fisher_pose(h, lm.astype(np.float64), oracle.params.fisher_information_threshold)
        assert len(lm) > 2000 and np.float32(info).view(np.uint32) == winfo[:1].view(np.uint32)[0] and safe == bool(wsafe[0])
        safe, info = ref.fisher_pose(h, np.zeros((0, 3)), oracle.params.fisher_information_threshold)
        assert not safe and info == 0.0
    finally:
        ref.fisher_destroy(h)


# ---- next rows 8f-2 and 8f-4 against the roadmap-explorer package's own code (oracle/_ref/libfitslam_ref_re.so) ---------------------
@pytest.mark.parametrize("shape,res,ox,oy", [((120, 160), 0.05, -3.0, -2.0), ((37, 300), 0.1, 0.0, 0.0), ((513, 257), 0.03, -7.7, 1.3)])
def test_reference_lethal_marker(oracle, ref, shape, res, ox, oy):
    """nav2_plugins/lethal_marker.cpp compiled from where it lies: the stamped map and the number of cached areas."""
    rng = np.random.default_rng(shape[0])
    H, W = shape
    meta = dict(res=res, ox=ox, oy=oy)
    a = rng.integers(0, 253, size=shape, dtype=np.uint8)
    areas = []
    for _ in range(40):
        wx = rng.uniform(ox - 0.5, ox + W * res + 0.5)  # some centres off the map: skipped, not counted
        wy = rng.uniform(oy - 0.5, oy + H * res + 0.5)
        kind = int(rng.integers(0, 2))
        extent = float(rng.choice([0.0, res * 0.4, 0.33, 1.0, 2.75, 40.0]))  # 40 m: every ray end clamped to the border
        areas.append((wx, wy, extent, float(rng.uniform(-7, 7)), kind))
    areas += [(ox, oy, 1.0, 0.3, 0), (ox + W * res - 1e-9, oy + H * res - 1e-9, 1.0, -2.0, 1), (ox + W * res, oy, 1.0, 0.0, 0)]
    for sub in (areas, areas[:1], areas[5:9], []):
        want, n_want = ref.mark_lethal(a, meta, sub)
        got, n_got = oracle.mark_lethal(a, meta, sub)
        assert n_got == n_want
        assert np.array_equal(got, want)
    assert n_want == 0 and np.array_equal(want, a)
    want, n_want = ref.mark_lethal(a, meta, areas)
    assert 0 < n_want < len(areas) and (want != a).sum() > 100


def test_reference_threshold_node(oracle, ref, capi):
    """The threshold half of the ingest (threshold_occupancy.cpp:24-49, the reference's own text): every int8 value at the node's
    default threshold and at others.  The oracle composes it with the static layer's translation; with the trinary static layer
    {-1, 0, 100} -> {255, 0, 254} is one to one, so equal costs mean equal thresholded grids.  (The static layer itself is a Nav2
    binary package absent from the reference tree: that half stays a restatement.)"""
    g = np.concatenate([np.arange(-128, 128, dtype=np.int16).astype(np.int8), np.random.default_rng(3).integers(-128, 128, 5000).astype(np.int8)])
    for thr in (85, 0, 1, 50, 100, 101, 127):
        p = capi.default_occupancy_params()
        p.lethal_obstacle_threshold = thr
        q = capi.default_occupancy_params()
        q.lethal_obstacle_threshold = -1  # no threshold node: the static layer alone
        t = ref.threshold_occupancy(g, thr)
        assert set(np.unique(t).tolist()) <= {-1, 0, 100}
        assert np.array_equal(oracle.occupancy_to_cost(g, p), oracle.occupancy_to_cost(t, q))
    t = ref.threshold_occupancy(np.array([-1, 0, 84, 85, 100, -2, -128, 127], np.int8), 85)
    assert t.tolist() == [-1, 0, 0, 100, 100, 0, 0, 100]


@pytest.mark.skipif(not ref_binding.available_re(), reason="oracle/_ref is built where /root/reference exists")
def test_reference_roadmap_is_connectable(oracle, synth):
    """FrontierRoadMap::isConnectable (RE/src/planners/FrontierRoadmap.cpp:671-692 over RE/src/Helpers.cpp getTracedCells),
    compiled from the reference, against the oracle: 3 000 node pairs x 3 radii."""
    re = ref_binding.ReferenceRE()
    a, meta = synth.make_costmap(300, 300, seed=13, n_disc=6, r_lo=20, r_hi=50, n_holes=10, inflation=2)
    rng = np.random.default_rng(5)
    H, W = a.shape
    n = 3000
    x1, y1 = meta["ox"] + rng.uniform(1, W - 1, n) * meta["res"], meta["oy"] + rng.uniform(1, H - 1, n) * meta["res"]
    d, th = rng.uniform(0.0, 5.0, n), rng.uniform(-np.pi, np.pi, n)
    x2 = np.clip(x1 + d * np.cos(th), meta["ox"] + meta["res"], meta["ox"] + (W - 1) * meta["res"])
    y2 = np.clip(y1 + d * np.sin(th), meta["oy"] + meta["res"], meta["oy"] + (H - 1) * meta["res"])
    seg = np.stack([x1, y1, x2, y2], axis=1)
    seg[:5, 2:] = seg[:5, :2]  # a node paired with itself
    for radius in (2.6, 1.0, 4.0):
        want = re.connectable(a, meta, seg, radius)
        got, _ = oracle.roadmap_connectable(a, meta, seg, radius)
        np.testing.assert_array_equal(got, want)
        assert 0 < want.sum() < n


@pytest.mark.skipif(not ref_binding.available_re(), reason="oracle/_ref is built where /root/reference exists")
def test_reference_sensor_simulator_rays(oracle, capi):
    """SensorSimulator::markRay (RE/src/SensorSimulator.cpp:212-246, compiled from the reference) under the fan loop of
    timerCallback, against the oracle: explored grids equal cell for cell, same number of rays."""
    re = ref_binding.ReferenceRE()
    rng = np.random.default_rng(17)
    p = capi.default_sensor_params()
    for H, W, res in ((200, 260, 0.05), (97, 131, 0.1), (64, 64, 1.0 / 16)):
        latest = rng.choice(np.array([0, 100, -1, 81, 80], np.int8), size=(H, W), p=[0.8, 0.05, 0.1, 0.02, 0.03])
        ox, oy = -W * res / 2 + 0.0071, -H * res / 2 - 0.0033
        explored_o = np.full((H, W), -1, np.int8)
        explored_r = explored_o.copy()
        for t in range(12):
            px, py = rng.uniform(ox - 0.5, ox + W * res + 0.5), rng.uniform(oy - 0.5, oy + H * res + 0.5)  # sometimes off the map
            yaw = rng.uniform(-7, 7)
            explored_o, n_o = oracle.sensor_mark_rays(latest, explored_o, res, ox, oy, px, py, yaw, p)
            explored_r, n_r = re.mark_rays(latest, explored_r, res, ox, oy, px, py, yaw, p)
            assert n_o == n_r == 121
            np.testing.assert_array_equal(explored_o, explored_r)
        assert (explored_o != -1).sum() > 100


# ---------------------------------------------------------------------------------------------------------------------------
# Rows F1, F2 and the keyframe mode of 8f-1: the ORB-SLAM3 wrapper's own mapPointsVisibleFromPose (both overloads), landmark
# service handler and type conversions (oracle/_ref/libfitslam_ref_orbw.so), against stand-in ORB_SLAM3 / Sophus / Eigen
# declarations.  Pinned: the control flow and the sequence of operations.  Not pinned: Sophus's quaternion storage and Eigen's
# Affine rotation() (the stand-ins keep (R, t), as oracle.cpp does).
# ---------------------------------------------------------------------------------------------------------------------------
orbw = pytest.mark.skipif(not ref_binding.available_orbw(), reason="oracle/_ref is built where /root/reference exists")


def _pose_q(capi, x, y, z, q):
    p = np.zeros(1, dtype=capi.POSE_DTYPE)
    p["x"], p["y"], p["z"] = x, y, z
    p["qx"], p["qy"], p["qz"], p["qw"] = q
    return p


def _yaw_q(yaw):
    return (0.0, 0.0, math.sin(yaw / 2), math.cos(yaw / 2))


@orbw
@pytest.mark.parametrize("exhaustive", [True, False])
def test_reference_landmark_query_keyframes(oracle, synth, capi, exhaustive):
    """A synthetic atlas of 200 keyframes over 60 k points: the landmarks of the response coordinate for coordinate and in order,
    the number of keyframes checked, with the handler's own 1000-landmark cutoff, smaller ones and none."""
    orb = ref_binding.ReferenceORBW()
    rng = np.random.default_rng(21)
    pts = rng.uniform(-25, 25, size=(60_000, 3)).astype(np.float32)
    pts[:, 2] = rng.uniform(0, 3, len(pts))
    kfs, idx = synth.make_keyframes(pts, 200, seed=9)
    perm = np.array([[0, -1, 0], [0, 0, -1], [1, 0, 0]], dtype=np.float64)
    seen_cut = seen_skip = 0
    for t in range(30):
        k = int(rng.integers(len(kfs)))
        t_ros = perm.T @ kfs[k]["twc"].astype(np.float64)
        if t % 3 == 2:  # a tilted camera (roll and pitch too), a quaternion not quite of unit length
            q = rng.normal(size=4)
            q = q / np.linalg.norm(q) * (1 + 1e-4 * rng.normal())
        else:
            q = _yaw_q(rng.uniform(-np.pi, np.pi))
        pose = _pose_q(capi, t_ros[0] + rng.normal(0, 1.5), t_ros[1] + rng.normal(0, 1.5), 0.3, q)
        for max_landmarks in (1000, 120, 0, 1 << 30):
            for max_distance in (5.0, 2.5):
                want, chk_w = orb.landmarks_in_view(pts, kfs, idx, pose, oracle.params, max_landmarks, max_distance, exhaustive)
                got, chk = oracle.landmarks_in_view_keyframes(pts, kfs, idx, pose, max_landmarks, max_distance, exhaustive)
                assert chk == chk_w and len(got) == len(want), (t, max_landmarks, max_distance, chk, chk_w, len(got), len(want))
                # the response carries float64 positions holding the handler's floats
                assert np.array_equal(got.astype(np.float64), want, equal_nan=True)
                seen_cut += int(max_landmarks == 1000 and len(want) > 1000)
                seen_skip += int(not exhaustive and chk_w < len(kfs))
    assert seen_cut > 0 or not exhaustive
    assert exhaustive or seen_skip > 0


@orbw
def test_reference_landmark_query_one_keyframe_is_the_exhaustive_query(oracle, capi):
    """Rows F1 / F2 as the other entry points answer them: one keyframe holding every map point, exhaustive search, no cutoff in
    reach; points behind the camera, on the image border, at the camera centre, non-finite."""
    orb = ref_binding.ReferenceORBW()
    rng = np.random.default_rng(4)
    pts = rng.uniform(-12, 12, size=(20_000, 3)).astype(np.float32)
    for t in range(12):
        q = rng.normal(size=4)
        q /= np.linalg.norm(q)
        if t % 2:
            q = np.array(_yaw_q(rng.uniform(-np.pi, np.pi)))
        pose = _pose_q(capi, rng.uniform(-5, 5), rng.uniform(-5, 5), rng.uniform(0, 2), q)
        pts[0] = (pose["x"][0], pose["y"][0], pose["z"][0])  # Zc = 0: 0 / 0 in the projection
        pts[1] = (np.nan, 0, 0)
        pts[2] = (np.inf, 1, 1)
        one = np.zeros(1, dtype=capi.KEYFRAME_DTYPE)
        one["Rwc"] = np.eye(3, dtype=np.float32).ravel()
        one["count"] = len(pts)
        idx = np.arange(len(pts), dtype=np.uint32)
        want, chk = orb.landmarks_in_view(pts, one, idx, pose, oracle.params, 1 << 30, 5.0, True)
        got = oracle.landmarks_in_view(pts, pose)
        assert chk == 1 and len(got) == len(want) > 0
        assert np.array_equal(got.astype(np.float64), want, equal_nan=True)


@orbw
def test_reference_rotation_to_euler_ros(oracle):
    """rotationORBToEulerROS: the relative yaw the prefilter reads, for rotations about the ORB y axis (a ROS yaw)."""
    orb = ref_binding.ReferenceORBW()
    perm = np.array([[0, -1, 0], [0, 0, -1], [1, 0, 0]], dtype=np.float64)
    for yaw in (0.0, 0.3, 1.5, 2.0, 3.0, -0.3, -1.5, -2.0, -3.0):
        c, s = math.cos(yaw), math.sin(yaw)
        R = perm @ np.array([[c, -s, 0], [s, c, 0], [0, 0, 1]]) @ perm.T
        e = orb.euler_ros(R.astype(np.float32))
        assert abs(e[2] - (yaw % (2 * math.pi))) < 1e-5, (yaw, e)
