"""Known-answer micro-cases for the CPU oracle: hand-derivable from the cited reference lines.
The reference ships no golden vectors for this path ("parity unpinned"); these pin the restatement instead,
together with the one reference test that applies (ORBW/tests/typeConversionsTests.cpp:8-33)."""
import math

import numpy as np
import pytest


def _meta(W, H, res=0.05, ox=0.0, oy=0.0):
    return dict(res=res, ox=ox, oy=oy, W=W, H=H)


def test_frontier_predicate_truth_table(oracle):
    # DEP/src/FrontierSearch.cpp:218-249: unknown cell, >= 1 free 4-neighbour, no lethal (>= 160, != 255) 4-neighbour
    a = np.full((5, 5), 255, np.uint8)
    a[2, 2] = 0
    m = oracle.frontier_mask(a)
    want = np.zeros((5, 5), np.uint8)
    want[1, 2] = want[3, 2] = want[2, 1] = want[2, 3] = 1  # 4-neighbours of the free cell; diagonals are not
    np.testing.assert_array_equal(m, want)
    a[0, 2] = 254          # a lethal neighbour vetoes (1,2)
    a[2, 0] = 159          # 159 < 160 is free, not lethal: (2,1) stays a frontier and (1,0),(3,0) become one
    a[4, 2] = 160          # 160 is lethal: vetoes (3,2)
    m = oracle.frontier_mask(a)
    assert m[1, 2] == 0 and m[3, 2] == 0 and m[2, 1] == 1 and m[2, 3] == 1 and m[1, 0] == 1 and m[3, 0] == 1


def test_straight_strip_chunks_21_21_18(oracle):
    # DEP/src/FrontierSearch.cpp:150-199: a 60-cell strip is cut in discovery order into 21 + 21 + 18
    W, H = 80, 9
    a = np.full((H, W), 254, np.uint8)
    a[3, 10:70] = 0        # free corridor
    a[4, 10:70] = 255      # the frontier strip below it (lethal everywhere else)
    a[5, 10:70] = 255      # second unknown row: has no free neighbour -> not frontier
    a[4, 9] = a[4, 70] = 255   # unknown, not lethal: the strip's end cells keep no lethal 4-neighbour
    meta = _meta(W, H)
    rx, ry = (10 + 0.5) * 0.05, (3 + 0.5) * 0.05
    st, fr, cells, seeds, cent = oracle.search_frontiers(a, meta, rx, ry)
    assert st == 0
    assert sorted(fr["size"].tolist(), reverse=True) == [21, 21, 18]
    assert len(cells) == 60 and len(seeds) == 1 and seeds[0] == 4 * W + 10  # first touch: robot cell's +W neighbour
    assert set(cells.tolist()) == {4 * W + x for x in range(10, 70)}
    assert np.all(fr["goal_my"] == 4)
    assert fr["chunk"].tolist() == [0, 1, 2] and fr["first_cell"].tolist() == [0, 21, 42]


def test_single_ray_counts(oracle, capi):
    # DEP/src/Helpers.cpp:7-96 + Helpers.hpp:50-77: 41 cells along +x on an all-unknown map; obstacles cut the count
    W, H = 200, 21
    meta = _meta(W, H)
    fr = np.zeros(1, capi.FRONTIER_DTYPE)
    fr["goal_x"], fr["goal_y"], fr["size"] = (100 + 0.5) * 0.05, (10 + 0.5) * 0.05, 20
    poly = [0, 0, W * 0.05, H * 0.05]
    a = np.full((H, W), 255, np.uint8)
    out, rays, wins = oracle.score_arrival(a, meta, fr, poly)
    assert rays.shape[1] == 63 and wins.shape[1] == 54   # theta loop with delta 0.10: 63 rays, k = 10
    assert oracle.ray_shape(0.05) == (63, 10, 54, 40)
    assert rays[0, 0] == 41
    for cost, want in ((254, 10), (240, 10), (239, 41 - 1), (0, 41 - 1)):
        b = a.copy()
        b[10, 110] = cost   # step 10 along the ray
        out, rays, _ = oracle.score_arrival(b, meta, fr, poly)
        assert rays[0, 0] == want, (cost, rays[0, 0])
    # limits: every ray counts all its cells; max window = 10 rays; 1.2x and 0.7x factors (CostCalculator.cpp:186-188)
    big = np.full((400, 400), 255, np.uint8)
    mx, mn, mv = oracle.arrival_limits(big, _meta(400, 400, ox=-10.0, oy=-10.0))
    assert mx == mv * 1.2 and mn == 0.70 * mx and 10 * 29 <= mv <= 10 * 41


def test_ray_endpoint_clamped_to_map_edge_stays_inside(oracle, capi):
    # getSizeInMetersX = (size_x - 1 + 0.5) * res: an end point clamped to origin + size still maps to the last cell
    W, H = 60, 60
    a = np.full((H, W), 255, np.uint8)
    meta = _meta(W, H)
    fr = np.zeros(1, capi.FRONTIER_DTYPE)
    fr["goal_x"], fr["goal_y"], fr["size"] = (55 + 0.5) * 0.05, (30 + 0.5) * 0.05, 20
    out, rays, _ = oracle.score_arrival(a, meta, fr, [-100, -100, 100, 100])
    assert out["ray_failed"][0] == 0 and rays[0, 0] == 5  # cells 55..59
    fr["goal_x"] = -0.01
    out, _, _ = oracle.score_arrival(a, meta, fr, [-100, -100, 100, 100])
    assert out["ray_failed"][0] == 1 and out["information"][0] == 0 and out["theta_s_star"][0] == 0.0


def test_fim_trace_closed_form_and_crowding(oracle):
    # trace(J^T J) = 2 + 2/|p|^2; crowd factor exp(1 - k^0.8)
    for p in ((1, 0, 0), (3, 0, 0), (0.3, 0, 0), (2.1, -3.3, 0.9), (20.7, 14.7, -14.7)):
        want = 2 + 2 / sum(x * x for x in p)
        assert abs(oracle.information_of_point_local(*p) - want) < 2e-5 * want
    assert math.isnan(oracle.information_of_point_local(0, 0, 0))
    assert oracle.crowd_factor(1) == 1.0
    assert abs(oracle.crowd_factor(2) - math.exp(1 - 2 ** 0.8)) < 1e-6
    assert abs(oracle.crowd_factor(3) - math.exp(1 - 3 ** 0.8)) < 1e-6


def test_lookup_table_shape_and_voxel_rounding(oracle):
    assert oracle.fisher_generate() == 0
    rec = oracle.fisher_records()
    assert oracle.fisher_num_entries() == len(rec)          # (0,0,0) was skipped as NaN, appended once with the max
    assert tuple(rec[-1][:3]) == (0.0, 0.0, 0.0) and rec[-1][3] == rec[:, 3].max()
    assert abs(rec[-1][3] - (2 + 2 / 0.09)) < 1e-3
    idx = np.rint(rec[:, :3].astype(np.float64) / np.float64(np.float32(0.3))).astype(int)
    assert idx[:, 0].min() == 0 and idx[:, 0].max() in (70, 71) and idx[:, 1].min() in (-50, -51) and idx[:, 1].max() in (49, 50, 51)  # float accumulation decides the last slab
    # getVoxelCoordinate: round half away from zero to the 0.3 m lattice
    vx = oracle.voxel_coordinate(0.449, -0.451, 0.15)
    assert abs(vx[0] - 0.3) < 1e-6 and abs(vx[1] + 0.6) < 1e-6 and abs(vx[2] - 0.3) < 1e-6
    assert math.isnan(oracle.voxel_information(25.0, 0, 0)) and not math.isnan(oracle.voxel_information(20.9, 0, 0))


def test_frustum_rule_and_crowded_voxel(oracle, capi):
    # a pose-frame point (x,y,z) is kept iff x >= 0, 0 <= 320 - fx*y/x <= 640, 0 <= 240 - fx*z/x <= 480
    oracle.fisher_generate()
    pose = np.zeros(1, capi.POSE_DTYPE)
    pose["qw"] = 1.0
    fx = 432.496042035043
    pts = np.array([[2, 0, 0], [2, 2 * 319 / fx, 0], [2, 2 * 321 / fx, 0], [-1, 0, 0], [2, 0, 2 * 239 / fx],
                    [2, 0, -2 * 241 / fx], [30, 0, 0]], np.float32)
    vis = oracle.landmarks_in_view(pts, pose)
    assert len(vis) == 4  # the centre point, the two just inside the borders, and the far one (no far plane)
    st, info, safe, nvis, nused = oracle.fisher_poses(pts, pose)
    assert nvis[0] == 4 and nused[0] == 3  # (30,0,0) is outside the table domain
    # k landmarks in one voxel: info * (1 + e^(1-2^0.8) + e^(1-3^0.8))
    same = np.array([[3.0, 0.01, 0.0], [3.01, -0.02, 0.01], [2.98, 0.0, -0.01]], np.float32)
    st, info, *_ = oracle.fisher_poses(same, pose)
    v = oracle.voxel_information(3.0, 0, 0)
    want = v * (1 + math.exp(1 - 2 ** 0.8) + math.exp(1 - 3 ** 0.8))
    assert abs(info[0] - want) < 1e-4 * want and abs(v - (2 + 2 / 9.0)) < 1e-4


def test_se3_ros_orb_roundtrip(oracle):
    # ORBW/tests/typeConversionsTests.cpp:8-33: se3ORBToROS(se3ROSToORB(A)).isApprox(A, 1e-5)
    rng = np.random.default_rng(0)
    for _ in range(20):
        q = rng.normal(size=4)
        q /= np.linalg.norm(q)
        w, x, y, z = q
        R = np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)],
                      [2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)],
                      [2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)]], np.float32)
        t = rng.uniform(-10, 10, 3).astype(np.float32)
        Ro, to = oracle.se3_roundtrip(R, t)
        A, B = np.hstack([R, t[:, None]]), np.hstack([Ro, to[:, None]])
        assert np.linalg.norm(A - B) <= 1e-5 * min(np.linalg.norm(A), np.linalg.norm(B))


def test_u1_utility_formula(oracle, capi):
    # SURVEY 8c-9 / DEP/src/FrontierCostsManager.cpp:152-198
    arr = np.zeros(3, capi.ARRIVAL_DTYPE)
    arr["information"] = [int(492 * 0.8), 100, 50]
    ach = np.array([1, 1, 0], np.uint8)
    pl, ph = np.array([10.0, 4.0, 1.0]), np.array([0.7, 0.1, 0.0])
    st, wc, ua, ud, err = oracle.normalise(arr, ach, pl, ph, None, 492.0)
    assert st == 0
    a0 = int(492 * 0.8) / 492.0
    p0 = 1 - (10.0 / 0.5 + 0.7 / 0.5) / (10.0 / 0.5 + math.pi / 0.5)
    assert ua[0] == a0 and ud[0] == p0 and wc[0] == 1 / (1.0 * (0.25 * a0 + 0.75 * p0))
    assert wc[2] == np.finfo(np.float64).max and ua[2] == -69.8 and ud[2] == -1.8
    arr["information"][0] = 600
    st, *_ = oracle.normalise(arr, ach, pl, ph, None, 492.0)
    assert st == capi.E_ARRIVAL_UTILITY


def test_selection_known_answers(oracle):
    # SURVEY 8c-10 / DEP/src/FullPathOptimizer.cpp:186-226
    el = np.ones(7, np.uint8)
    srt, sel = oracle.select(el, np.arange(1.0, 8.0))
    assert (sel.n_local, sel.n_global, sel.closest_global) == (5, 0, 6) and srt.tolist() == list(range(7))
    srt, sel = oracle.select(np.ones(5, np.uint8), np.array([5.0, 4, 3, 2, 1]))
    assert (sel.n_local, sel.n_global, sel.closest_global) == (4, 0, 0) and srt.tolist() == [4, 3, 2, 1, 0]
    srt, sel = oracle.select(np.ones(4, np.uint8), np.array([20.0, 30, 1, 2]))
    assert (sel.n_local, sel.n_global, sel.global_begin, sel.closest_global) == (2, 2, 2, 0)
    srt, sel = oracle.select(np.ones(1, np.uint8), np.array([3.0]))
    assert (sel.n_local, sel.n_global, sel.global_begin, sel.closest_global) == (0, 1, 0, 0)
    srt, sel = oracle.select(np.zeros(3, np.uint8), np.array([3.0, 1, 2]))
    assert sel.n_sorted == 0 and sel.closest_global == -1


def test_uid_is_libstdcpp_hash_xor(oracle):
    # Frontier.hpp:151-161; std::hash<double>(0.0) == 0
    assert oracle.generate_uid(0.0, 0.0) == 0
    assert oracle.generate_uid(1.5, 0.0) != 0 and oracle.generate_uid(1.5, 2.5) != oracle.generate_uid(2.5, 1.5)


def test_robot_out_of_bounds_and_nearest_free_cell(oracle, capi):
    a = np.full((20, 20), 255, np.uint8)
    a[10, 10] = 0
    st, fr, *_ = oracle.search_frontiers(a, _meta(20, 20), -1.0, 0.2)
    assert st == capi.E_ROBOT_OUT_OF_BOUNDS and len(fr) == 0
    ok, cell = oracle.nearest_free_cell(a, 7 * 20 + 10, 160)   # 3 cells below the free cell: 8-connected BFS finds it
    assert ok and cell == 10 * 20 + 10
    ok, _ = oracle.nearest_free_cell(np.full((5, 5), 255, np.uint8), 12, 160)
    assert not ok
