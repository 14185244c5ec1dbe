"""The two writers next to the path (SURVEY.md 8f-3 stamping half, 8f-4):
LethalMarker ray fans stamped with 253 (DEP/src/nav2_plugins/lethal_marker.cpp, FS2 keepout_layer.cpp) and
SensorSimulator::markRay (RE/src/SensorSimulator.cpp:193-246).  Oracle against hand-derived cells on CPU; the CUDA
kernels against the oracle, bit for bit, on the GPU."""
import math

import numpy as np
import pytest


# ----------------------------------------------------------------------------------------------- oracle, known answers
def _meta(res=1.0, ox=0.0, oy=0.0):
    return {"res": res, "ox": ox, "oy": oy}


def test_oracle_lethal_circle_known_cells(oracle):
    a = np.zeros((21, 21), np.uint8)
    out, n = oracle.mark_lethal(a, _meta(), [(10.5, 10.5, 3.0, 0.0, 0)])
    assert n == 1
    assert set(np.unique(out)) == {0, 253}
    # ray 90 (angle pi/2 - pi/2 = 0) ends at (13, 10); ray 0 at (10, 7); ray 180 at (10, 13); ray 270 (cos(pi) = -1) at (7, 10)
    # the end points truncate towards zero, not towards the centre: the 225 degree ray ends at (int)(10 - 2.12) = 7 on both
    # axes, the 45 degree ray at (int)(10 + 2.12) = 12
    for x, y in [(10, 10), (11, 10), (12, 10), (13, 10), (10, 7), (10, 13), (7, 10), (7, 7), (12, 12), (7, 12)]:
        assert out[y, x] == 253, (x, y)
    for x, y in [(14, 10), (10, 6), (10, 14), (6, 10), (13, 13), (13, 11), (11, 13), (6, 6)]:
        assert out[y, x] == 0, (x, y)
    ys, xs = np.nonzero(out)
    assert xs.min() == 7 and xs.max() == 13 and ys.min() == 7 and ys.max() == 13
    # a centre off the map is skipped (lethal_marker.cpp:222-223); the extent truncates to whole cells (3.9 -> 3)
    out2, n2 = oracle.mark_lethal(a, _meta(), [(-0.5, 3.0, 3.0, 0.0, 0), (25.0, 3.0, 3.0, 0.0, 1)])
    assert n2 == 0 and not out2.any()
    out3, _ = oracle.mark_lethal(a, _meta(), [(10.5, 10.5, 3.9, 0.0, 0)])
    np.testing.assert_array_equal(out3, out)


def test_oracle_lethal_circle_clamps_to_the_map(oracle):
    a = np.zeros((12, 12), np.uint8)
    out, n = oracle.mark_lethal(a, _meta(), [(1.5, 1.5, 5.0, 0.0, 0)])
    assert n == 1
    # rays towards -x / -y end on column / row 0 (cellEnforceBoundaries), the +x ray reaches (6, 1)
    assert out[1, 0] == 253 and out[0, 1] == 253 and out[0, 0] == 253 and out[1, 6] == 253 and out[1, 7] == 0


def test_oracle_lethal_triangle_known_cells(oracle):
    a = np.zeros((41, 41), np.uint8)
    out, n = oracle.mark_lethal(a, _meta(), [(10.5, 10.5, 8.0, 0.0, 1)])
    assert n == 1
    # base centre (18, 10), half base 8 tan(22.5 deg) = 3.31: the 20 rays end between (18, 13) and (18, 6)
    for x, y in [(10, 10), (18, 13), (18, 6), (14, 10), (18, 10)]:
        assert out[y, x] == 253, (x, y)
    for x, y in [(19, 10), (18, 14), (18, 5), (9, 10), (10, 11), (10, 9)]:
        assert out[y, x] == 0, (x, y)
    assert out[:, 18].sum() // 253 == 8  # one base cell per row 6..13


def test_oracle_sensor_single_ray(capi, oracle):
    p = capi.default_sensor_params()
    assert (p.sensor_min_angle, p.sensor_max_angle, p.angular_resolution, p.sensor_max_range) == (-math.pi / 4, math.pi / 4, 0.013, 3.0)
    latest = np.zeros((5, 20), np.int8)
    latest[2, 7] = 100
    latest[2, 5] = 80  # not > 80: the ray goes on
    explored = np.full((5, 20), -1, np.int8)
    p.sensor_min_angle = p.sensor_max_angle = 0.0
    p.sensor_max_range = 10.0
    out, n = oracle.sensor_mark_rays(latest, explored, 1.0, 0.0, 0.0, 2.5, 2.5, 0.0, p)
    assert n == 1
    np.testing.assert_array_equal(out[2], [-1, -1, 0, 0, 0, 80, 0, 100] + [-1] * 12)
    assert (np.delete(out, 2, axis=0) == -1).all()
    # no obstacle: the ray stops at the map border (range 30 > map)
    latest[2, 7] = 0
    p.sensor_max_range = 30.0
    out, _ = oracle.sensor_mark_rays(latest, explored, 1.0, 0.0, 0.0, 2.5, 2.5, 0.0, p)
    np.testing.assert_array_equal(out[2], [-1, -1] + list(latest[2, 2:]))
    # the robot's own cell off the map: r = 0 already breaks
    out, _ = oracle.sensor_mark_rays(latest, explored, 1.0, 0.0, 0.0, -3.0, 2.5, 0.0, p)
    assert (out == -1).all()


def test_oracle_sensor_ray_count_is_the_accumulated_loop(capi, oracle):
    p = capi.default_sensor_params()
    n_loop, a = 0, p.sensor_min_angle
    while a <= p.sensor_max_angle:
        n_loop += 1
        a += p.angular_resolution
    latest = np.zeros((64, 64), np.int8)
    out, n = oracle.sensor_mark_rays(latest, np.full_like(latest, -1), 0.05, -1.6, -1.6, 0.0, 0.0, 0.3, p)
    assert n == n_loop == 121
    # a fan of +-45 degrees around yaw 0.3 with range 3 m on a 3.2 m map: cells behind the robot stay unknown
    assert out[32, 32] == 0 and out[32, 0] == -1 and (out == 0).sum() > 500


# ----------------------------------------------------------------------------------------------- CUDA against the oracle
def _costmap(rng, H, W):
    a = rng.choice(np.array([0, 254, 255, 100], np.uint8), size=(H, W), p=[0.6, 0.1, 0.25, 0.05])
    return np.ascontiguousarray(a)


def _areas(rng, meta, H, W, n):
    out = []
    for _ in range(n):
        wx = meta["ox"] + rng.uniform(-0.1, 1.1) * W * meta["res"]
        wy = meta["oy"] + rng.uniform(-0.1, 1.1) * H * meta["res"]
        kind = int(rng.integers(0, 2))
        out.append((wx, wy, float(rng.uniform(0.0, 4.0)), float(rng.uniform(-7.0, 7.0)), kind))
    return out


@pytest.mark.gpu
@pytest.mark.parametrize("shape,res", [((200, 317), 0.05), ((64, 64), 0.1), ((97, 33), 0.025)])
def test_gpu_mark_lethal_matches_oracle(capi, oracle, shape, res):
    H, W = shape
    rng = np.random.default_rng(H * 7 + W)
    a = _costmap(rng, H, W)
    meta = {"res": res, "ox": -W * res / 2 + 0.013, "oy": -H * res / 3}
    ctx = capi.Context()
    try:
        ctx.set_costmap(a, meta["res"], meta["ox"], meta["oy"])
        want = a
        for batch in (1, 7, 40):
            areas = _areas(rng, meta, H, W, batch)
            want, n_want = oracle.mark_lethal(want, meta, areas)
            assert ctx.mark_lethal(areas) == n_want
            np.testing.assert_array_equal(ctx.costmap_copy(), want)
        assert (want == 253).sum() > 100
        # the reference callers: addNewMarkedAreaFOV(x, y, yaw, 3.5) (ExplorationBT.cpp:622), addNewMarkedArea(x, y, 1.7) (:621)
        areas = [(meta["ox"] + W * res / 2, meta["oy"] + H * res / 2, 3.5, 0.7, 1), (meta["ox"] + W * res / 4, meta["oy"] + H * res / 4, 1.7, 0.0, 0)]
        want, n_want = oracle.mark_lethal(want, meta, areas)
        assert ctx.mark_lethal(areas) == n_want == 2
        np.testing.assert_array_equal(ctx.costmap_copy(), want)
        assert ctx.mark_lethal([]) == 0
        with pytest.raises(Exception):
            ctx.mark_lethal([(0.0, 0.0, -1.0, 0.0, 0)])
    finally:
        ctx.close()


@pytest.mark.gpu
def test_gpu_search_sees_the_stamps(capi, oracle, synth):
    """A stamped area is lethal to the search (253 >= the 160 / 254 tests of the reference's cell classes apply as on any costmap)."""
    a, meta, pts, ov = synth.make_config("C1")
    areas = [(meta["robot_x"] + 2.0, meta["robot_y"], 3.5, 0.0, 1), (meta["robot_x"] - 3.0, meta["robot_y"] + 1.0, 1.7, 0.0, 0)]
    want_map, _ = oracle.mark_lethal(a, meta, areas)
    c1, c2 = capi.Context(), capi.Context()
    try:
        c1.set_costmap(a, meta["res"], meta["ox"], meta["oy"])
        c1.mark_lethal(areas)
        c2.set_costmap(want_map, meta["res"], meta["ox"], meta["oy"])
        s1, f1, cells1 = c1.search_frontiers(meta["robot_x"], meta["robot_y"])
        s2, f2, cells2 = c2.search_frontiers(meta["robot_x"], meta["robot_y"])
        assert s1 == s2 and f1.tobytes() == f2.tobytes() and cells1.tobytes() == cells2.tobytes()
        assert len(f1) > 0 and (want_map == 253).sum() > 500
    finally:
        c1.close(); c2.close()


def _occupancy(rng, H, W):
    g = np.zeros((H, W), np.int8)
    g[rng.random((H, W)) < 0.04] = 100
    g[rng.random((H, W)) < 0.02] = rng.integers(70, 95, 1).astype(np.int8)[0]
    g[rng.random((H, W)) < 0.05] = -1
    return g


@pytest.mark.gpu
@pytest.mark.parametrize("shape,res", [((240, 320), 0.05), ((50, 61), 0.1), ((400, 400), 0.02)])
def test_gpu_sensor_rays_match_oracle(capi, oracle, shape, res):
    import torch
    H, W = shape
    rng = np.random.default_rng(H + 31 * W)
    latest = _occupancy(rng, H, W)
    ox, oy = -W * res / 2 + 0.0071, -H * res / 2 - 0.0033
    p = capi.default_sensor_params()
    ctx = capi.Context()
    try:
        want = np.full((H, W), -1, np.int8)
        got = want.copy()
        t_latest = torch.from_numpy(latest).cuda()
        t_got = torch.from_numpy(want.copy()).cuda()
        for step in range(24):
            # poses on and around the map, also off it; every few steps a different fan / range
            px = ox + rng.uniform(-0.05, 1.05) * W * res
            py = oy + rng.uniform(-0.05, 1.05) * H * res
            yaw = rng.uniform(-math.pi, math.pi)
            if step % 6 == 5:
                p.sensor_min_angle, p.sensor_max_angle = sorted(rng.uniform(-math.pi, math.pi, 2))
                p.angular_resolution = float(rng.uniform(0.004, 0.05))
                p.sensor_max_range = float(rng.uniform(0.0, 8.0))
            want, n_want = oracle.sensor_mark_rays(latest, want, res, ox, oy, px, py, yaw, p)
            assert ctx.sensor_mark_rays(latest, got, res, ox, oy, px, py, yaw, p) == n_want  # host buffers
            np.testing.assert_array_equal(got, want)
            assert ctx.sensor_mark_rays(t_latest.data_ptr(), t_got.data_ptr(), res, ox, oy, px, py, yaw, p, shape=(H, W)) == n_want
            np.testing.assert_array_equal(t_got.cpu().numpy(), want)  # device buffers
        assert (want != -1).sum() > 200
        # cells only ever take the latest grid's value
        seen = want != -1
        np.testing.assert_array_equal(want[seen], latest[seen])
    finally:
        ctx.close()


@pytest.mark.gpu
def test_gpu_sensor_rays_full_size_properties(capi):
    """4096^2 (the C2 grid): idempotent, only copies latest, nothing beyond the range, stops behind a wall."""
    import torch
    H = W = 4096
    res, ox, oy = 0.05, -W * 0.025, -H * 0.025
    latest = torch.zeros((H, W), dtype=torch.int8, device="cuda")
    latest[:, 2100] = 100  # a wall 52 cells = 2.6 m east of the robot
    explored = torch.full((H, W), -1, dtype=torch.int8, device="cuda")
    p = capi.default_sensor_params()
    p.sensor_min_angle, p.sensor_max_angle, p.sensor_max_range = -math.pi, math.pi, 6.0
    ctx = capi.Context()
    try:
        n = ctx.sensor_mark_rays(latest.data_ptr(), explored.data_ptr(), res, ox, oy, 0.0, 0.0, 0.0, p, shape=(H, W))
        assert n == 484
        first = explored.clone()
        ctx.sensor_mark_rays(latest.data_ptr(), explored.data_ptr(), res, ox, oy, 0.0, 0.0, 0.0, p, shape=(H, W))
        assert torch.equal(first, explored)
        seen = explored != -1
        assert torch.equal(explored[seen], latest[seen])
        ys, xs = torch.nonzero(seen, as_tuple=True)
        assert float(torch.hypot((xs - 2048).float(), (ys - 2048).float()).max()) * res <= 6.0 + 2 * res
        assert int(xs.max()) == 2100 and int((explored == 100).sum()) > 20  # nothing behind the wall
    finally:
        ctx.close()
