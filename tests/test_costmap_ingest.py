"""Costmap ingest: nav_msgs/OccupancyGrid int8 -> uint8 costmap, i.e. TraversabilityThresholdNode
(threshold_occupancy.cpp:25-49) followed by nav2's StaticLayer::interpretValue.  Oracle against hand-derived values on
CPU; the CUDA translation (full map, odd sizes, dirty rectangles, device input) against the oracle on the GPU."""
import numpy as np
import pytest


def test_oracle_known_answers(capi, oracle):
    p = capi.default_occupancy_params()
    assert (p.lethal_obstacle_threshold, p.lethal_cost_threshold, p.unknown_cost_value, p.trinary_costmap, p.track_unknown_space) == (85, 100, -1, 1, 1)
    g = np.array([-1, 0, 1, 50, 84, 85, 99, 100, 127, -2, -128], np.int8)
    # threshold node: >= 85 -> 100, -1 -> -1, everything else (negative values too) -> 0; static layer: 255 / 254 / 0
    np.testing.assert_array_equal(oracle.occupancy_to_cost(g, p), [255, 0, 0, 0, 0, 254, 254, 254, 254, 0, 0])
    # no threshold node: the static layer sees the raw bytes as unsigned char: -2 = 254 and -128 = 128 are >= 100
    p.lethal_obstacle_threshold = -1
    np.testing.assert_array_equal(oracle.occupancy_to_cost(g, p), [255, 0, 0, 0, 0, 0, 0, 254, 254, 254, 254])
    # scaled costs: (unsigned char)(value / 100.0 * 254)
    p.trinary_costmap = 0
    np.testing.assert_array_equal(oracle.occupancy_to_cost(g, p), [255, 0, 2, 127, 213, 215, 251, 254, 254, 254, 254])
    # unknown not tracked: -1 is free space
    p.track_unknown_space = 0
    assert oracle.occupancy_to_cost(g, p)[0] == 0


def _grid(rng, H, W):
    g = rng.integers(-1, 101, (H, W)).astype(np.int8)
    g[rng.random((H, W)) < 0.3] = -1
    g[rng.random((H, W)) < 0.01] = rng.integers(-128, 127, 1).astype(np.int8)[0]
    return g


@pytest.mark.gpu
@pytest.mark.parametrize("shape", [(300, 401), (512, 512), (37, 16)])
@pytest.mark.parametrize("variant", ["default", "raw", "scaled"])
def test_gpu_ingest_matches_oracle(capi, oracle, shape, variant):
    import torch
    H, W = shape
    rng = np.random.default_rng(H * 1000 + W)
    g = _grid(rng, H, W)
    p = capi.default_occupancy_params()
    if variant != "default":
        p.lethal_obstacle_threshold = -1
    if variant == "scaled":
        p.trinary_costmap = 0
    res, ox, oy = 0.05, -W * 0.025, -H * 0.025
    ctx = capi.Context()
    try:
        ctx.set_occupancy_grid(g, res, ox, oy, p)
        want = oracle.occupancy_to_cost(g, p)
        np.testing.assert_array_equal(ctx.costmap_copy(), want)
        # dirty rectangles (host input), then the whole grid again from a device buffer
        for _ in range(6):
            w, h = int(rng.integers(1, min(W, 90))), int(rng.integers(1, min(H, 90)))
            x0, y0 = int(rng.integers(0, W - w + 1)), int(rng.integers(0, H - h + 1))
            g[y0:y0 + h, x0:x0 + w] = _grid(rng, h, w)
            ctx.set_occupancy_grid(g, res, ox, oy, p, dirty=[x0, y0, w, h])
            np.testing.assert_array_equal(ctx.costmap_copy(), oracle.occupancy_to_cost(g, p))
        g2 = _grid(rng, H, W)
        t = torch.from_numpy(g2).cuda()
        ctx.set_occupancy_grid(t.data_ptr(), res, ox, oy, p, shape=(H, W))
        np.testing.assert_array_equal(ctx.costmap_copy(), oracle.occupancy_to_cost(g2, p))
        ctx.set_occupancy_grid(t.data_ptr(), res, ox, oy, p, dirty=[3, 1, min(W - 3, 33), min(H - 1, 20)], shape=(H, W))
        np.testing.assert_array_equal(ctx.costmap_copy(), oracle.occupancy_to_cost(g2, p))
    finally:
        ctx.close()


@pytest.mark.gpu
def test_search_on_ingested_grid_equals_search_on_costmap(capi, oracle, synth):
    a, meta, pts, ov = synth.make_config("C1")
    # an occupancy grid whose translation is the trinary part of the C1 costmap
    g = np.where(a == 255, -1, np.where(a >= 254, 100, 0)).astype(np.int8)
    p = capi.default_occupancy_params()
    want_map = oracle.occupancy_to_cost(g, p)
    c1, c2 = capi.Context(), capi.Context()
    try:
        c1.set_occupancy_grid(g, meta["res"], meta["ox"], meta["oy"], p)
        c2.set_costmap(want_map, meta["res"], meta["ox"], meta["oy"])
        s1, f1, cells1 = c1.search_frontiers(meta["robot_x"], meta["robot_y"])
        s2, f2, cells2 = c2.search_frontiers(meta["robot_x"], meta["robot_y"])
        assert s1 == s2 == 0 and len(f1) > 0 and f1.tobytes() == f2.tobytes() and cells1.tobytes() == cells2.tobytes()
    finally:
        c1.close(); c2.close()
