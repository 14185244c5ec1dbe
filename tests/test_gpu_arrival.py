"""Stage R parity: setArrivalInformationForFrontier / setMaxArrivalInformation on the GPU vs the oracle.
Integer outputs must be bit-exact; theta_s_star is a double product and must be bit-equal too."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _frontiers_from_oracle(oracle, synth, name, reference_rays, seed=None):
    a, meta, pts, ov = synth.make_config(name, seed=seed, reference_rays=reference_rays)
    synth.apply_overrides(oracle.params, ov)
    st, fr, cells, seeds, cent = oracle.search_frontiers(a, meta, meta["robot_x"], meta["robot_y"], capacity=1 << 17)
    assert st == 0 and len(fr) > 0
    return a, meta, ov, fr


def _compare(capi, oracle, a, meta, ov, fr, poly):
    params = capi.default_params()
    for k, v in ov.items():
        setattr(params, k, v)
    ctx = capi.Context(params)
    try:
        ctx.set_costmap(a, meta["res"], meta["ox"], meta["oy"])
        assert ctx.ray_shape() == oracle.ray_shape(meta["res"])
        assert ctx.arrival_limits() == oracle.arrival_limits(a, meta)
        got, grays, gwins = ctx.score_arrival(fr, poly, want_rays=True, want_windows=True)
        want, wrays, wwins = oracle.score_arrival(a, meta, fr, poly)
        np.testing.assert_array_equal(grays, wrays)
        np.testing.assert_array_equal(gwins, wwins)
        for f in ("information", "max_index", "achievable", "footprint_in_lethal", "ray_failed"):
            np.testing.assert_array_equal(got[f], want[f], err_msg=f)
        np.testing.assert_array_equal(got["theta_s_star"].view(np.uint64), want["theta_s_star"].view(np.uint64))
        return got
    finally:
        ctx.close()


@pytest.mark.parametrize("name,reference_rays", [("C1", True), ("C1", False), ("C3", True), ("C2", True), ("C2", False)])
def test_arrival_matches_oracle(capi, oracle, synth, name, reference_rays):
    a, meta, ov, fr = _frontiers_from_oracle(oracle, synth, name, reference_rays)
    got = _compare(capi, oracle, a, meta, ov, fr, meta["poly"])
    assert got["information"].max() > 0


def test_arrival_inflated_map_and_tight_polygon(capi, oracle, synth):
    # inflation ring: costs 1..253 exercise the 240..254 obstacle band; a polygon smaller than the map clamps rays
    a, meta = synth.make_costmap(1000, 1000, seed=7, n_disc=12, r_lo=30, r_hi=90, n_holes=20, inflation=6)
    st, fr, *_ = oracle.search_frontiers(a, meta, meta["robot_x"], meta["robot_y"])
    assert len(fr) > 10
    gx, gy = fr["goal_x"], fr["goal_y"]
    poly = [np.percentile(gx, 20), np.percentile(gy, 20), np.percentile(gx, 80), np.percentile(gy, 80)]
    _compare(capi, oracle, a, meta, {}, fr, poly)


def test_arrival_goals_at_map_edges_and_outside(capi, oracle, synth):
    # rays clamped at the map border, footprint disc leaving the map, a goal outside the map (ray failure)
    rng = np.random.default_rng(3)
    W, H, res = 300, 200, 0.05
    a = rng.choice(np.array([0, 255, 254, 253, 240, 239, 100], dtype=np.uint8), size=(H, W),
                   p=[0.3, 0.5, 0.04, 0.04, 0.04, 0.04, 0.04])
    meta = dict(res=res, ox=-3.0, oy=1.0)
    cells = [(0, 0), (W - 1, 0), (0, H - 1), (W - 1, H - 1), (5, 7), (W - 3, 100), (150, H - 2), (150, 0)]
    fr = np.zeros(len(cells) + 3, dtype=capi.FRONTIER_DTYPE)
    for i, (mx, my) in enumerate(cells):
        fr[i]["goal_x"], fr[i]["goal_y"] = meta["ox"] + (mx + 0.5) * res, meta["oy"] + (my + 0.5) * res
        fr[i]["size"] = 3 + i * 2
    fr[-3]["goal_x"], fr[-3]["goal_y"], fr[-3]["size"] = meta["ox"] - 1.0, meta["oy"] + 1.0, 5   # left of the map
    fr[-2]["goal_x"], fr[-2]["goal_y"], fr[-2]["size"] = meta["ox"] + W * res + 0.5, meta["oy"] + 1.0, 5
    fr[-1]["goal_x"], fr[-1]["goal_y"], fr[-1]["size"] = meta["ox"] + 1.0, meta["oy"] + H * res, 30
    poly = [meta["ox"] - 10, meta["oy"] - 10, meta["ox"] + 100, meta["oy"] + 100]
    got = _compare(capi, oracle, a, meta, {}, fr, poly)
    assert got["ray_failed"][-3:].all()


def test_arrival_random_goals_dense_noise(capi, oracle):
    # stress: random costs, goals anywhere (not only frontier medians), odd width (pitch != W)
    rng = np.random.default_rng(11)
    W, H, res = 777, 333, 0.05
    a = rng.integers(0, 256, size=(H, W), dtype=np.uint8)
    a[rng.random((H, W)) < 0.6] = 255
    meta = dict(res=res, ox=12.5, oy=-40.0)
    n = 4000
    fr = np.zeros(n, dtype=capi.FRONTIER_DTYPE)
    fr["goal_x"] = meta["ox"] + rng.uniform(0, W * res, n)
    fr["goal_y"] = meta["oy"] + rng.uniform(0, H * res, n)
    fr["size"] = rng.integers(1, 22, n)
    poly = [meta["ox"], meta["oy"], meta["ox"] + W * res, meta["oy"] + H * res]
    _compare(capi, oracle, a, meta, {}, fr, poly)


def test_sharded_scoring_union_equals_full(capi, oracle, synth):
    a, meta, ov, fr = _frontiers_from_oracle(oracle, synth, "C1", True)
    want, _, _ = oracle.score_arrival(a, meta, fr, meta["poly"])
    merged = np.zeros(len(fr), dtype=capi.ARRIVAL_DTYPE)
    for r in range(3):
        ctx = capi.Context(oracle.params)
        ctx.set_shard(r, 3)
        ctx.set_costmap(a, meta["res"], meta["ox"], meta["oy"])
        got, _, _ = ctx.score_arrival(fr, meta["poly"])
        B = (len(fr) + 2) // 3
        merged[r * B:(r + 1) * B] = got[r * B:(r + 1) * B]
        other = np.ones(len(fr), bool)
        other[r * B:(r + 1) * B] = False
        assert not got[other]["information"].any()
        ctx.close()
    np.testing.assert_array_equal(merged["information"], want["information"])
    np.testing.assert_array_equal(merged["max_index"], want["max_index"])
    np.testing.assert_array_equal(merged["achievable"], want["achievable"])
