"""FrontierRoadMap::isConnectable (RE/src/planners/FrontierRoadmap.cpp:671-694), SURVEY.md 8f-2: the oracle against
hand-derived cases on CPU, the CUDA batch against the oracle on the GPU (bit-exact: integer cell walks)."""
import numpy as np
import pytest


def _open_map(W=200, H=120, res=0.05):
    a = np.zeros((H, W), np.uint8)
    meta = dict(res=res, ox=-W * res / 2, oy=-H * res / 2)
    return a, meta


def _w(meta, mx, my):
    return meta["ox"] + (mx + 0.5) * meta["res"], meta["oy"] + (my + 0.5) * meta["res"]


def test_oracle_known_answers(oracle):
    a, meta = _open_map()
    R = 2.0  # max_length = (unsigned)(3.0 / 0.05) = 60 steps; unknown limit = 2.0 / 0.05 * 0.3 = 12
    seg = lambda x0, y0, x1, y1: [*_w(meta, x0, y0), *_w(meta, x1, y1)]
    # free line: connectable, no unknown cells
    ok, unk = oracle.roadmap_connectable(a, meta, [seg(10, 10, 50, 30)], R)
    assert ok[0] == 1 and unk[0] == 0
    # a 253 or a 254 cell on the line blocks it; a 252 cell does not
    for cost, want in ((253, 0), (254, 0), (252, 1)):
        b = a.copy()
        b[10, 30] = cost
        ok, _ = oracle.roadmap_connectable(b, meta, [seg(10, 10, 50, 10)], R)
        assert ok[0] == want, cost
    # 12 unknown cells are tolerated, 13 are not (strict > against 12.0)
    for k, want in ((12, 1), (13, 0)):
        b = a.copy()
        b[10, 20:20 + k] = 255
        ok, unk = oracle.roadmap_connectable(b, meta, [seg(10, 10, 50, 10)], R)
        assert unk[0] == k and ok[0] == want
    # the walk stops after max_length steps: an obstacle at step 61 of a 100-cell line is never seen, one at step 60 is
    for x, want in ((10 + 61, 1), (10 + 60, 0)):
        b = a.copy()
        b[10, x] = 254
        ok, _ = oracle.roadmap_connectable(b, meta, [seg(10, 10, 110, 10)], R)
        assert ok[0] == want, x
    # an end off the map: getTracedCells fails, not connectable
    ok, unk = oracle.roadmap_connectable(a, meta, [[meta["ox"] - 1.0, 0.0, 0.0, 0.0], [0.0, 0.0, 0.0, -meta["oy"] + 5.0]], R)
    assert list(ok) == [0, 0] and list(unk) == [-1, -1]
    # zero-length segment: one cell
    b = a.copy()
    b[40, 40] = 255
    ok, unk = oracle.roadmap_connectable(b, meta, [seg(40, 40, 40, 40)], R)
    assert ok[0] == 1 and unk[0] == 1


def _random_segments(rng, meta, W, H, n, reach):
    x0 = rng.uniform(meta["ox"] - 0.3, meta["ox"] + W * meta["res"] + 0.3, n)
    y0 = rng.uniform(meta["oy"] - 0.3, meta["oy"] + H * meta["res"] + 0.3, n)
    ang = rng.uniform(0, 2 * np.pi, n)
    ln = rng.uniform(0, reach, n)
    seg = np.stack([x0, y0, x0 + ln * np.cos(ang), y0 + ln * np.sin(ang)], axis=1)
    seg[::17, 2:] = seg[::17, :2]          # zero-length
    seg[1::19, 3] = seg[1::19, 1]          # axis-aligned
    seg[2::23, 2] = seg[2::23, 0]
    return seg


@pytest.mark.gpu
@pytest.mark.parametrize("radius", [2.6, 0.4, 7.0])
def test_gpu_connectable_matches_oracle(capi, oracle, synth, radius):
    a, meta, pts, ov = synth.make_config("C1")
    H, W = a.shape
    rng = np.random.default_rng(7)
    b = a.copy()
    idx = rng.integers(0, H * W, 4000)
    b.reshape(-1)[idx] = rng.choice(np.array([252, 253, 254, 255, 0, 128], np.uint8), len(idx))  # every band on some line
    seg = _random_segments(rng, meta, W, H, 20000, 3.0 * radius)
    ctx = capi.Context()
    try:
        ctx.set_costmap(b, meta["res"], meta["ox"], meta["oy"])
        ok, unk = ctx.roadmap_connectable(seg, radius)
        wok, wunk = oracle.roadmap_connectable(b, meta, seg, radius)
        np.testing.assert_array_equal(unk, wunk)
        np.testing.assert_array_equal(ok, wok)
        assert 0 < ok.sum() < len(ok) and (wunk < 0).any()
        ok2, none = ctx.roadmap_connectable(seg[:5], radius, want_unknown=False)
        assert none is None and list(ok2) == list(wok[:5])
    finally:
        ctx.close()
