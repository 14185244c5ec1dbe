"""Stage S parity: FrontierSearch::searchFrom on the GPU vs the oracle.  Everything is integer or exact-grid data,
so the bar is bit-exact: frontier list (goal cell, goal point, size, component, chunk, offsets) in discovery order
and the every_frontier cell list in acceptance order."""
import numpy as np
import pytest

from oracle import ref_binding

pytestmark = pytest.mark.gpu


def _compare(capi, oracle, a, meta, rx, ry, ov=None, expect_frontiers=True):
    params = capi.default_params()
    for k, v in (ov or {}).items():
        setattr(params, k, v)
        setattr(oracle.params, k, v)
    wst, wfr, wcells, wseeds, wcent = oracle.search_frontiers(a, meta, rx, ry, capacity=1 << 18, cell_capacity=1 << 23)
    ctx = capi.Context(params)
    try:
        ctx.set_costmap(a, meta["res"], meta["ox"], meta["oy"])
        for _ in range(2):  # twice: the second run must find the claim array clean
            st, fr, cells = ctx.search_frontiers(rx, ry, capacity=1 << 18, cell_capacity=1 << 23)
            assert st == wst
            assert len(cells) == len(wcells)
            np.testing.assert_array_equal(cells, wcells)
            assert len(fr) == len(wfr)
            for f in ("goal_mx", "goal_my", "size", "component", "chunk", "first_cell"):
                np.testing.assert_array_equal(fr[f], wfr[f], err_msg=f)
            np.testing.assert_array_equal(fr["goal_x"].view(np.uint64), wfr["goal_x"].view(np.uint64))
            np.testing.assert_array_equal(fr["goal_y"].view(np.uint64), wfr["goal_y"].view(np.uint64))
        if expect_frontiers:
            assert len(fr) > 0
        if ref_binding.available() and a.size <= 1 << 22:
            # ... and against the REFERENCE itself (oracle/_ref: DEP/src/FrontierSearch.cpp compiled where it lies)
            n, rfr, rcells = ref_binding.Reference().search_frontiers(a, meta, params, rx, ry, capacity=1 << 18, cell_capacity=1 << 23)
            assert max(n, 0) == len(fr)
            np.testing.assert_array_equal(fr["goal_x"].view(np.uint64), rfr[:, 0].copy().view(np.uint64))
            np.testing.assert_array_equal(fr["goal_y"].view(np.uint64), rfr[:, 1].copy().view(np.uint64))
            np.testing.assert_array_equal(fr["size"].astype(np.float64), rfr[:, 2])
            W = a.shape[1]
            cw = np.stack([meta["ox"] + (cells % W + 0.5) * meta["res"], meta["oy"] + (cells // W + 0.5) * meta["res"]], axis=1)
            np.testing.assert_array_equal(cw.view(np.uint64), rcells.copy().view(np.uint64))
        return fr, cells
    finally:
        ctx.close()


@pytest.mark.parametrize("name", ["C1", "C3", "C2"])
def test_search_matches_oracle_on_config(capi, oracle, synth, name):
    a, meta, pts, ov = synth.make_config(name, reference_rays=True)
    _compare(capi, oracle, a, meta, meta["robot_x"], meta["robot_y"], ov)


@pytest.mark.parametrize("seed", [1, 2, 3])
def test_search_inflated_maps_small_radius(capi, oracle, synth, seed):
    # inflation ring (costs 1..253: traversable < 254 but lethal >= 160), search radius smaller than the explored
    # region (cells beyond it are visited but not expanded), odd map width (pitch != W, partial mask words)
    a, meta = synth.make_costmap(1001, 777, seed=seed, n_disc=14, r_lo=25, r_hi=70, n_holes=30, inflation=5)
    _compare(capi, oracle, a, meta, meta["robot_x"], meta["robot_y"], dict(max_frontier_distance=6.0 + seed))


def test_search_cluster_size_parameters(capi, oracle, synth):
    a, meta = synth.make_costmap(800, 800, seed=5, n_disc=10, r_lo=30, r_hi=80, n_holes=20)
    for mn, mx in ((1, 20), (3, 7), (0, 40), (5, 62)):
        _compare(capi, oracle, a, meta, meta["robot_x"], meta["robot_y"],
                 dict(min_frontier_cluster_size=mn, max_frontier_cluster_size=mx))


def test_search_strip_and_map_border(capi, oracle):
    # the 60-cell strip of tests/test_oracle_micro.py, and a frontier running along the map border
    W, H = 80, 9
    a = np.full((H, W), 254, np.uint8)
    a[3, 10:70] = 0
    a[4, 10:70] = 255
    a[5, 10:70] = 255
    a[4, 9] = a[4, 70] = 255
    meta = dict(res=0.05, ox=0.0, oy=0.0)
    fr, cells = _compare(capi, oracle, a, meta, (10 + 0.5) * 0.05, (3 + 0.5) * 0.05)
    assert sorted(fr["size"].tolist(), reverse=True) == [21, 21, 18]
    b = np.full((40, 50), 255, np.uint8)
    b[1:39, 1:49] = 0          # free interior, unknown ring on the border rows/columns
    b[10:12, 20:30] = 254
    _compare(capi, oracle, b, dict(res=0.1, ox=-2.5, oy=-2.0), 0.0, 0.0)


def test_search_robot_on_lethal_or_unknown_cell(capi, oracle, synth):
    a, meta = synth.make_costmap(600, 600, seed=11, n_disc=8, r_lo=30, r_hi=70, n_holes=10)
    cx, cy = 300, 300
    a[cy - 4:cy + 5, cx - 4:cx + 5] = 254   # nearestFreeCell has to walk out of a lethal block
    _compare(capi, oracle, a, meta, meta["robot_x"], meta["robot_y"])
    a[cy - 4:cy + 5, cx - 4:cx + 5] = 255
    a[cy + 3, cx - 2] = 200                  # a non-free (>= lethal_threshold) traversable cell is not "free"
    _compare(capi, oracle, a, meta, meta["robot_x"], meta["robot_y"])


def test_search_out_of_bounds_and_empty(capi, oracle):
    a = np.zeros((64, 64), np.uint8)
    meta = dict(res=0.05, ox=0.0, oy=0.0)
    ctx = capi.Context()
    ctx.set_costmap(a, 0.05, 0.0, 0.0)
    st, fr, cells = ctx.search_frontiers(-1.0, 1.0)
    assert st == capi.E_ROBOT_OUT_OF_BOUNDS and len(fr) == 0
    st, fr, cells = ctx.search_frontiers(1.0, 64 * 0.05 + 0.01)
    assert st == capi.E_ROBOT_OUT_OF_BOUNDS
    ctx.close()
    _compare(capi, oracle, a, meta, 1.0, 1.0, expect_frontiers=False)  # fully known map: no frontier at all


def test_search_random_noise_maps(capi, oracle):
    # adversarial: salt-and-pepper maps make many tiny components and many same-level ties in the wavefront
    rng = np.random.default_rng(4)
    for W, H in ((97, 61), (256, 200)):
        a = rng.choice(np.array([0, 255, 254, 100, 200], np.uint8), size=(H, W), p=[0.55, 0.3, 0.05, 0.05, 0.05])
        a[H // 2, W // 2] = 0
        meta = dict(res=0.05, ox=-1.0, oy=2.0)
        _compare(capi, oracle, a, meta, meta["ox"] + (W // 2 + 0.5) * 0.05, meta["oy"] + (H // 2 + 0.5) * 0.05)


@pytest.mark.parametrize("name", ["C1", "C3"])
def test_search_global_memory_wavefronts(capi, oracle, synth, name, monkeypatch):
    # FITSLAM_WAVEFRONT=global selects the global-memory wavefront kernels (the fallback for levels wider than the
    # shared-memory queue and for components above 4096 cells); both paths must give the oracle's lists
    monkeypatch.setenv("FITSLAM_WAVEFRONT", "global")
    a, meta, pts, ov = synth.make_config(name, reference_rays=True)
    _compare(capi, oracle, a, meta, meta["robot_x"], meta["robot_y"], ov)


@pytest.mark.parametrize("wide", [True, False])
def test_search_wide_levels_fall_back(capi, oracle, monkeypatch, wide):
    # an open 3000 x 3000 map: the diamond wavefront grows past the (here lowered) shared-memory queue capacity, so the
    # shared-memory outer wavefront hands over to the global-memory kernel mid-search, which resumes from the global
    # queue; one huge ring component (> 4096 cells) goes to the global-memory replay likewise.
    # wide=True: levels above the capacity spill to the global-memory queue inside the same kernel instead
    monkeypatch.setenv("FITSLAM_OW_QCAP", "3000")
    if not wide:
        monkeypatch.setenv("FITSLAM_OW_NOWIDE", "1")
    # (resolution 1/16 m keeps every cell centre an exact double; test_search_straight_frontiers_at_5cm is the same square
    # at the reference's 0.05 m, where the median sort hangs on the last bit of atan2)
    W = H = 3000
    a = np.full((H, W), 255, np.uint8)
    a[200:2800, 200:2800] = 0
    meta = dict(res=0.0625, ox=-93.75, oy=-93.75)
    _compare(capi, oracle, a, meta, 0.0, 0.0, dict(max_frontier_distance=250.0))


@pytest.mark.parametrize("thr", [1, 100, 127, 128, 129, 200, 253, 254, 255])
def test_search_lethal_thresholds(capi, oracle, thr):
    # the mask kernel compares four cost bytes at a time against the lethal threshold: every kind of threshold the context
    # admits (1 .. 254: below / at / above 128, the ends of the range) on a map that holds every cost value, odd width
    if thr > 254:
        p = capi.default_params()
        p.lethal_threshold = thr
        with pytest.raises(capi.FitslamError):
            capi.Context(p)
        return
    rng = np.random.default_rng(thr)
    H, W = 203, 301
    a = rng.integers(0, 256, size=(H, W)).astype(np.uint8)
    a[rng.random((H, W)) < 0.35] = 255
    a[rng.random((H, W)) < 0.35] = 0
    a[90:110, 140:160] = 0  # the robot stands in free space
    meta = dict(res=0.05, ox=-7.5, oy=-5.0)
    _compare(capi, oracle, a, meta, meta["ox"] + 150.5 * 0.05, meta["oy"] + 100.5 * 0.05,
             dict(lethal_threshold=thr, max_frontier_distance=30.0), expect_frontiers=False)


@pytest.mark.parametrize("name", ["C1", "C3", "C2"])
def test_search_on_the_cluster_wavefront(capi, oracle, synth, name, monkeypatch):
    # FITSLAM_OW_CLUSTER_MIN=1: every search runs on the thread-block cluster kernel, whatever the width of its levels (the
    # default keeps narrow searches on one CTA)
    monkeypatch.setenv("FITSLAM_OW_CLUSTER_MIN", "1")
    a, meta, pts, ov = synth.make_config(name, reference_rays=True)
    _compare(capi, oracle, a, meta, meta["robot_x"], meta["robot_y"], ov)
    ctx = capi.Context(synth.apply_overrides(capi.default_params(), ov))
    try:
        ctx.set_costmap(a, meta["res"], meta["ox"], meta["oy"])
        ctas = [ctx.cycle(meta["robot_x"], meta["robot_y"], 0.0, meta["poly"], capi.FISHER_OFF)[1].wavefront_ctas for _ in range(2)]
        assert all(n in (8, 16) for n in ctas), ctas
    finally:
        ctx.close()


def test_search_cluster_wavefront_falls_back(capi, oracle, monkeypatch):
    # the cluster wavefront forced on the same open map with a winner list of 100 entries per CTA: the level in which a CTA
    # appends more than that ends the kernel, and the global-memory wavefront reruns the search (the touch keys already written
    # are the ones it writes again)
    monkeypatch.setenv("FITSLAM_OW_CLUSTER_MIN", "1")
    monkeypatch.setenv("FITSLAM_OW_QCAP", "100")
    W = H = 3000
    a = np.full((H, W), 255, np.uint8)
    a[200:2800, 200:2800] = 0
    meta = dict(res=0.0625, ox=-93.75, oy=-93.75)
    _compare(capi, oracle, a, meta, 0.0, 0.0, dict(max_frontier_distance=250.0))


@pytest.mark.parametrize("shape", ["h", "v", "square"])
def test_search_straight_frontiers_at_5cm(capi, oracle, shape):
    # straight axis-aligned frontiers found from their middle at the reference's resolution (0.05 m): the chunks are two
    # collinear branches, the centroid is off the line by rounding noise (~1e-15), and the order of the angles is decided in
    # their last bit.  The device decides what a correctly rounded atan2 decides for certain and hands the rest to the host's
    # libm; goal cells equal the oracle's (and, where oracle/_ref is built, the reference's own) everywhere.
    if shape == "square":
        W = H = 1200
        a = np.full((H, W), 255, np.uint8)
        a[100:H - 100, 100:W - 100] = 0
        cx, cy = W // 2, H // 2
    else:
        W, H = (900, 30) if shape == "h" else (30, 900)
        a = np.full((H, W), 255, np.uint8)
        if shape == "h":
            a[H // 2:, :] = 0
            a[H // 2 + 5:, :] = 254
            cx, cy = W // 2, H // 2 + 1
        else:
            a[:, W // 2:] = 0
            a[:, W // 2 + 5:] = 254
            cx, cy = W // 2 + 1, H // 2
    meta = dict(res=0.05, ox=-W * 0.05 / 2, oy=-H * 0.05 / 2)
    fr, cells = _compare(capi, oracle, a, meta, meta["ox"] + (cx + 0.5) * 0.05, meta["oy"] + (cy + 0.5) * 0.05, dict(max_frontier_distance=100.0))
    assert len(fr) >= 40


def test_search_host_sorted_chunks_equal_device_sorted(capi, oracle, synth, monkeypatch):
    # FITSLAM_FIX_ALL=1 sends every chunk's median sort to the host path (gather, libm + std::sort on the host, scatter):
    # the same lists as the oracle, and as the default path
    a, meta, pts, ov = synth.make_config("C1", reference_rays=True)
    monkeypatch.setenv("FITSLAM_FIX_ALL", "1")
    _compare(capi, oracle, a, meta, meta["robot_x"], meta["robot_y"], ov)
