"""Stage N parity: U1 utility block (weighted_cost bit-exact: plain IEEE double arithmetic) and
getFilteredFrontiersN (sorted order, local/global split) on the GPU vs the oracle."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _scored_ctx(capi, oracle, synth, name="C1"):
    a, meta, pts, ov = synth.make_config(name, reference_rays=True)
    synth.apply_overrides(oracle.params, ov)
    st, fr, *_ = oracle.search_frontiers(a, meta, meta["robot_x"], meta["robot_y"], capacity=1 << 17)
    ctx = capi.Context(oracle.params)
    ctx.set_costmap(a, meta["res"], meta["ox"], meta["oy"])
    arr, _, _ = ctx.score_arrival(fr, meta["poly"])
    return ctx, a, meta, fr, arr


def _compare(capi, oracle, ctx, arr, ach, pl, ph, bl):
    max_gt = ctx.arrival_limits()[0]
    wst, wwc, wua, wud, werr = oracle.normalise(arr, ach, pl, ph, bl, max_gt)
    st, wc, ua, ud, gach, srt, sel = ctx.normalise_select(pl, ph, bl)
    assert st == wst
    if wst != 0:
        return st
    for g, w in ((wc, wwc), (ua, wua), (ud, wud)):
        np.testing.assert_array_equal(g.view(np.uint64), w.view(np.uint64))
    elig = ach.astype(bool) & (~bl.astype(bool) if bl is not None else True)
    wsrt, wsel = oracle.select(elig.astype(np.uint8), pl)
    np.testing.assert_array_equal(srt, wsrt)
    for f in ("n_sorted", "n_local", "global_begin", "n_global", "closest_global"):
        assert getattr(sel, f) == getattr(wsel, f), f
    return st


def test_normalise_select_matches_oracle(capi, oracle, synth):
    ctx, a, meta, fr, arr = _scored_ctx(capi, oracle, synth)
    rng = np.random.default_rng(2)
    ach, pl, ph = oracle.plan_euclidean(meta["robot_x"], meta["robot_y"], 0.3, fr, arr["achievable"])
    # the plan step can only clear achievability; the GPU call starts from arrival.achievable, so feed lengths that
    # keep the two views equal: frontiers closer than 0.5 m are given a huge length instead
    ach = arr["achievable"].copy()
    pl = np.where(pl > 1e300, 1e6 + rng.uniform(0, 1, len(pl)), pl * (1 + 0.2 * rng.uniform(0, 1, len(pl))))
    ph = np.where(ph > 1e300, 1.0, ph)
    assert ach.sum() > 5
    assert _compare(capi, oracle, ctx, arr, ach, pl, ph, None) == 0
    bl = (rng.random(len(fr)) < 0.2).astype(np.uint8)
    assert _compare(capi, oracle, ctx, arr, ach, pl, ph, bl) == 0
    ctx.close()


@pytest.mark.parametrize("lengths", [[1, 2, 3, 4, 5, 6, 7], [1, 2, 3, 4, 5], [3.0], [20, 30, 1, 2], [13, 14, 15],
                                     [1, 2, 3, 4, 5, 6, 7, 20, 30], []])
def test_selection_known_answers(capi, oracle, synth, lengths):
    # 7 local-range frontiers -> local = 5 shortest, closest_global = the 7th; exactly 5 -> 4 + pop
    ctx, a, meta, fr, arr = _scored_ctx(capi, oracle, synth)
    n = len(fr)
    ach = arr["achievable"].copy()
    idx = np.nonzero(ach)[0]
    assert len(idx) >= len(lengths)
    bl = np.ones(n, dtype=np.uint8)
    bl[idx[:len(lengths)]] = 0
    pl = np.full(n, 5.0)
    pl[idx[:len(lengths)]] = lengths
    ph = np.full(n, 0.5)
    _compare(capi, oracle, ctx, arr, ach, pl, ph, bl)
    st, wc, ua, ud, gach, srt, sel = ctx.normalise_select(pl, ph, bl)
    if lengths == [1, 2, 3, 4, 5, 6, 7]:
        assert (sel.n_local, sel.n_global, sel.closest_global) == (5, 0, idx[6])
    if lengths == [1, 2, 3, 4, 5]:
        assert (sel.n_local, sel.n_global, sel.closest_global) == (4, 0, idx[4])
    if lengths == [3.0]:
        assert (sel.n_local, sel.n_global, sel.global_begin, sel.closest_global) == (0, 1, 0, idx[0])
    ctx.close()


def test_arrival_utility_error_status(capi, oracle, synth):
    # max_arrival_info_gt shrinks when max_arrival_factor is tiny -> utility > 1 -> the reference throws
    a, meta, pts, ov = synth.make_config("C1", reference_rays=True)
    synth.apply_overrides(oracle.params, ov)
    oracle.params.max_arrival_factor = 0.2
    oracle.params.min_arrival_factor = 0.1
    st, fr, *_ = oracle.search_frontiers(a, meta, meta["robot_x"], meta["robot_y"])
    ctx = capi.Context(oracle.params)
    ctx.set_costmap(a, meta["res"], meta["ox"], meta["oy"])
    arr, _, _ = ctx.score_arrival(fr, meta["poly"])
    n = len(fr)
    pl = np.linspace(1, 30, n)
    ph = np.full(n, 0.2)
    st = _compare(capi, oracle, ctx, arr, arr["achievable"].copy(), pl, ph, None)
    assert st == capi.E_ARRIVAL_UTILITY
    ctx.close()
