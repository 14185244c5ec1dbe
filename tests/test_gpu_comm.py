"""The NCCL communicator of the C ABI (csrc/comm.cu): with a communicator set, fitslam_set_costmap /
fitslam_set_map_points are collective (rank 0 uploads, the others receive over NVLink) and fitslam_cycle merges the ranks'
score records inside the library.  A one-rank communicator runs everywhere; the two-rank case needs two GPUs."""
import multiprocessing as mp
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _cycle_bytes(ctx, capi, meta, mode):
    st, res = ctx.cycle(meta["robot_x"], meta["robot_y"], 0.3, meta["poly"], mode)
    assert st == 0
    out = ctx.cycle_fetch(res.n_frontiers)
    return res, {k: v.tobytes() for k, v in out.items()}


def test_one_rank_communicator_changes_nothing(capi, synth):
    a, meta, pts, ov = synth.make_config("C1")
    params = synth.apply_overrides(capi.default_params(), ov)
    plain = capi.Context(params)
    comm = capi.Context(params)
    try:
        comm.comm_init(capi.Context.comm_unique_id(), 0, 1)
        for c in (plain, comm):
            c.set_costmap(a, meta["res"], meta["ox"], meta["oy"])
            c.fisher_generate_lookup()
            c.set_map_points(pts)
        r0, want = _cycle_bytes(plain, capi, meta, capi.FISHER_ALL_YAWS)
        r1, got = _cycle_bytes(comm, capi, meta, capi.FISHER_ALL_YAWS)
        assert r0.n_frontiers == r1.n_frontiers > 0 and got == want
        rect = synth.reveal_step(a, meta, a.shape[1] // 2, a.shape[0] // 2, np.random.default_rng(3))
        for c in (plain, comm):
            c.set_costmap(a, meta["res"], meta["ox"], meta["oy"], dirty=rect)
        assert _cycle_bytes(comm, capi, meta, capi.FISHER_BEST_YAW)[1] == _cycle_bytes(plain, capi, meta, capi.FISHER_BEST_YAW)[1]
        comm.comm_destroy()
        assert _cycle_bytes(comm, capi, meta, capi.FISHER_BEST_YAW)[1] == _cycle_bytes(plain, capi, meta, capi.FISHER_BEST_YAW)[1]
    finally:
        plain.close()
        comm.close()


def test_comm_errors(capi):
    c = capi.Context()
    try:
        with pytest.raises(capi.FitslamError) as e:
            c.comm_init(bytes(128), 2, 2)  # rank out of range
        assert e.value.status == capi.E_INVALID
        c.cycle_merge()  # no communicator: nothing to do
        c.comm_destroy()
    finally:
        c.close()


def _rank_main(rank, world, uid, q):
    import importlib
    import sys
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    from conftest import load_package
    capi = load_package().capi
    synth = importlib.import_module("fitslam_b200.synth")
    a, meta, pts, ov = synth.make_config("C1")
    params = synth.apply_overrides(capi.default_params(), ov)
    ctx = capi.Context(params, device=rank)
    ctx.comm_init(uid, rank, world)
    H, W = a.shape
    # rank 0 alone holds the data
    ctx.set_costmap(a if rank == 0 else None, meta["res"], meta["ox"], meta["oy"], shape=(H, W))
    ctx.fisher_generate_lookup()
    ctx.set_map_points(pts if rank == 0 else None, n=len(pts))
    res, got = _cycle_bytes(ctx, capi, meta, capi.FISHER_ALL_YAWS)
    rect = synth.reveal_step(a, meta, W // 2, H // 2, np.random.default_rng(3))  # same seed on every rank: same rectangle
    ctx.set_costmap(a if rank == 0 else None, meta["res"], meta["ox"], meta["oy"], dirty=rect, shape=(H, W))
    res2, got2 = _cycle_bytes(ctx, capi, meta, capi.FISHER_BEST_YAW)
    q.put((rank, res.n_frontiers, got, res2.n_frontiers, got2))
    ctx.close()


def test_two_rank_broadcast_and_merge(capi, synth):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    uid = capi.Context.comm_unique_id()
    mpc = mp.get_context("spawn")
    q = mpc.Queue()
    procs = [mpc.Process(target=_rank_main, args=(r, 2, uid, q)) for r in range(2)]
    for p in procs:
        p.start()
    outs = sorted(q.get(timeout=300) for _ in procs)
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    # one GPU, no communicator: the same bits
    a, meta, pts, ov = synth.make_config("C1")
    ctx = capi.Context(synth.apply_overrides(capi.default_params(), ov))
    try:
        ctx.set_costmap(a, meta["res"], meta["ox"], meta["oy"])
        ctx.fisher_generate_lookup()
        ctx.set_map_points(pts)
        res, want = _cycle_bytes(ctx, capi, meta, capi.FISHER_ALL_YAWS)
        rect = synth.reveal_step(a, meta, a.shape[1] // 2, a.shape[0] // 2, np.random.default_rng(3))
        ctx.set_costmap(a, meta["res"], meta["ox"], meta["oy"], dirty=rect)
        res2, want2 = _cycle_bytes(ctx, capi, meta, capi.FISHER_BEST_YAW)
    finally:
        ctx.close()
    for rank, F, got, F2, got2 in outs:
        assert F == res.n_frontiers and F2 == res2.n_frontiers
        assert got == want and got2 == want2, rank
