"""The N > 1 path on CPU: world_size-2 `gloo` processes exercise the host-side sharding (fit-slam_b200/shard.py, the
mirror of shard_block / shard_range in csrc/ctx.h) and the collectives of the path: the integer all-reduce of the cycle's
zero-padded record rows (frontier f on rank f % world) and, for the stand-alone scoring calls, the in-place all-gather of the
16-byte score records.  Each rank scores ITS block of frontiers -- here with the CPU oracle standing in
for the kernels, which is what a checker is for -- and after the all-gather every rank must hold exactly the records a
single rank computes for the whole list."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out_dir):
    import torch
    import torch.distributed as dist
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from conftest import load_package
    import importlib
    pkg = load_package()
    capi = pkg.capi
    shard = importlib.import_module("fitslam_b200.shard")
    synth = importlib.import_module("fitslam_b200.synth")
    from oracle.binding import Oracle
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        a, meta = synth.make_costmap(400, 360, seed=21, n_disc=8, r_lo=25, r_hi=60, n_holes=10)
        orc = Oracle(capi)
        st, fr, cells, seeds, cent = orc.search_frontiers(a, meta, meta["robot_x"], meta["robot_y"])  # replicated
        assert st == 0 and len(fr) > 3
        F = len(fr)
        poly = [meta["ox"], meta["oy"], meta["ox"] + a.shape[1] * meta["res"], meta["oy"] + a.shape[0] * meta["res"]]
        first, count = shard.shard_range(F, rank, world)
        B = shard.shard_block(F, world)
        rec = np.zeros(world * B, shard.RECORD_DTYPE)
        if count:
            arr, _, _ = orc.score_arrival(a, meta, fr[first:first + count], poly)
            rec[first:first + count] = shard.pack_records(arr)
        t = torch.from_numpy(rec.view(np.int32).reshape(-1, 4))
        shard.allgather_records_(t, rank, world)
        # the single-rank answer
        arr_all, _, _ = orc.score_arrival(a, meta, fr, poly)
        want = np.zeros(world * B, shard.RECORD_DTYPE)
        want[:F] = shard.pack_records(arr_all)
        got = t.numpy().reshape(-1).view(shard.RECORD_DTYPE)
        assert np.array_equal(got, want), "merged records differ from the single-rank records"
        np.save(os.path.join(out_dir, f"rec{rank}.npy"), t.numpy())
        # the cycle's sharding: frontier f on rank f % world, zero-padded rows, merged by an integer all-reduce
        own = np.array([shard.owner(f, world) == rank for f in range(F)])
        rec2 = np.zeros(F, shard.RECORD_DTYPE)
        arr_own, _, _ = orc.score_arrival(a, meta, fr[own], poly)
        fisher = (np.arange(F, dtype=np.float32) * -0.37 + 1.25)[own]  # any bit patterns: they must survive the sum
        rec2[own] = shard.pack_records(arr_own, fisher=fisher)
        t2 = torch.from_numpy(rec2.view(np.int32).reshape(-1, 4))
        shard.allreduce_records_(t2, world)
        want2 = shard.pack_records(arr_all, fisher=np.arange(F, dtype=np.float32) * -0.37 + 1.25)
        got2 = t2.numpy().reshape(-1).view(shard.RECORD_DTYPE)
        assert got2.tobytes() == want2.tobytes(), "all-reduced records differ from the single-rank records"
        np.save(os.path.join(out_dir, f"red{rank}.npy"), t2.numpy())
    finally:
        dist.destroy_process_group()


def test_shard_ranges_cover_exactly_once():
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from conftest import load_package
    import importlib
    load_package()
    shard = importlib.import_module("fitslam_b200.shard")
    for n in (0, 1, 2, 7, 8, 9, 1788, 20000):
        for w in (1, 2, 3, 4, 8):
            seen = np.zeros(n, np.int32)
            b = shard.shard_block(n, w)
            assert b * w >= n and (n == 0 or (b - 1) * w < n)
            for r in range(w):
                first, count = shard.shard_range(n, r, w)
                assert first == min(n, r * b) and first + count <= n and count <= b
                seen[first:first + count] += 1
            assert np.all(seen == 1)
    assert shard.RECORD_DTYPE.itemsize == 16


def test_two_rank_allgather_of_score_records(tmp_path):
    import torch.multiprocessing as mp
    port = _free_port()
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    r0 = np.load(tmp_path / "rec0.npy")
    r1 = np.load(tmp_path / "rec1.npy")
    assert np.array_equal(r0, r1) and r0.any()
    d0 = np.load(tmp_path / "red0.npy")
    d1 = np.load(tmp_path / "red1.npy")
    assert np.array_equal(d0, d1) and d0.any()
