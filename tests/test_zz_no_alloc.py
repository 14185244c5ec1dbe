"""fitslam_landmarks_in_view_static: the landmark query with its scratch in a fixed device array inside the library (no
cudaMalloc / cudaFree, nothing staged, nothing synchronised, no counterpart in
the reference).  Through the C ABI, with device pointers in and out, it must return exactly what fitslam_landmarks_in_view
returns (the same floats, compared with == like the other landmark tests) on calls in a row that share the scratch.

This file is collected last on purpose; it passes alone in a fresh process and after the
other tests."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _poses(capi, rng, n):
    p = np.zeros(n, dtype=capi.POSE_DTYPE)
    p["x"] = rng.uniform(-15, 15, n)
    p["y"] = rng.uniform(-15, 15, n)
    p["z"] = 1.0
    q = rng.normal(size=(n, 4))
    q /= np.linalg.norm(q, axis=1, keepdims=True)
    p["qx"], p["qy"], p["qz"], p["qw"] = q.T
    return p


def _static_query(ctx, torch, d_xyz, n_points, pose, capacity):
    """One call of the new entry point: device buffers made by torch (the caller's memory), results read back after a sync."""
    d_pose = torch.from_numpy(np.frombuffer(np.ascontiguousarray(pose).tobytes(), dtype=np.uint8).copy()).cuda()
    d_out = torch.full((max(capacity, 1), 3), -7.0, dtype=torch.float32, device="cuda")
    d_count = torch.full((1,), 0xDEAD, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()  # the uploads ran on torch's stream, the query runs on the context's
    st = ctx.landmarks_in_view_static(d_xyz.data_ptr() if d_xyz is not None else None, n_points, d_pose.data_ptr(), d_out.data_ptr(),
                                      capacity, d_count.data_ptr())
    torch.cuda.synchronize()
    return st, d_out.cpu().numpy(), int(d_count.cpu().numpy().view(np.uint32)[0])


def test_static_scratch_query_equals_the_allocating_one(capi):
    import torch
    rng = np.random.default_rng(2024)
    ctx = capi.Context()
    try:
        # two clouds of different sizes, so that the second call finds a longer run of stale tile counts in the shared scratch
        for n_points in (200_000, 60_001):
            pts = rng.uniform(-20, 20, size=(n_points, 3)).astype(np.float32)
            poses = _poses(capi, rng, 3)
            pts[0] = (poses[0]["x"], poses[0]["y"], poses[0]["z"])  # a landmark at the camera centre
            pts[1] = (np.nan, 0, 0)
            pts[2] = (np.inf, 1, 1)
            ctx.set_map_points(pts)
            d_xyz = torch.from_numpy(pts).cuda()
            total = 0
            for i in (0, 1, 2, 0):  # calls in a row on the same scratch; pose 0 again at the end
                st0, want, n_want = ctx.landmarks_in_view(poses[i])
                assert st0 == 0
                st, got, n_got = _static_query(ctx, torch, d_xyz, n_points, poses[i:i + 1], n_points)
                assert st == 0, ctx.lib.fitslam_last_error(ctx.h)
                assert n_got == n_want == len(want)
                assert np.array_equal(got[:n_got], want, equal_nan=True)
                assert np.all(got[n_got:] == -7.0)  # nothing written past the count
                total += n_got
                if n_want > 3:  # a short output buffer: the count is still whole, the first entries are written
                    st, part, n2 = _static_query(ctx, torch, d_xyz, n_points, poses[i:i + 1], 3)
                    assert st == 0 and n2 == n_want and np.array_equal(part[:3], want[:3], equal_nan=True)
            assert total > 0
    finally:
        ctx.close()


def test_static_scratch_limits_and_errors(capi):
    import torch
    ctx = capi.Context()
    try:
        pose = _poses(capi, np.random.default_rng(1), 1)
        d_xyz = torch.zeros((16, 3), dtype=torch.float32, device="cuda")
        # beyond the fixed scratch: refused with a clear error, nothing enqueued, the outputs untouched
        st, out, cnt = _static_query(ctx, torch, d_xyz, capi.LIV_STATIC_MAX_POINTS + 1, pose, 4)
        assert st == capi.E_CAPACITY and cnt == 0xDEAD and np.all(out == -7.0)
        msg = ctx.lib.fitslam_last_error(ctx.h).decode()
        assert "exceeds the static scratch" in msg and str(capi.LIV_STATIC_MAX_POINTS) in msg
        # no points: the count is zeroed
        st, out, cnt = _static_query(ctx, torch, None, 0, pose, 4)
        assert st == 0 and cnt == 0
        # null output pointers
        assert ctx.lib.fitslam_landmarks_in_view_static(ctx.h, d_xyz.data_ptr(), 16, None, None, 0, None) == capi.E_INVALID
        # the largest size the scratch admits runs (all points in front of the camera of an identity pose)
        n = capi.LIV_STATIC_MAX_POINTS
        big = torch.zeros((n, 3), dtype=torch.float32, device="cuda")
        big[:, 0] = 5.0
        ident = np.zeros(1, dtype=capi.POSE_DTYPE)
        ident["qw"] = 1.0
        st, out, cnt = _static_query(ctx, torch, big, n, ident, 2)
        assert st == 0 and cnt == n and np.array_equal(out[:2], np.array([[5, 0, 0], [5, 0, 0]], np.float32))
    finally:
        ctx.close()
