"""Timing of fitslam_landmarks_in_view on the C2 point cloud (200 k points) and a 1 M-point cloud: device output buffer,
CUDA events on the context stream around the whole call (3 launches + the count read-back)."""
import os, sys, ctypes as C
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
pkg, capi, synth = bench.load()
ctx = capi.Context()
stream = torch.cuda.ExternalStream(ctx.stream)
rng = np.random.default_rng(11)
for n in (200_000, 1_000_000):
    pts = rng.uniform(-100, 100, size=(n, 3)).astype(np.float32); pts[:, 2] = rng.uniform(0, 3, n)
    ctx.set_map_points(pts)
    pose = np.zeros(1, capi.POSE_DTYPE); pose["z"] = 0.3; pose["qw"] = 1.0
    out = torch.zeros((n, 3), dtype=torch.float32, device="cuda"); cnt = C.c_uint32(0)
    call = lambda: ctx.lib.fitslam_landmarks_in_view(ctx.h, pose.ctypes.data, out.data_ptr(), n, C.byref(cnt))
    for _ in range(3): assert call() == 0
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(20): call()
    e1.record(stream); e1.synchronize()
    ms = e0.elapsed_time(e1) / 20
    print(f"P={n}: {ms*1e3:.1f} us per query, {cnt.value} visible, {(2*12*n + 12*cnt.value)/ms/1e6:.0f} GB/s algorithmic (two 12 B reads per point + 12 B per visible landmark)")
