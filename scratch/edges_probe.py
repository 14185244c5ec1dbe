"""Timing of fitslam_roadmap_connectable on the C2 costmap: n random node pairs, device buffers, CUDA events."""
import os, sys, ctypes as C
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
pkg, capi, synth = bench.load()
wl = bench.Workload("C2", capi, synth)
a, meta = wl.a, wl.meta
H, W = a.shape
ctx = capi.Context(wl.params(capi.default_params()), device=0)
ctx.set_costmap(a, meta["res"], meta["ox"], meta["oy"])
rng = np.random.default_rng(3)
radius = 2.6
for n in (10_000, 1_000_000):
    x0 = rng.uniform(meta["ox"], meta["ox"] + W * meta["res"], n); y0 = rng.uniform(meta["oy"], meta["oy"] + H * meta["res"], n)
    ang = rng.uniform(0, 2 * np.pi, n); ln = rng.uniform(0, 1.5 * radius, n)
    seg = torch.from_numpy(np.stack([x0, y0, x0 + ln * np.cos(ang), y0 + ln * np.sin(ang)], 1)).cuda()
    ok = torch.zeros(n, dtype=torch.uint8, device="cuda"); unk = torch.zeros(n, dtype=torch.int32, device="cuda")
    stream = torch.cuda.ExternalStream(ctx.stream)
    call = lambda: ctx.lib.fitslam_roadmap_connectable(ctx.h, C.c_void_p(seg.data_ptr()), n, radius, C.c_void_p(ok.data_ptr()), C.c_void_p(unk.data_ptr()))
    for _ in range(3): assert call() == 0
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(10): call()
    e1.record(stream); e1.synchronize()
    ms = e0.elapsed_time(e1) / 10
    cells = float((torch.clamp((seg[:, 2:] - seg[:, :2]).abs().max(1).values / meta["res"], max=78) + 1).sum())
    print(f"n={n}: {ms:.3f} ms per batch, {n / ms / 1e3:.1f} M pairs/s, ~{cells / 1e6:.1f} M cells read = {cells / ms / 1e6:.1f} GB/s algorithmic, connectable {int(ok.sum())}")
# parity of the last batch against the oracle on the same map
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
from oracle.binding import Oracle
orc = Oracle(capi)
wok, wunk = orc.roadmap_connectable(a, meta, seg.cpu().numpy(), radius)
print("mismatches vs oracle:", int((ok.cpu().numpy() != wok).sum()), int((unk.cpu().numpy() != wunk).sum()), "of", n)
