"""Timing of fitslam_set_occupancy_grid with a device-resident int8 grid: W x H bytes read + W x H bytes written."""
import os, sys, ctypes as C
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
pkg, capi, synth = bench.load()
ctx = capi.Context()
p = capi.default_occupancy_params()
stream = torch.cuda.ExternalStream(ctx.stream)
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for n in (4096, 16384):
    g = torch.randint(-1, 101, (n, n), dtype=torch.int8, device="cuda")
    call = lambda: ctx.lib.fitslam_set_occupancy_grid(ctx.h, C.c_void_p(g.data_ptr()), n, n, 0.05, 0.0, 0.0, C.byref(p), None)
    for _ in range(3): assert call() == 0
    tot = 0.0
    for _ in range(10):
        with torch.cuda.stream(stream):
            flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream); call(); e1.record(stream); e1.synchronize()
        tot += e0.elapsed_time(e1)
    ms = tot / 10
    print(f"{n}x{n}: {ms*1e3:.1f} us per grid, {2*n*n/ms/1e6:.0f} GB/s (read + write), L2 flushed between calls")
