import os, sys, time, importlib
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
pkg, capi, synth = bench.load()
wl = bench.Workload("C2", capi, synth)
ctx = capi.Context(wl.params(capi.default_params()), device=0)
a, meta, pts = wl.a, wl.meta, wl.pts
a_pin = torch.from_numpy(a).pin_memory().numpy(); pts_pin = torch.from_numpy(pts).pin_memory().numpy()
ctx.set_costmap(a_pin, meta["res"], meta["ox"], meta["oy"]); ctx.fisher_generate_lookup(); ctx.set_map_points(pts_pin)
rx, ry, poly = meta["robot_x"], meta["robot_y"], meta["poly"]
T = np.zeros(4)
for it in range(8):
    t0 = time.perf_counter(); ctx.set_costmap(a_pin, meta["res"], meta["ox"], meta["oy"])
    t1 = time.perf_counter(); ctx.set_map_points(pts_pin)
    t2 = time.perf_counter(); st, res = ctx.cycle(rx, ry, wl.ryaw, poly, wl.fisher_mode)
    t3 = time.perf_counter(); out = ctx.cycle_fetch(res.n_frontiers)
    t4 = time.perf_counter()
    if it >= 3: T += np.array([t1 - t0, t2 - t1, t3 - t2, t4 - t3])
print("ms: set_costmap %.2f set_map_points %.2f cycle %.2f fetch %.2f" % tuple(T / 5 * 1e3), [round(v, 2) for v in res.stage_ms])
