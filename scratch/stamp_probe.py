"""Timing of the two writers (SURVEY 8f-3 stamping half, 8f-4) on the C2-sized grid: CUDA events around the C-ABI call with device
buffers (sensor rays) / the resident costmap (lethal marker), and the oracle on one host core beside it."""
import math, os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
import conftest
pkg = conftest.load_package()
capi = pkg.capi
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from oracle.binding import Oracle
orc = Oracle(capi)
H = W = 4096
res, ox, oy = 0.05, -W * 0.025, -H * 0.025
rng = np.random.default_rng(5)
a = np.zeros((H, W), np.uint8)
ctx = capi.Context()
ctx.set_costmap(a, res, ox, oy)
areas = [(float(rng.uniform(-90, 90)), float(rng.uniform(-90, 90)), 3.5, float(rng.uniform(-3, 3)), int(i % 2)) for i in range(256)]
def timed(fn, reps=20):
    fn(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps): fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / reps * 1e3
ms = timed(lambda: ctx.mark_lethal(areas))
t0 = time.perf_counter(); want, n = orc.mark_lethal(a, {"res": res, "ox": ox, "oy": oy}, areas); cpu = (time.perf_counter() - t0) * 1e3
ok = np.array_equal(ctx.costmap_copy(), want)
print(f"mark_lethal: 256 areas (128 circles of 360 rays, 128 triangles of 20), {int((want == 253).sum())} cells: {ms:.3f} ms per call (host tables + upload + kernel + sync), oracle {cpu - 0:.1f} ms incl. 16 MB map copy, equal {ok}")
latest = np.zeros((H, W), np.int8); latest[rng.random((H, W)) < 0.01] = 100
t_latest = torch.from_numpy(latest).cuda(); t_exp = torch.full((H, W), -1, dtype=torch.int8, device="cuda")
p = capi.default_sensor_params(); p.sensor_min_angle, p.sensor_max_angle, p.sensor_max_range = -math.pi, math.pi, 12.0
ms_dev = timed(lambda: ctx.sensor_mark_rays(t_latest.data_ptr(), t_exp.data_ptr(), res, ox, oy, 1.0, 2.0, 0.3, p, shape=(H, W)))
exp_h = np.full((H, W), -1, np.int8)
ms_host = timed(lambda: ctx.sensor_mark_rays(latest, exp_h, res, ox, oy, 1.0, 2.0, 0.3, p))
t0 = time.perf_counter(); want, n = orc.sensor_mark_rays(latest, np.full((H, W), -1, np.int8), res, ox, oy, 1.0, 2.0, 0.3, p); cpu = (time.perf_counter() - t0) * 1e3
print(f"sensor_mark_rays: {n} rays x 241 steps: device buffers {ms_dev:.3f} ms, host buffers (rows in reach staged) {ms_host:.3f} ms, oracle {cpu:.1f} ms incl. 16 MB copy, equal {np.array_equal(exp_h, want)} {np.array_equal(t_exp.cpu().numpy(), want)}")
