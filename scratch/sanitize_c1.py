"""Two C1 cycles + one roadmap edge batch, for compute-sanitizer (no torch import: ctypes + numpy only)."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
from conftest import load_package
import importlib
pkg = load_package()
capi = pkg.capi
synth = importlib.import_module("fitslam_b200.synth")
a, meta, pts, ov = synth.make_config("C1")
ctx = capi.Context(synth.apply_overrides(capi.default_params(), ov))
ctx.set_costmap(a, meta["res"], meta["ox"], meta["oy"]); ctx.fisher_generate_lookup(); ctx.set_map_points(pts)
for mode in (capi.FISHER_ALL_YAWS, capi.FISHER_BEST_YAW):
    st, r = ctx.cycle(meta["robot_x"], meta["robot_y"], 0.3, meta["poly"], mode)
    out = ctx.cycle_fetch(r.n_frontiers)
    print("cycle", st, r.n_frontiers, float(out["fisher"].sum()))
seg = np.random.default_rng(1).uniform(-20, 20, (2000, 4))
ok, unk = ctx.roadmap_connectable(seg, 2.6)
print("edges", int(ok.sum()))
ctx.close()
