"""Cycle probe: `python scratch/cycle_probe.py C2 5 [off|best|all]` times fitslam_cycle with the given Fisher mode
(FITSLAM_PROF=1 prints the search's stream timeline on stderr)."""
import importlib, os, sys, time
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
from conftest import load_package
pkg = load_package()
capi = pkg.capi
synth = importlib.import_module("fitslam_b200.synth")
name = sys.argv[1] if len(sys.argv) > 1 else "C2"
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 5
mode = {"off": capi.FISHER_OFF, "best": capi.FISHER_BEST_YAW, "all": capi.FISHER_ALL_YAWS}[sys.argv[3] if len(sys.argv) > 3 else "all"]
a, meta, pts, ov = synth.make_config(name)
ctx = capi.Context(synth.apply_overrides(capi.default_params(), ov))
ctx.set_costmap(a, meta["res"], meta["ox"], meta["oy"])
ctx.fisher_generate_lookup()
ctx.set_map_points(pts)
for i in range(reps):
    t = time.perf_counter()
    st, res = ctx.cycle(meta["robot_x"], meta["robot_y"], 0.3, meta["poly"], mode)
    print(name, "cycle", i, "status", st, "frontiers", res.n_frontiers, "wall ms %.3f" % ((time.perf_counter() - t) * 1e3))
ctx.close()
