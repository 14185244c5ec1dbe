"""Search-only probe of the outer wavefront: `python scratch/ow_probe.py C2 5` runs searchFrom 5 times on config C2
(FITSLAM_PROF=1 prints the kernel's level / pass / cycle counts and the stream timeline on stderr)."""
import importlib, os, sys, time
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
from conftest import load_package
pkg = load_package()
capi = pkg.capi
synth = importlib.import_module("fitslam_b200.synth")
name = sys.argv[1] if len(sys.argv) > 1 else "C2"
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 5
a, meta, pts, ov = synth.make_config(name)
ctx = capi.Context(synth.apply_overrides(capi.default_params(), ov))
ctx.set_costmap(a, meta["res"], meta["ox"], meta["oy"])
for i in range(reps):
    t = time.perf_counter()
    st, fr, cells = ctx.search_frontiers(meta["robot_x"], meta["robot_y"], capacity=1 << 17, cell_capacity=1 << 22)
    print(name, "search", i, "status", st, "frontiers", len(fr), "cells", len(cells), "wall ms %.3f" % ((time.perf_counter() - t) * 1e3))
ctx.close()
