"""Extraction-only A/B: `FITSLAM_LIB=... python scratch/extract_ab.py C4 6` prints the median of the search stage (Fisher off)."""
import importlib, os, sys
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
from conftest import load_package
pkg = load_package()
capi = pkg.capi
synth = importlib.import_module("fitslam_b200.synth")
name, reps = sys.argv[1], int(sys.argv[2])
a, meta, pts, ov = synth.make_config(name)
ctx = capi.Context(synth.apply_overrides(capi.default_params(), ov))
ctx.set_costmap(a, meta["res"], meta["ox"], meta["oy"])
ext = []
for i in range(reps + 2):
    st, res = ctx.cycle(meta["robot_x"], meta["robot_y"], 0.3, meta["poly"], capi.FISHER_OFF)
    if i >= 2:
        ext.append(list(res.stage_ms)[:8])
ext = np.array(ext)
print(os.environ.get("FITSLAM_LIB", "head"), name, "frontiers", res.n_frontiers, "stage_ms median", np.round(np.median(ext, axis=0), 3).tolist())
ctx.close()
