"""One rank's share of a cycle on one GPU: `python scratch/shard_probe.py C4 8 5` scores the frontiers of rank 0 of 8
(fitslam_set_shard, no communicator) and prints the median cycle time and the stage times."""
import importlib, os, sys
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
from conftest import load_package
pkg = load_package()
capi = pkg.capi
synth = importlib.import_module("fitslam_b200.synth")
name, nranks, reps = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
a, meta, pts, ov = synth.make_config(name)
ctx = capi.Context(synth.apply_overrides(capi.default_params(), ov))
ctx.set_costmap(a, meta["res"], meta["ox"], meta["oy"])
ctx.fisher_generate_lookup()
ctx.set_map_points(pts)
ctx.set_shard(0, nranks)
rows = []
for i in range(reps + 2):
    st, res = ctx.cycle(meta["robot_x"], meta["robot_y"], 0.3, meta["poly"], capi.FISHER_ALL_YAWS)
    if i >= 2:
        rows.append(list(res.stage_ms)[:8])
m = np.median(np.array(rows), axis=0)
print(os.path.basename(os.environ.get("FITSLAM_LIB", "head")), name, "rank 0 of", nranks, "frontiers", res.n_frontiers,
      "cycle ms %.3f" % (m[5] + m[6]), "stage_ms", np.round(m, 3).tolist())
ctx.close()
