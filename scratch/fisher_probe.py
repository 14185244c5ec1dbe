"""Fisher-only probe: `python scratch/fisher_probe.py C2 3` scores every (frontier, yaw) pose of config C2 three times."""
import importlib, os, sys, time
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
from conftest import load_package
pkg = load_package()
capi = pkg.capi
synth = importlib.import_module("fitslam_b200.synth")
name = sys.argv[1] if len(sys.argv) > 1 else "C2"
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
nf = int(sys.argv[3]) if len(sys.argv) > 3 else 0
a, meta, pts, ov = synth.make_config(name)
prm = synth.apply_overrides(capi.default_params(), ov)
ctx = capi.Context(prm)
ctx.set_costmap(a, meta["res"], meta["ox"], meta["oy"])
ctx.fisher_generate_lookup()
ctx.set_map_points(pts)
st, fr, cells = ctx.search_frontiers(meta["robot_x"], meta["robot_y"], capacity=1 << 17, cell_capacity=1 << 22)
if nf: fr = fr[:nf]
Y = 16
th = (np.arange(Y) * prm.delta_theta + prm.camera_fov / 2)
poses = np.concatenate([synth.frontier_poses(fr, t, prm.candidate_pose_z) for t in th])
for i in range(reps):
    t = time.perf_counter()
    info, safe, nvis = ctx.score_fisher(poses)
    print(name, "fisher", i, len(poses), "poses, wall ms %.3f" % ((time.perf_counter() - t) * 1e3), "sum", float(info.sum()), "nvis mean", nvis.mean())
ctx.close()
