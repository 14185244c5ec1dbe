"""Repeats the cycle many times on several maps and checks that every repetition gives the bits of the first one, and
that the overlapped default equals the unoverlapped order of launches (FITSLAM_OW_PARTS=1 + FITSLAM_NO_EARLY_SCORING=1,
run in a child process).  Looks for races between the side streams of the search."""
import os, sys, hashlib, subprocess, json
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
pkg, capi, synth = bench.load()


def digest(out):
    h = hashlib.sha256()
    for k in sorted(out):
        h.update(out[k].tobytes())
    return h.hexdigest()


def run(reps):
    res = {}
    for name, seed in (("C1", None), ("C2", None), ("C3", None)):
        a, meta, pts, ov = synth.make_config(name)
        params = synth.apply_overrides(capi.default_params(), ov)
        ctx = capi.Context(params)
        ctx.set_costmap(a, meta["res"], meta["ox"], meta["oy"]); ctx.fisher_generate_lookup(); ctx.set_map_points(pts)
        first = None
        for mode in (capi.FISHER_ALL_YAWS, capi.FISHER_BEST_YAW):
            for i in range(reps):
                st, r = ctx.cycle(meta["robot_x"], meta["robot_y"], 0.3, meta["poly"], mode)
                assert st == 0
                d = digest(ctx.cycle_fetch(r.n_frontiers))
                key = f"{name}/{mode}"
                if key not in res:
                    res[key] = d
                assert d == res[key], f"{key}: repetition {i} differs"
        ctx.close()
    return res


if __name__ == "__main__":
    reps = int(sys.argv[1]) if len(sys.argv) > 1 else 50
    mine = run(reps)
    if os.environ.get("STRESS_CHILD"):
        print(json.dumps(mine))
    else:
        env = dict(os.environ, STRESS_CHILD="1", FITSLAM_OW_PARTS="1", FITSLAM_NO_EARLY_SCORING="1")
        out = subprocess.run([sys.executable, __file__, "2"], env=env, capture_output=True, text=True)
        assert out.returncode == 0, out.stderr[-2000:]
        ref = json.loads(out.stdout.strip().splitlines()[-1])
        assert ref == mine, "overlapped and unoverlapped runs differ"
        print(f"ok: {reps} repetitions x {len(mine)} (map, mode) pairs identical, and equal to the unoverlapped run")
