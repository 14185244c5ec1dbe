#!/bin/bash
# measurement refresh: tests, bench (both arms), ncu launch list, ncu full capture of the top kernels
set -x
O=gpurun_out
timeout 900 python -m pytest tests -m gpu -q > $O/pytest_gpu_v24.log 2>&1; tail -1 $O/pytest_gpu_v24.log
timeout 900 python bench.py > $O/bench_v24.json 2> $O/bench_v24.err; tail -c 600 $O/bench_v24.json
timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_v24_ref.json 2> $O/bench_v24_ref.err; tail -c 400 $O/bench_v24_ref.json
FITSLAM_NO_EARLY_SCORING=1 timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu > $O/bench_v24_noearly.json 2>/dev/null
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file $O/launches_v24.csv python bench.py --steps 2 --warmup 3 --no-cpu > $O/ncu_launch.log 2>&1; tail -2 $O/ncu_launch.log
FITSLAM_OW_PARTS=1 FITSLAM_NO_EARLY_SCORING=1 timeout 1200 ncu --set full --import-source on --clock-control none -k regex:"rank_bfs_smem|fisher_kernel|replay_warp|raycast_kernel|mask_kernel|chunk_kernel" --launch-skip 24 -c 8 -o $O/r1_v24_top -f python bench.py --steps 2 --warmup 3 --no-cpu > $O/ncu_full.log 2>&1; tail -2 $O/ncu_full.log
