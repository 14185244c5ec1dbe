#!/bin/bash
# Round-2 profile capture on the GPU box (one gpurun call): launch list of the bench command, ncu --set full of the top kernels
# at C2 (each stage one launch per cycle: FITSLAM_OW_PARTS=1 FITSLAM_NO_EARLY_SCORING=1) and of the wavefront kernel at C4.
OUT=gpurun_out
python bench.py --steps 2 --warmup 3 --no-cpu --no-c4 > $OUT/plain_bench.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file $OUT/r2_launches_cycle.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu --no-c4 > $OUT/ncu_launch.log 2>&1
export FITSLAM_OW_PARTS=1 FITSLAM_NO_EARLY_SCORING=1
python scratch/cycle_probe.py C2 2 all > $OUT/plain_c2.log 2>&1 &&
timeout 900 ncu --set full --import-source on --clock-control none \
    -k regex:"ow_kernel|ow_prep_kernel|fisher_kernel|replay_line_kernel|replay_warp_kernel|raycast_kernel|mask_kernel|chunk_kernel" -c 80 -f \
    -o $OUT/r2_top_c2 python scratch/cycle_probe.py C2 2 all > $OUT/ncu_c2.log 2>&1
python scratch/cycle_probe.py C4 2 all > $OUT/plain_c4.log 2>&1 &&
timeout 900 ncu --set full --import-source on --clock-control none -k regex:"ow_kernel|ow_cluster_kernel" -c 2 -f \
    -o $OUT/r2_top_c4 python scratch/cycle_probe.py C4 2 all > $OUT/ncu_c4.log 2>&1
for f in $OUT/ncu_launch.log $OUT/ncu_c2.log $OUT/ncu_c4.log; do tail -n 2 $f; done
