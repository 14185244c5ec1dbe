#!/bin/bash
# ncu --set full of the top kernels at the head of the round (each stage one launch per cycle: FITSLAM_OW_PARTS=1
# FITSLAM_NO_EARLY_SCORING=1): C2 kernels, then the cluster wavefront at C4.  Each capture only after the plain run exited 0.
OUT=gpurun_out
export FITSLAM_OW_PARTS=1 FITSLAM_NO_EARLY_SCORING=1
python scratch/cycle_probe.py C2 2 all > $OUT/plain_c2.log 2>&1 &&
timeout 900 ncu --set full --import-source on --clock-control none \
    -k regex:"ow_kernel|ow_prep_kernel|fisher_kernel|replay_line_kernel|replay_warp_kernel|raycast_kernel|mask_kernel|chunk_kernel" -c 40 -f \
    -o $OUT/r2_top_c2 python scratch/cycle_probe.py C2 2 all > $OUT/ncu_c2.log 2>&1
python scratch/cycle_probe.py C4 2 off > $OUT/plain_c4.log 2>&1 &&
timeout 900 ncu --set full --import-source on --clock-control none -k regex:"ow_cluster_kernel" -c 1 -f \
    -o $OUT/r2_top_c4 python scratch/cycle_probe.py C4 2 off > $OUT/ncu_c4.log 2>&1
for f in $OUT/ncu_c2.log $OUT/ncu_c4.log; do tail -n 2 $f; done
ls -la $OUT/*.ncu-rep
