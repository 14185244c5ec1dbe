#!/bin/bash
# usage: scratch/ncu_lines.sh <report.ncu-rep> <object stem, e.g. fisher> <kernel substring> <units> [top]
set -e
rep=$1; stem=$2; kern=$3; units=$4; top=${5:-50}
tmp=/tmp/sass_$stem; mkdir -p $tmp; (cd $tmp && rm -f *.cubin && cuobjdump -xelf all /root/repo/fit-slam_b200/libfitslam_cuda.so > /dev/null)
nvdisasm -g -c $tmp/$stem.sm_100a.cubin > $tmp/$stem.disasm 2>/dev/null
ncu -i $rep --page source --csv > $tmp/source.csv 2>/dev/null
python /root/repo/scratch/sass_lines.py $tmp/$stem.disasm $kern $tmp/source.csv $units $top
