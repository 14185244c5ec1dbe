#!/bin/bash
for p in "$@"; do
  FITSLAM_OW_PARTS=$p timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print('parts', $p, {k:round(v,3) for k,v in d['stage_ms'].items() if k in ('extract','outer_wavefront_kernel','fisher')}, round(d['ms_per_step'],3))"
done
