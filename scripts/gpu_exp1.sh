#!/bin/bash
OUT=gpurun_out; mkdir -p $OUT
B="python bench.py --steps 3 --warmup 3 --inner 100 --no-cpu-baseline --no-e2e"
for hint in 0 1; do
  for wl in "--workload c2" "--workload c4 --size 150"; do
    echo "== TB_L2_HINT=$hint $wl"
    TB_L2_HINT=$hint $B $wl 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print(d['value']/1e9, d['roofline']['frac'], d['roofline']['avg_launch_us'], d['config']['tiling'])
    "
  done
done
