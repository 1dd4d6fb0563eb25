#!/bin/bash
B="bench.py --steps 3 --warmup 3 --inner 100 --no-cpu-baseline --no-e2e"
run() { python $B "$@" | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print(round(d['value']/1e9,2), round(d['roofline']['frac'],4), round(d['roofline']['avg_launch_us'],1))"; }
for blk in 0 4 8; do
  export TB_TILE_BLOCK=$blk
  echo "== block $blk: c2 100^3"; run --workload c2
  echo "== block $blk: c4 150^3"; run --workload c4 --size 150
  echo "== block $blk: c4 60x240x240"; run --workload c4 --shape 60,240,240
done
