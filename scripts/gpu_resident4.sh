#!/bin/bash
fmt() { grep '^{' | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); t=d['config'].get('tiling',{}); print(round(d['value']/1e9,2), 'us', round(d['roofline']['avg_launch_us'],2), 'tiles', t.get('n_tiles'))"; }
for n in 18 37 55 74 148; do
  echo "c3 $n robots resident:  $(TB_RESIDENT_MAX_TILES=100000 timeout 300 python bench.py --workload c3 --robots $n --steps 3 --warmup 3 --no-cpu-baseline --no-e2e 2>/dev/null | fmt)"
  echo "c3 $n robots streaming: $(TB_RESIDENT=0 timeout 300 python bench.py --workload c3 --robots $n --steps 3 --warmup 3 --no-cpu-baseline --no-e2e 2>/dev/null | fmt)"
done
