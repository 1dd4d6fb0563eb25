#!/bin/bash
run() { timeout 200 titan_b200/bin/titan_facade_dump bench $1 $2 2>&1 | grep '^{' | cut -c60-200; }
echo "c3_64 streaming:           $(run c3_64 2000)"
echo "c3_64 resident (512 tiles): $(TB_RESIDENT_MAX_TILES=512 run c3_64 2000)"
echo "c1 resident:               $(run c1 10000)"
fmt() { grep '^{' | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); t=d['config'].get('tiling',{}); print(round(d['value']/1e9,2), 'us', round(d['roofline']['avg_launch_us'],1), 'tiles', t.get('n_tiles'))"; }
for n in 16 37 74 148 296; do
  echo "c3 $n robots streaming: $(timeout 300 python bench.py --workload c3 --robots $n --steps 3 --warmup 3 --no-cpu-baseline --no-e2e 2>/dev/null | fmt)"
  echo "c3 $n robots resident:  $(TB_RESIDENT_MAX_TILES=100000 timeout 300 python bench.py --workload c3 --robots $n --steps 3 --warmup 3 --no-cpu-baseline --no-e2e 2>/dev/null | fmt)"
done
