#!/bin/bash
python -m pytest tests -m gpu -q 2>&1 | tail -3
fmt() { grep '^{' | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); t=d['config'].get('tiling',{}); print(round(d['value']/1e9,2), 'us', round(d['roofline']['avg_launch_us'],2), 'tiles', t.get('n_tiles'), 'dict', t.get('dict_records'), 'tpm', t.get('threads_per_mass'))"; }
for n in 1 4 16; do
  echo "c3 $n robots resident:  $(timeout 300 python bench.py --workload c3 --robots $n --steps 3 --warmup 3 --no-cpu-baseline --no-e2e 2>/dev/null | fmt)"
  echo "c3 $n robots streaming: $(TB_RESIDENT=0 timeout 300 python bench.py --workload c3 --robots $n --steps 3 --warmup 3 --no-cpu-baseline --no-e2e 2>/dev/null | fmt)"
done
timeout 120 titan_b200/bin/titan_facade_dump bench c1 10000 2>&1 | grep '^{' | cut -c60-200
timeout 120 titan_b200/bin/titan_facade_dump bench lattice10 10000 2>&1 | grep '^{' | cut -c60-200
