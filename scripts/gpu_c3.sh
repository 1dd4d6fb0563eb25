#!/bin/bash
fmt() { grep '^{' | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); t=d['config'].get('tiling',{}); print(round(d['value']/1e9,2), 'frac', round(d['roofline']['frac'],3), 'us', round(d['roofline']['avg_launch_us'],1), 'dict', t.get('dict_records'), 'smem', t.get('smem_bytes'), 'grid', t.get('grid'))"; }
timeout 900 python -m pytest tests -m gpu -q -x -k "not facade" 2>&1 | tail -3
echo "c3: $(timeout 600 python bench.py --workload c3 --steps 5 --warmup 3 --no-cpu-baseline --no-e2e 2>/dev/null | fmt)"
echo "c3 plain records: $(TB_DICT=0 timeout 600 python bench.py --workload c3 --steps 5 --warmup 3 --no-cpu-baseline --no-e2e 2>/dev/null | fmt)"
echo "c2: $(timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-e2e 2>/dev/null | fmt)"
