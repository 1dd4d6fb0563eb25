#!/bin/bash
fmt() { grep '^{' | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); t=d['config'].get('tiling',{}); print(round(d['value']/1e9,2), 'frac', round(d['roofline']['frac'],3), 'us', round(d['roofline']['avg_launch_us'],1), 'kernel', d['roofline']['kernel'])"; }
timeout 900 python -m pytest tests -m gpu -q 2>&1 | tail -3
echo "c2 euler:  $(timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-e2e 2>/dev/null | fmt)"
echo "c2 verlet: $(timeout 300 python bench.py --verlet --steps 5 --warmup 3 --no-cpu-baseline --no-e2e 2>/dev/null | fmt)"
echo "c2 verlet generic kernel: $(timeout 300 python bench.py --verlet --csr --steps 3 --warmup 3 --no-cpu-baseline --no-e2e 2>/dev/null | fmt)"
echo "c2 euler generic kernel: $(timeout 300 python bench.py --csr --steps 3 --warmup 3 --no-cpu-baseline --no-e2e 2>/dev/null | fmt)"
