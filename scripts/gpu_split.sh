#!/bin/bash
# split dictionary: parity tests + C5 / distinct-params benches
timeout 600 python -m pytest tests/test_engine_parity.py -m gpu -q -x -k "dictionary or removed or fast_within or determinism" > gpurun_out/r02a_pytest_split.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r02a_pytest_split.log
fmt() { grep '^{' | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); t=d['config'].get('tiling',{}); print(round(d['value']/1e9,2), 'frac', round(d['roofline']['frac'],3), 'us', round(d['roofline']['avg_launch_us'],1), 'dict', t.get('dict_records'), 'smem', t.get('smem_bytes'), 'grid', t.get('grid'), 'e2e', d.get('e2e') and round(d['e2e']['value']/1e9,1))"; }
echo "c2 default:          $(timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline 2>/dev/null | fmt)"
echo "c2 distinct k plain: $(TB_DICT=0 timeout 300 python bench.py --distinct-params --steps 5 --warmup 3 --no-cpu-baseline 2>/dev/null | fmt)"
echo "c2 distinct k split: $(timeout 300 python bench.py --distinct-params --steps 5 --warmup 3 --no-cpu-baseline 2>/dev/null | fmt)"
echo "c5 plain:            $(TB_DICT=0 timeout 600 python bench.py --workload c5 --no-cpu-baseline --steps 5 2>/dev/null | fmt)"
echo "c5 split:            $(timeout 600 python bench.py --workload c5 --no-cpu-baseline --steps 5 2>/dev/null | fmt)"
