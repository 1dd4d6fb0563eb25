#!/bin/bash
# dictionary records vs one record per spring (TB_DICT=0), same library
B="bench.py --steps 3 --warmup 3 --inner 100 --no-cpu-baseline --no-e2e"
run() { python $B "$@" 2>gpurun_out/err.log | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print(round(d['value']/1e9,2), round(d['roofline']['avg_launch_us'],1), d['config'].get('tiling'))"; grep "tb profile" gpurun_out/err.log; }
for rep in 1 2; do
echo "== c2 dict: $(run --workload c2)"
echo "== c2 plain: $(TB_DICT=0 run --workload c2)"
done
echo "== c2 dict lite-profile: $(TB_PROFILE=2 run --workload c2)"
echo "== c2 plain lite-profile: $(TB_DICT=0 TB_PROFILE=2 run --workload c2)"
echo "== 150 dict: $(run --workload c4 --size 150)"
echo "== 150 plain: $(TB_DICT=0 run --workload c4 --size 150)"
echo "== c3 dict: $(run --workload c3)"
echo "== c3 plain: $(TB_DICT=0 run --workload c3)"
