#!/bin/bash
B="bench.py --steps 3 --warmup 3 --inner 100 --no-cpu-baseline --no-e2e"
run() { python $B "$@" | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print(round(d['value']/1e9,2), round(d['roofline']['frac'],4), round(d['roofline']['avg_launch_us'],1))"; }
echo "== c2"; run --workload c2
echo "== c4 150^3 plain"; run --workload c4 --size 150
echo "== c4 150^3 TB_FORCE_HALO"; TB_FORCE_HALO=1 run --workload c4 --size 150
TB_PROFILE=1 python $B --workload c4 --size 150 2>&1 | grep "tb profile"
scripts/gpu_multi.sh r1m 2 --no-e2e --inner 100 --steps 3
