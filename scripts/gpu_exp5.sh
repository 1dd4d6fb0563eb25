#!/bin/bash
B="bench.py --steps 3 --warmup 3 --inner 100 --no-cpu-baseline --no-e2e"
run() { python $B "$@" | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print(round(d['value']/1e9,2), round(d['roofline']['frac'],4), round(d['roofline']['avg_launch_us'],1))"; }
echo "== c4 150^3 plain"; run --workload c4 --size 150
echo "== c4 150^3 TB_FORCE_HALO"; TB_FORCE_HALO=1 run --workload c4 --size 150
echo "== c4 95x189x189 plain"; run --workload c4 --shape 95,189,189
echo "== c4 95x189x189 TB_FORCE_HALO"; TB_FORCE_HALO=1 run --workload c4 --shape 95,189,189
echo "== c4 95x190x190 plain"; run --workload c4 --shape 95,190,190
