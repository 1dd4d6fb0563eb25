#!/bin/bash
# tile-ring depth (TB_STAGES) on the current library
B="bench.py --steps 3 --warmup 3 --inner 100 --no-cpu-baseline --no-e2e"
run() { python $B "$@" 2>gpurun_out/err.log | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print(round(d['value']/1e9,2), round(d['roofline']['avg_launch_us'],1), d['config'].get('tiling',{}).get('smem_bytes'), d['config'].get('tiling',{}).get('grid'))"; grep "tb profile" gpurun_out/err.log; }
for rep in 1 2; do
for st in 2 3; do
echo "== c2 stages=$st: $(TB_STAGES=$st run --workload c2)"
echo "== 150 stages=$st: $(TB_STAGES=$st run --workload c4 --size 150)"
done; done
echo "== c3 stages=2: $(TB_STAGES=2 run --workload c3)"
echo "== c3 stages=3: $(TB_STAGES=3 run --workload c3)"
echo "== c2 stages=2 lite: $(TB_STAGES=2 TB_PROFILE=2 run --workload c2)"
echo "== c2 stages=3 lite: $(TB_STAGES=3 TB_PROFILE=2 run --workload c2)"
echo "== c2 stages=3 FORCE_HALO: $(TB_FORCE_HALO=1 run --workload c2)"
