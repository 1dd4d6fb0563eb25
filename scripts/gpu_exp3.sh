#!/bin/bash
B="bench.py --steps 3 --warmup 3 --inner 100 --no-cpu-baseline --no-e2e"
for dbg in 0 8 16 24; do
echo "== 2 GPU c4 189 TB_DEBUG_SKIP=$dbg"; TB_DEBUG_SKIP=$dbg timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 $B --gpus 2 2>&1 | grep "value" | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print(d['value']/1e9, d['roofline']['avg_launch_us'])"
done
echo "== 1 GPU c4 189x189 slab (95 planes as a plain lattice would be size-equivalent): c4 --size 150"; python $B --workload c4 --size 150 | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print(d['value']/1e9, d['roofline']['avg_launch_us'])"
