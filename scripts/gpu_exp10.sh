#!/bin/bash
# where does the partitioned step lose time?  2 GPUs, C4 189^3, exchange pieces switched off (results invalid, timing only)
export PYTHONPATH=$PWD:$PWD/tests
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
run() { timeout 300 $TR --master-port $1 bench.py --gpus 2 --steps 3 --warmup 3 --no-e2e --no-cpu-baseline 2>/dev/null | grep '^{' | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print(round(d['value']/1e9,2), round(d['roofline']['avg_launch_us'],1))"; }
echo "full exchange: $(run 29601)"
echo "no export stores (8): $(TB_DEBUG_SKIP=8 run 29602)"
echo "no wait for peers (16): $(TB_DEBUG_SKIP=16 run 29603)"
echo "neither (24): $(TB_DEBUG_SKIP=24 run 29604)"
echo "full exchange again: $(run 29605)"
