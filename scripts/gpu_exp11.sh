#!/bin/bash
# thin slabs as on 8 GPUs (38 planes of 300x300 per rank), on 2 GPUs and on 1 GPU
export PYTHONPATH=$PWD:$PWD/tests
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
fmt() { grep '^{' | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print(round(d['value']/1e9,2), round(d['roofline']['avg_launch_us'],1), d['config']['tiling']['n_tiles'])"; }
run2() { timeout 300 $TR --master-port $1 bench.py --gpus 2 --workload c4 --shape 76,300,300 --steps 3 --warmup 3 --no-e2e --no-cpu-baseline 2>/dev/null | fmt; }
run1() { timeout 300 python bench.py --workload c4 --shape 38,300,300 --steps 3 --warmup 3 --no-e2e --no-cpu-baseline 2>/dev/null | fmt; }
echo "1 GPU 38x300x300: $(run1)"
echo "1 GPU 38x300x300 FORCE_HALO: $(TB_FORCE_HALO=1 run1)"
echo "2 GPU 76x300x300 full exchange: $(run2 29601)"
echo "2 GPU no export stores (8): $(TB_DEBUG_SKIP=8 run2 29602)"
echo "2 GPU neither (24): $(TB_DEBUG_SKIP=24 run2 29604)"
