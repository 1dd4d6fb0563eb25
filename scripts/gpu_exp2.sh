#!/bin/bash
OUT=gpurun_out; mkdir -p $OUT
B="bench.py --steps 2 --warmup 3 --inner 100 --no-cpu-baseline --no-e2e"
echo "== 1 GPU c4 150"; TB_PROFILE=1 python $B --workload c4 --size 150 2>&1 | grep "tb profile"
echo "== 2 GPU c4 189"; TB_PROFILE=1 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 $B --gpus 2 2>&1 | grep "tb profile\|value" | cut -c1-600
