#!/bin/bash
# scripts/gpu_c3_multi.sh TAG N : BASELINE config C3 (4096 independent actuated robots) sharded over N GPUs, no communication
TAG=$1; N=$2; OUT=gpurun_out; mkdir -p $OUT
export PYTHONPATH=$PWD:$PWD/tests
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29514 bench.py --gpus $N --steps 5 --warmup 3 --workload c3 > $OUT/${TAG}_bench_c3_n$N.log 2>&1; echo "c3 rc=$?"; grep '^{' $OUT/${TAG}_bench_c3_n$N.log | cut -c1-500
