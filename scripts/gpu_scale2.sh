#!/bin/bash
# scripts/gpu_scale2.sh TAG N : parity worker + default bench (c4: one lattice partitioned over N GPUs) at N GPUs
TAG=$1; N=$2
OUT=gpurun_out; mkdir -p $OUT
export PYTHONPATH=$PWD:$PWD/tests
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
TB_TEST_NX=93 timeout 300 $TR --master-port 29512 tests/dist_gpu_worker.py > $OUT/${TAG}_parity_n$N.log 2>&1; echo "parity rc=$?"; grep "world" $OUT/${TAG}_parity_n$N.log
timeout 900 $TR --master-port 29513 bench.py --gpus $N --steps 5 --warmup 3 > $OUT/${TAG}_bench_c4_n$N.log 2>&1; echo "c4 rc=$?"; grep '^{' $OUT/${TAG}_bench_c4_n$N.log | cut -c1-900
