#!/bin/bash
# Usage (GPU box with N GPUs): scripts/gpu_multi.sh TAG N [extra bench args]   -- parity worker, then bench at N GPUs
TAG=${1:-rXX}; N=${2:-2}; shift 2
OUT=gpurun_out; mkdir -p $OUT
export PYTHONPATH=$PWD:$PWD/tests
nvidia-smi topo -m > $OUT/${TAG}_topo.txt 2>&1
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 tests/dist_gpu_worker.py > $OUT/${TAG}_parity_n$N.log 2>&1
echo "parity rc=$?"; tail -4 $OUT/${TAG}_parity_n$N.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus $N --steps 5 --warmup 3 "$@" > $OUT/${TAG}_bench_n$N.log 2>&1
echo "bench rc=$?"; tail -3 $OUT/${TAG}_bench_n$N.log | cut -c1-1500
