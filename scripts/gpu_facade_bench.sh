#!/bin/bash
# scripts/gpu_facade_bench.sh TAG : the same driver source (tests/cpp/titan_api_driver.cu `bench c2`) against Titan's own CUDA
# engine and against this repo's drop-in facade, on the same B200; wall clock around Simulation::pause
TAG=$1; OUT=gpurun_out; mkdir -p $OUT
timeout 300 tests/cpp/bin/titan_facade_dump bench c2 1000 > $OUT/${TAG}_facade_bench_c2.log 2>&1; echo "facade rc=$?"
timeout 300 tests/cpp/bin/titan_facade_dump bench lattice37 1000 >> $OUT/${TAG}_facade_bench_c2.log 2>&1
timeout 600 oracle/_ref/titan_ref_cuda bench c2 20 > $OUT/${TAG}_titan_cuda_bench_c2.log 2>&1; echo "titan cuda rc=$?"
for c in "c1 10000" "c3_64 1000"; do
  timeout 300 tests/cpp/bin/titan_facade_dump bench $c >> $OUT/${TAG}_facade_bench_c2.log 2>&1
  timeout 600 oracle/_ref/titan_ref_cuda bench $c >> $OUT/${TAG}_titan_cuda_bench_c2.log 2>&1
done
grep '^{' $OUT/${TAG}_facade_bench_c2.log $OUT/${TAG}_titan_cuda_bench_c2.log
timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.read()); print(round(d['value'] / 1e9, 2), d['clocks'])"
