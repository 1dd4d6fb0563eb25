#!/usr/bin/env bash
# r01w: golden dumps of the REFERENCE's own CUDA engine for the magnets scene and for its VERLET build, then the GPU
# test-suite and the default bench line.  Run with: gpurun --timeout 1500 -- bash scripts/gpu_r01w.sh
set -u
mkdir -p gpurun_out/golden
G=gpurun_out/golden
timeout 120 oracle/_ref/titan_ref_cuda dump magnets $G/ref_magnets.bin > gpurun_out/r01w_ref_dump.log 2>&1
timeout 120 oracle/_ref/titan_ref_cuda dump magnets $G/ref_magnets_run2.bin >> gpurun_out/r01w_ref_dump.log 2>&1
for s in c1_readme contact_friction actuated local_constraints; do
  timeout 120 oracle/_ref/titan_ref_cuda_verlet dump $s $G/ref_verlet_$s.bin >> gpurun_out/r01w_ref_dump.log 2>&1
done
ls -la $G >> gpurun_out/r01w_ref_dump.log
# goldens produced by an earlier call are already in tests/golden/ref_cuda; otherwise use the fresh ones
for f in $G/ref_magnets.bin $G/ref_verlet_*.bin; do
  [ -f tests/golden/ref_cuda/$(basename $f) ] || cp $f tests/golden/ref_cuda/
done
timeout 1200 python -m pytest tests -m gpu -q -x --timeout 600 > gpurun_out/r01w_pytest_gpu.log 2>&1
echo "pytest exit $?" >> gpurun_out/r01w_pytest_gpu.log
timeout 120 python -m pytest tests -q -m "not gpu" -k "reference_cuda or verlet or builders" > gpurun_out/r01w_pytest_cpu.log 2>&1
timeout 600 python bench.py > gpurun_out/r01w_bench_default.json 2> gpurun_out/r01w_bench_default.err
tail -3 gpurun_out/r01w_pytest_gpu.log; tail -3 gpurun_out/r01w_pytest_cpu.log; cat gpurun_out/r01w_bench_default.json
