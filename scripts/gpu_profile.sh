#!/bin/bash
# Usage (on the GPU box, from the repo root): scripts/gpu_profile.sh TAG
# 1. plain bench (must exit 0), 2. ncu launch list, 3. one ncu --set full capture of the step kernel.
TAG=${1:-rXX}
OUT=gpurun_out
mkdir -p $OUT
B="python bench.py --steps 1 --warmup 3 --inner 60 --no-cpu-baseline --no-e2e"
$B > $OUT/${TAG}_plain.log 2> $OUT/${TAG}_plain.err || { echo "plain bench failed"; tail -5 $OUT/${TAG}_plain.err; exit 1; }
TB_PROFILE=1 $B > /dev/null 2> $OUT/${TAG}_phases.log
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/${TAG}_launches.csv $B > $OUT/${TAG}_ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:tiled_step -s 100 -c 2 -f -o $OUT/${TAG}_prof $B > $OUT/${TAG}_ncu2.log 2>&1
ls -la $OUT | tail -8
