#!/bin/bash
# A/B of engine builds on ONE box: scripts/gpu_ab.sh "<bench args>" lib1.so lib2.so ...   (each also with TB_FORCE_HALO)
ARGS="$1"; shift
B="bench.py --steps 3 --warmup 3 --inner 100 --no-cpu-baseline --no-e2e $ARGS"
run() { python $B | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print(round(d['value']/1e9,2), round(d['roofline']['avg_launch_us'],1))"; }
for rep in 1 2; do
for lib in "$@"; do
  export TITAN_B200_LIB=$PWD/$lib
  echo "== $lib plain: $(run)"
  echo "== $lib FORCE_HALO: $(TB_FORCE_HALO=1 run)"
done; done
