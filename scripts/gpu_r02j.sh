#!/bin/bash
# final round-1 verification on the rebuilt tree + the low-degree straight line of the resident kernel (C1)
OUT=gpurun_out; mkdir -p $OUT
for i in 1 2 3; do timeout 120 titan_b200/bin/titan_facade_dump bench c1 20000 2>&1 | grep '^{' | cut -c1-260; done
timeout 120 titan_b200/bin/titan_facade_dump bench lattice10 10000 2>&1 | grep '^{' | cut -c1-260
bash scripts/gpu_round_end.sh r02j
