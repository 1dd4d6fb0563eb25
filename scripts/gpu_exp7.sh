#!/bin/bash
B="bench.py --steps 3 --warmup 3 --inner 100 --no-cpu-baseline --no-e2e"
run() { python $B "$@" | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print(round(d['value']/1e9,2), round(d['roofline']['frac'],4), round(d['roofline']['avg_launch_us'],1), d['config']['tiling']['threads_per_mass'], d['config']['tiling']['grid'], d['config']['tiling']['smem_bytes'])"; }
for tpm in 1 2 4; do
  echo "== TB_TPM=$tpm c2: $(TB_TPM=$tpm run --workload c2)"
  echo "== TB_TPM=$tpm c4-150: $(TB_TPM=$tpm run --workload c4 --size 150)"
done
echo "== c3: $(run --workload c3)"
echo "== c3 TPM1: $(TB_TPM=1 run --workload c3)"
