#!/bin/bash
# Is the per-GPU rate of a multi-GPU box lower because of the box (power budget shared by busy GPUs) or because of the code?
# One GPU alone, then the same single-GPU bench on every GPU at once (no communication at all), with power / clocks sampled.
N=${1:-4}; TAG=${2:-box}; OUT=gpurun_out; mkdir -p $OUT
nvidia-smi --query-gpu=index,power.limit,power.max_limit,clocks.max.sm,clocks.max.mem --format=csv > $OUT/${TAG}_limits.csv; cat $OUT/${TAG}_limits.csv
B="python bench.py --no-titan-cuda --no-cpu-baseline --no-e2e --steps 5 --warmup 3"
nvidia-smi --query-gpu=index,clocks.sm,clocks.mem,power.draw,clocks_throttle_reasons.active --format=csv,noheader -lms 500 > $OUT/${TAG}_smi_alone.csv & SMI=$!
CUDA_VISIBLE_DEVICES=0 timeout 300 $B > $OUT/${TAG}_alone.json 2> $OUT/${TAG}_alone.err; echo "alone rc=$?"
kill $SMI
nvidia-smi --query-gpu=index,clocks.sm,clocks.mem,power.draw,clocks_throttle_reasons.active --format=csv,noheader -lms 500 > $OUT/${TAG}_smi_all.csv & SMI=$!
PIDS=""
for ((g=0; g<N; g++)); do CUDA_VISIBLE_DEVICES=$g timeout 300 $B > $OUT/${TAG}_all_$g.json 2> $OUT/${TAG}_all_$g.err & PIDS="$PIDS $!"; done
wait $PIDS     # the bench processes only: a bare job-number wait also waited for the sampler and ran into the call's limit (r3z)
sleep 1; kill $SMI
python - $OUT/${TAG}_alone.json $OUT/${TAG}_all_*.json <<'PY'
import sys, json
for f in sys.argv[1:]:
    try:
        d = json.loads([l for l in open(f).read().splitlines() if l.startswith('{')][-1])
        print(f, round(d['value']/1e9, 2), 'G  weak_base', (d.get('weak_base') or {}).get('value'), d['clocks'])
    except Exception as e: print(f, 'ERR', e)
PY
for f in $OUT/${TAG}_smi_alone.csv $OUT/${TAG}_smi_all.csv; do echo $f; sort -t, -k4 -n -r $f | head -4; done
