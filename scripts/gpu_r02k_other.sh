#!/bin/bash
# the other single-GPU workloads on the final tree
OUT=gpurun_out; mkdir -p $OUT
for w in c3 c5; do
  timeout 600 python bench.py --workload $w --steps 5 --warmup 3 --no-cpu-baseline > $OUT/r02k_bench_$w.json 2> $OUT/r02k_bench_$w.err; echo "$w rc=$?"
  python - $OUT/r02k_bench_$w.json <<'PY'
import sys, json
d = json.loads(open(sys.argv[1]).read())
print(round(d['value']/1e9, 2), 'frac', round(d['roofline']['frac'], 3), 'e2e', round(d['e2e']['value']/1e9, 2) if d.get('e2e') else None, d['config']['workload'][:60])
PY
done
timeout 120 titan_b200/bin/titan_facade_dump bench c2 2000 2>&1 | grep '^{' | cut -c1-260
