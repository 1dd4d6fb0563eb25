#!/bin/bash
# scripts/gpu_final.sh TAG : the record run -- default bench (both arms), ncu launch list + full capture, other workloads
TAG=$1; OUT=gpurun_out; mkdir -p $OUT
python bench.py --impl reference --steps 3 --warmup 1 > $OUT/${TAG}_bench_reference_arm.json 2> $OUT/${TAG}_ref.err; echo "reference rc=$?"
python bench.py > $OUT/${TAG}_bench_default.json 2> $OUT/${TAG}_default.err; echo "default rc=$?"
python bench.py --distinct-params --no-cpu-baseline > $OUT/${TAG}_bench_c2_distinct_params.json 2>/dev/null; echo "distinct rc=$?"
python bench.py --workload c4 --size 150 --no-cpu-baseline > $OUT/${TAG}_bench_150.json 2>/dev/null; echo "150 rc=$?"
python bench.py --workload c3 --no-cpu-baseline > $OUT/${TAG}_bench_c3.json 2>/dev/null; echo "c3 rc=$?"
python bench.py --workload c5 --no-cpu-baseline --steps 5 > $OUT/${TAG}_bench_c5.json 2>/dev/null; echo "c5 rc=$?"
bash scripts/gpu_profile.sh $TAG > /dev/null 2>&1
TB_PROFILE=2 python bench.py --steps 1 --warmup 3 --inner 60 --no-cpu-baseline --no-e2e > /dev/null 2> $OUT/${TAG}_phases_lite.log
for f in $OUT/${TAG}_bench_*.json; do python - "$f" <<'PY'
import sys, json
for l in open(sys.argv[1]):
    if l.startswith('{'):
        d = json.loads(l); e = d.get('e2e') or {}; c = d.get('cpu_baseline') or {}
        print(sys.argv[1].split('/')[-1], round(d['value'] / 1e9, 2), 'frac', d.get('roofline', {}).get('frac'), 'e2e', e.get('value') and round(e['value'] / 1e9, 2), 'cpu', c.get('value') and round(c['value'] / 1e6, 1))
PY
done
