#!/bin/bash
# round-end sequence on the final tree: GPU suite, smoke, default bench line (timed), reference arm (timed)
TAG=${1:-r4c}; OUT=gpurun_out; mkdir -p $OUT
S=$SECONDS; timeout 900 python -m pytest tests -x -q -m gpu > $OUT/${TAG}_pytest_gpu.log 2>&1; echo "pytest rc=$? ($((SECONDS-S)) s)"; tail -2 $OUT/${TAG}_pytest_gpu.log
S=$SECONDS; timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2; echo "smoke ($((SECONDS-S)) s)"
S=$SECONDS; timeout 900 python bench.py > $OUT/${TAG}_bench_default.json 2> $OUT/${TAG}_default.err; echo "default rc=$? ($((SECONDS-S)) s)"; tail -3 $OUT/${TAG}_default.err
S=$SECONDS; timeout 600 python bench.py --impl reference > $OUT/${TAG}_bench_reference_arm.json 2> $OUT/${TAG}_ref.err; echo "reference rc=$? ($((SECONDS-S)) s)"
python - $OUT/${TAG}_bench_default.json $OUT/${TAG}_bench_reference_arm.json <<'PY'
import sys, json
d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print('value', round(d['value']/1e9, 2), 'frac', round(d['roofline']['frac'], 3), 'distinct', d['roofline'].get('distinct_params'))
print('e2e', d['e2e']); print('e2e_cabi', d.get('e2e_cabi')); print('weak_base', d.get('weak_base')); print('titan', d.get('titan_cuda_baseline'))
print('cpu', d['cpu_baseline']); print(d['clocks'], 'launches', d['gpu_launches'])
r = json.loads(open(sys.argv[2]).read().strip().splitlines()[-1])
print('reference arm', r.get('value'), r.get('cpu_baseline'), r.get('steps'))
PY
