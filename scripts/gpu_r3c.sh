#!/bin/bash
# default bench line with its extras, the reference arm, the facade tests
TAG=${1:-r3c}; OUT=gpurun_out; mkdir -p $OUT
timeout 600 python -m pytest tests/test_facade.py tests/test_zz_no_alloc.py -x -q -m gpu > $OUT/${TAG}_pytest_facade.log 2>&1; echo "facade tests rc=$?"; tail -2 $OUT/${TAG}_pytest_facade.log
timeout 300 tests/cpp/bin/titan_facade_dump bench_e2e c2 200 5 > $OUT/${TAG}_facade_e2e.log 2>&1; echo "facade e2e rc=$?"; grep "{" $OUT/${TAG}_facade_e2e.log | cut -c1-260
timeout 900 python bench.py > $OUT/${TAG}_bench_default.json 2> $OUT/${TAG}_default.err; echo "default rc=$?"; tail -3 $OUT/${TAG}_default.err
timeout 600 python bench.py --impl reference --gpus 1 --steps 3 --warmup 1 > $OUT/${TAG}_bench_reference_arm.json 2> $OUT/${TAG}_ref.err; echo "reference rc=$?"
python - $OUT/${TAG}_bench_default.json <<'PY'
import sys, json
d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print('value', round(d['value']/1e9, 2), 'frac', round(d['roofline']['frac'], 3), 'distinct', d['roofline'].get('distinct_params'))
print('e2e', d['e2e']); print('e2e_cabi', d.get('e2e_cabi')); print('weak_base', d.get('weak_base')); print('titan', d.get('titan_cuda_baseline'))
print('cpu', d['cpu_baseline']); print(d['clocks'])
PY
