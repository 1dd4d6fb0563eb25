#!/bin/bash
# what the driver runs at round end, in its order: GPU tests, smoke, reference arm, default bench
TAG=$1; OUT=gpurun_out; mkdir -p $OUT
timeout 900 python -m pytest tests -x -q -m gpu > $OUT/${TAG}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -2 $OUT/${TAG}_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $OUT/${TAG}_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 $OUT/${TAG}_smoke.log
timeout 600 python bench.py --impl reference --gpus 1 --steps 3 --warmup 1 > $OUT/${TAG}_bench_reference_arm.json 2> $OUT/${TAG}_ref.err; echo "reference rc=$?"
timeout 600 python bench.py --gpus 1 > $OUT/${TAG}_bench_default.json 2> $OUT/${TAG}_default.err; echo "default rc=$?"
cut -c1-400 $OUT/${TAG}_bench_reference_arm.json; python - "$OUT/${TAG}_bench_default.json" <<'PY'
import sys, json
d = json.loads(open(sys.argv[1]).read())
print(round(d['value']/1e9, 2), 'frac', round(d['roofline']['frac'], 3), 'e2e', round(d['e2e']['value']/1e9, 2), 'cpu', round(d['cpu_baseline']['value']/1e6, 1), d['cpu_baseline']['cores'], 'launches', d['gpu_launches'], d['clocks'])
PY
