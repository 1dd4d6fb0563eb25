#!/bin/bash
# 2 GPUs: GPU suite (the CUDA-IPC partition test runs here), bench.py --gpus 2 (parity_n before timing), reference arm at N = 2
TAG=${1:-rn2}; OUT=gpurun_out; mkdir -p $OUT
timeout 1200 python -m pytest tests -x -q -m gpu > $OUT/${TAG}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -2 $OUT/${TAG}_pytest_gpu.log
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533"
timeout 900 $T bench.py --gpus 2 --steps 10 --warmup 3 > $OUT/${TAG}_bench_c4_n2.json 2> $OUT/${TAG}_n2.err; echo "bench n2 rc=$?"; tail -2 $OUT/${TAG}_n2.err
timeout 900 $T bench.py --impl reference --gpus 2 --steps 2 --warmup 1 > $OUT/${TAG}_bench_reference_n2.json 2> $OUT/${TAG}_ref2.err; echo "ref n2 rc=$?"
python - $OUT/${TAG}_bench_c4_n2.json $OUT/${TAG}_bench_reference_n2.json <<'PY'
import sys, json
for f in sys.argv[1:]:
    try:
        d = json.loads([l for l in open(f).read().splitlines() if l.startswith('{')][-1])
        print(f, round(d['value']/1e9, 3), 'G', d.get('parity_n'), (d.get('cpu_baseline') or {}).get('cores'), (d.get('roofline') or {}).get('frac'), (d.get('e2e') or {}).get('value'))
    except Exception as e: print(f, 'ERR', e)
PY
