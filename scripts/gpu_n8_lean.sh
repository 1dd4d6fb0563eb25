#!/bin/bash
# 8 GPUs, bench lines only (charged 8x): C4 300^3 with parity_n, then C3
TAG=${1:-rn8}; OUT=gpurun_out; mkdir -p $OUT
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29533"
S=$SECONDS; timeout 400 $T bench.py --gpus 8 --steps 5 --warmup 3 > $OUT/${TAG}_bench_c4_n8.json 2> $OUT/${TAG}_c4.err; echo "bench c4 n8 rc=$? wall $((SECONDS-S)) s"
S=$SECONDS; timeout 200 $T bench.py --gpus 8 --workload c3 --steps 5 --warmup 3 --no-e2e > $OUT/${TAG}_bench_c3_n8.json 2> $OUT/${TAG}_c3.err; echo "bench c3 n8 rc=$? wall $((SECONDS-S)) s"
python - $OUT/${TAG}_bench_c4_n8.json $OUT/${TAG}_bench_c3_n8.json <<'PY'
import sys, json
for f in sys.argv[1:]:
    try:
        d = json.loads([l for l in open(f).read().splitlines() if l.startswith('{')][-1])
        print(f, round(d['value']/1e9, 3), 'G', d.get('parity_n'), (d.get('roofline') or {}).get('frac'), (d.get('e2e') or {}).get('value'), d.get('clocks'))
    except Exception as e: print(f, 'ERR', e)
PY
