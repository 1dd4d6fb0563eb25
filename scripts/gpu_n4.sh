#!/bin/bash
# 4 GPUs: bench.py --gpus 4 on C4 (238^3), then the same with the peer stores / the peer wait switched off (timing experiments,
# wrong results: --no-parity), and one lattice over 4 GPUs through titan::Simulation (bench_multi).
TAG=${1:-rn4}; OUT=gpurun_out; mkdir -p $OUT
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29533"
timeout 400 $T bench.py --gpus 4 --steps 5 --warmup 3 > $OUT/${TAG}_bench_c4_n4.json 2> $OUT/${TAG}_c4.err; echo "bench c4 n4 rc=$?"
for d in 8 24; do
TB_DEBUG_SKIP=$d timeout 400 $T bench.py --gpus 4 --steps 5 --warmup 3 --no-parity --no-e2e > $OUT/${TAG}_bench_c4_n4_skip$d.json 2> $OUT/${TAG}_c4_skip$d.err; echo "bench c4 n4 skip $d rc=$?"
done
timeout 400 tests/cpp/bin/titan_facade_dump bench_multi lattice240 300 4 > $OUT/${TAG}_facade_multi_240_n4.json 2> $OUT/${TAG}_facade.err; echo "facade multi rc=$?"; tail -3 $OUT/${TAG}_facade.err; cat $OUT/${TAG}_facade_multi_240_n4.json
timeout 200 tests/cpp/bin/titan_facade_tests multi 1 | tail -5
python - $OUT/${TAG}_bench_c4_n4*.json <<'PY'
import sys, json
for f in sys.argv[1:]:
    try:
        d = json.loads([l for l in open(f).read().splitlines() if l.startswith('{')][-1])
        print(f, round(d['value']/1e9, 3), 'G', d.get('parity_n'), (d.get('roofline') or {}).get('frac'), (d.get('e2e') or {}).get('value'), d['clocks'])
    except Exception as e: print(f, 'ERR', e)
PY
