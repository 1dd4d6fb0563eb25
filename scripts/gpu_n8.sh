#!/bin/bash
# 8 GPUs (charged 8x: keep it short): multi-GPU tests, bench.py --gpus 8 on C4 300^3 (parity_n before timing) and on C3,
# one lattice over 8 GPUs through titan::Simulation(std::vector<int>) (tests/cpp/titan_api_driver.cu bench_multi).
TAG=${1:-rn8}; OUT=gpurun_out; mkdir -p $OUT
nvidia-smi --query-gpu=index,name,memory.total --format=csv,noheader | head -8
timeout 300 python -m pytest tests/test_group.py tests/test_partitioned.py tests/test_facade.py -x -q -m gpu > $OUT/${TAG}_pytest_multi.log 2>&1; echo "pytest rc=$?"; tail -2 $OUT/${TAG}_pytest_multi.log
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29533"
S=$SECONDS; timeout 600 $T bench.py --gpus 8 --steps 5 --warmup 3 > $OUT/${TAG}_bench_c4_n8.json 2> $OUT/${TAG}_c4.err; echo "bench c4 n8 rc=$? wall $((SECONDS-S)) s"; tail -3 $OUT/${TAG}_c4.err
timeout 300 $T bench.py --gpus 8 --workload c3 --steps 5 --warmup 3 > $OUT/${TAG}_bench_c3_n8.json 2> $OUT/${TAG}_c3.err; echo "bench c3 n8 rc=$?"; tail -3 $OUT/${TAG}_c3.err
timeout 400 tests/cpp/bin/titan_facade_dump bench_multi lattice240 300 8 > $OUT/${TAG}_facade_multi_240_n8.json 2> $OUT/${TAG}_facade.err; echo "facade multi rc=$?"; tail -3 $OUT/${TAG}_facade.err; cat $OUT/${TAG}_facade_multi_240_n8.json
python - $OUT/${TAG}_bench_c4_n8.json $OUT/${TAG}_bench_c3_n8.json <<'PY'
import sys, json
for f in sys.argv[1:]:
    try:
        d = json.loads([l for l in open(f).read().splitlines() if l.startswith('{')][-1])
        print(f, round(d['value']/1e9, 3), 'G', d.get('parity_n'), (d.get('roofline') or {}).get('frac'), (d.get('e2e') or {}).get('value'), d['config'].get('wall_ms'))
    except Exception as e: print(f, 'ERR', e)
PY
