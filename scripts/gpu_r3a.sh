#!/bin/bash
# round-3 baseline on today's box: GPU suite, bench lines of the workloads the verdict names, source-level ncu of the step kernel
TAG=${1:-r3a}; OUT=gpurun_out; mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.max.sm,power.limit --format=csv > $OUT/${TAG}_gpu.txt
timeout 300 python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu > $OUT/${TAG}_pytest_zz_alone.log 2>&1; echo "zz alone rc=$?"; tail -1 $OUT/${TAG}_pytest_zz_alone.log
python -m pytest tests -q -m gpu --collect-only 2>/dev/null | grep "::" | tail -2
TB_PARITY_LOG=$OUT/${TAG}_parity_errors.jsonl timeout 1200 python -m pytest tests -q -m gpu > $OUT/${TAG}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -2 $OUT/${TAG}_pytest_gpu.log
B="python bench.py --no-cpu-baseline --no-e2e"
timeout 300 $B > $OUT/${TAG}_c2.json 2> $OUT/${TAG}_c2.err; echo "c2 rc=$?"
timeout 300 $B --distinct-params > $OUT/${TAG}_c2_distinct.json 2>> $OUT/${TAG}_c2.err; echo "distinct rc=$?"
timeout 300 $B --size 150 > $OUT/${TAG}_c2_150.json 2>> $OUT/${TAG}_c2.err; echo "150 rc=$?"
timeout 300 $B --workload c3 > $OUT/${TAG}_c3.json 2>> $OUT/${TAG}_c2.err; echo "c3 rc=$?"
timeout 600 python bench.py --workload c5 --steps 10 > $OUT/${TAG}_c5.json 2>> $OUT/${TAG}_c2.err; echo "c5 rc=$?"
for f in c2 c2_distinct c2_150 c3 c5; do python - $OUT/${TAG}_$f.json <<'PY'
import sys, json
try:
    d = json.loads(open(sys.argv[1]).read()); print(sys.argv[1], round(d['value']/1e9, 2), 'frac', round(d['roofline']['frac'], 3), d.get('clocks'))
except Exception as e: print(sys.argv[1], 'ERR', e)
PY
done
P="python bench.py --steps 1 --warmup 3 --inner 60 --no-cpu-baseline --no-e2e"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:tiled_step -s 100 -c 1 -f -o $OUT/${TAG}_c2_prof $P > $OUT/${TAG}_ncu_c2.log 2>&1; echo "ncu c2 rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:tiled_step -s 100 -c 1 -f -o $OUT/${TAG}_c2d_prof $P --distinct-params > $OUT/${TAG}_ncu_c2d.log 2>&1; echo "ncu c2d rc=$?"
for r in c2 c2d; do ncu -i $OUT/${TAG}_${r}_prof.ncu-rep --page source --csv > $OUT/${TAG}_${r}_source.csv 2>/dev/null; python profiles/ncu_summary.py $OUT/${TAG}_${r}_prof.ncu-rep > $OUT/${TAG}_${r}_summary.txt 2>&1; ncu -i $OUT/${TAG}_${r}_prof.ncu-rep --page raw --csv > $OUT/${TAG}_${r}_raw.csv 2>/dev/null; rm -f $OUT/${TAG}_${r}_prof.ncu-rep; done
ls -la $OUT | tail -20
timeout 600 python scripts/drift_table.py $OUT/${TAG}_drift.json > $OUT/${TAG}_drift.log 2>&1; echo "drift rc=$?"; tail -3 $OUT/${TAG}_drift.log
timeout 200 tests/cpp/bin/titan_facade_dump bench_e2e c2 200 5 > $OUT/${TAG}_facade_e2e.log 2>&1; echo "facade e2e rc=$?"; tail -1 $OUT/${TAG}_facade_e2e.log
timeout 200 tests/cpp/bin/titan_facade_dump bench c1 10000 >> $OUT/${TAG}_facade_e2e.log 2>&1; tail -1 $OUT/${TAG}_facade_e2e.log
timeout 300 oracle/_ref/titan_ref_cuda bench c2 200 > $OUT/${TAG}_titan_cuda.log 2>&1; echo "titan cuda rc=$?"; tail -1 $OUT/${TAG}_titan_cuda.log
