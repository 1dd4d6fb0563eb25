#!/bin/bash
# kernel iteration: GPU suite (-x), bench lines of C2 / distinct-k / 150^3 / C3 / C5, one source-level ncu capture of the C2 step kernel
TAG=${1:-rk}; OUT=gpurun_out; mkdir -p $OUT
timeout 1200 python -m pytest tests -x -q -m gpu > $OUT/${TAG}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -2 $OUT/${TAG}_pytest_gpu.log
B="python bench.py --no-cpu-baseline --no-e2e --no-extras"
timeout 300 $B > $OUT/${TAG}_c2.json 2> $OUT/${TAG}_err.log; echo "c2 rc=$?"
timeout 300 $B --distinct-params > $OUT/${TAG}_c2_distinct.json 2>> $OUT/${TAG}_err.log; echo "distinct rc=$?"
timeout 300 $B --size 150 > $OUT/${TAG}_c2_150.json 2>> $OUT/${TAG}_err.log; echo "150 rc=$?"
timeout 300 $B --workload c3 > $OUT/${TAG}_c3.json 2>> $OUT/${TAG}_err.log; echo "c3 rc=$?"
timeout 600 python bench.py --workload c5 --steps 10 > $OUT/${TAG}_c5.json 2>> $OUT/${TAG}_err.log; echo "c5 rc=$?"
for f in c2 c2_distinct c2_150 c3 c5; do python - $OUT/${TAG}_$f.json <<'PY'
import sys, json
try:
    d = json.loads(open(sys.argv[1]).read()); print(sys.argv[1], round(d['value']/1e9, 2), 'frac', round(d['roofline']['frac'], 3), (d.get('clocks') or {}).get('reasons'), d['config'].get('tiling', {}).get('smem_bytes'))
except Exception as e: print(sys.argv[1], 'ERR', e)
PY
done
if [ "$2" != "noncu" ]; then
P="python bench.py --steps 1 --warmup 3 --inner 60 --no-cpu-baseline --no-e2e --no-extras"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:tiled_step -s 100 -c 1 -f -o $OUT/${TAG}_c2_prof $P > $OUT/${TAG}_ncu_c2.log 2>&1; echo "ncu c2 rc=$?"
ncu -i $OUT/${TAG}_c2_prof.ncu-rep --page source --csv > $OUT/${TAG}_c2_source.csv 2>/dev/null
python profiles/ncu_summary.py $OUT/${TAG}_c2_prof.ncu-rep > $OUT/${TAG}_c2_summary.txt 2>&1
python profiles/ncu_source_summary.py $OUT/${TAG}_c2_source.csv 10 > $OUT/${TAG}_c2_source_summary.txt 2>&1
rm -f $OUT/${TAG}_c2_prof.ncu-rep
grep -E "time_duration|bank_conflicts|wavefronts_mem_shared|registers|l1tex__throughput|dram__bytes|issue_active" $OUT/${TAG}_c2_summary.txt; head -3 $OUT/${TAG}_c2_source_summary.txt
fi
