#!/bin/bash
# A/B of the bank-aware record slots (TB_BANK_SLOTS=0 = first-occurrence order) and the inert HALO variant vs the plain kernel
TAG=${1:-rs}; OUT=gpurun_out; mkdir -p $OUT
timeout 1200 python -m pytest tests -x -q -m gpu > $OUT/${TAG}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -2 $OUT/${TAG}_pytest_gpu.log
B="python bench.py --no-cpu-baseline --no-e2e --no-extras"
run() { local name=$1; shift; timeout 600 "$@" > $OUT/${TAG}_$name.json 2>> $OUT/${TAG}_err.log; python - $OUT/${TAG}_$name.json <<'PY'
import sys, json
try:
    d = json.loads(open(sys.argv[1]).read()); print(sys.argv[1], round(d['value']/1e9, 2), 'frac', round(d['roofline']['frac'], 3), 'us', round(d['roofline']['avg_launch_us'], 2), (d.get('clocks') or {}).get('reasons'))
except Exception as e: print(sys.argv[1], 'ERR', e)
PY
}
run c2 $B
run c2_oldslots env TB_BANK_SLOTS=0 $B
run distinct $B --distinct-params
run distinct_oldslots env TB_BANK_SLOTS=0 $B --distinct-params
run plainrec env TB_DICT=0 $B
run plainrec_oldslots env TB_DICT=0 TB_BANK_SLOTS=0 $B
run c3 $B --workload c3
run c5 python bench.py --workload c5 --steps 10
run c5_oldslots env TB_BANK_SLOTS=0 python bench.py --workload c5 --steps 10
run plain150 $B --size 150
run halo150 env TB_FORCE_HALO=1 $B --size 150
