#!/bin/bash
# ncu --set full of the kernel instantiations the bench lines other than C2 run: summary + source-level summary each
TAG=${1:-rv}; OUT=gpurun_out; mkdir -p $OUT
P="python bench.py --steps 1 --warmup 3 --inner 60 --no-cpu-baseline --no-e2e --no-extras"
cap() {  # name, kernel regex, skip, command...
  local name=$1 rx=$2 skip=$3; shift 3
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:$rx -s $skip -c 1 -f -o $OUT/${TAG}_${name}_prof "$@" > $OUT/${TAG}_ncu_${name}.log 2>&1; echo "ncu $name rc=$?"
  ncu -i $OUT/${TAG}_${name}_prof.ncu-rep --page source --csv > $OUT/${TAG}_${name}_source.csv 2>/dev/null
  python profiles/ncu_summary.py $OUT/${TAG}_${name}_prof.ncu-rep > $OUT/${TAG}_${name}_summary.txt 2>&1
  python profiles/ncu_source_summary.py $OUT/${TAG}_${name}_source.csv 10 > $OUT/${TAG}_${name}_source_summary.txt 2>&1
  rm -f $OUT/${TAG}_${name}_prof.ncu-rep $OUT/${TAG}_${name}_source.csv
  grep -E "^kernel|time_duration|bank_conflicts|wavefronts_mem_shared|registers|l1tex__throughput|dram__bytes|issue_active|warps_active|fp64" $OUT/${TAG}_${name}_summary.txt | cut -c1-150
}
cap distinct tiled_step 100 $P --distinct-params
cap c3 tiled_step 100 $P --workload c3
cap halo150 tiled_step 100 env TB_FORCE_HALO=1 $P --size 150
cap plain150 tiled_step 100 $P --size 150
cap csr csr_step 100 $P --size 60 --csr
cat > /tmp/res.py <<'PY'
import sys; sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
import scenes, titan_b200 as tb
sc = scenes.c1_readme(); eng = tb.Engine(0, tb.MATH_FAST); scenes.upload(eng, sc)
for i in range(3): eng.run(2000, 0.2 * i)
eng.close()
PY
cap resident resident_step 1 python /tmp/res.py
