#!/bin/bash
# final tree of round 2: round-end sequence (GPU suite, smoke, default bench line, reference arm), then the ncu evidence of
# the same bench command: launch list (kernel shares) and one --set full capture of the step kernel with its source summary
TAG=${1:-r4f}; OUT=gpurun_out; mkdir -p $OUT
bash scripts/gpu_r4c.sh $TAG
P="python bench.py --steps 1 --warmup 3 --inner 60 --no-cpu-baseline --no-e2e --no-extras"
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/${TAG}_launches.csv $P > $OUT/${TAG}_ncu_launches.log 2>&1; echo "ncu launches rc=$?"
python - $OUT/${TAG}_launches.csv <<'PY'
import csv, sys, collections
t = collections.defaultdict(float); n = collections.Counter()
rows = [r for r in csv.reader(l for l in open(sys.argv[1]) if l.startswith('"'))]
hdr = rows[0]; k = hdr.index('Kernel Name'); v = hdr.index('Metric Value')
for r in rows[1:]:
    name = r[k].split('(')[0][:60]; t[name] += float(r[v].replace(',', '')); n[name] += 1
tot = sum(t.values())
for name in sorted(t, key=t.get, reverse=True)[:6]: print('%-62s %4d launches %10.1f us %5.1f %%' % (name, n[name], t[name] / 1e3 if tot > 1e7 else t[name], 100 * t[name] / tot))
PY
timeout 400 ncu --set full --clock-control none --import-source on -k regex:tiled_step -s 100 -c 1 -f -o $OUT/${TAG}_c2_prof $P > $OUT/${TAG}_ncu_c2.log 2>&1; echo "ncu full rc=$?"
ncu -i $OUT/${TAG}_c2_prof.ncu-rep --page source --csv > $OUT/${TAG}_c2_source.csv 2>/dev/null
python profiles/ncu_summary.py $OUT/${TAG}_c2_prof.ncu-rep > $OUT/${TAG}_c2_summary.txt 2>&1
python profiles/ncu_source_summary.py $OUT/${TAG}_c2_source.csv 10 > $OUT/${TAG}_c2_source_summary.txt 2>&1
rm -f $OUT/${TAG}_c2_prof.ncu-rep $OUT/${TAG}_c2_source.csv
grep -E "^kernel|time_duration|bank_conflicts|wavefronts_mem_shared|registers|l1tex__throughput|dram__bytes|issue_active|warps_active|fp64|local" $OUT/${TAG}_c2_summary.txt | cut -c1-150
