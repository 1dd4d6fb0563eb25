#!/bin/bash
# magnet pass: parity subset, per-step time on a 60^3 lattice (tiled / generic), ncu launch list + one full capture
OUT=gpurun_out; mkdir -p $OUT
timeout 600 python -m pytest tests -q -m gpu 2>&1 | tail -4
timeout 300 python scripts/gpu_magnet_scale.py 60 2>&1 | tee $OUT/r02k_magnet_scale.jsonl | cut -c1-200
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $OUT/r02k_magnet_launches.csv python scripts/gpu_magnet_prof.py 60 > $OUT/ncu.log 2>&1
timeout 500 ncu --set full --clock-control none --import-source on -k regex:magnet_force -s 2 -c 1 -f -o $OUT/r02k_magnet python scripts/gpu_magnet_prof.py 60 > $OUT/ncu2.log 2>&1; tail -2 $OUT/ncu2.log
