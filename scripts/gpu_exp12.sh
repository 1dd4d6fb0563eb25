#!/bin/bash
# layered brick shapes (lattice.brick_layers) vs one shape, on the slabs an 8-GPU rank of the 300^3 lattice owns, 1 GPU
fmt() { grep '^{' | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print(round(d['value']/1e9,2), round(d['roofline']['avg_launch_us'],1), d['config']['tiling']['n_tiles'], d['config']['tiling'].get('ell_unique_tiles'))"; }
run1() { timeout 300 python bench.py --workload c4 --shape $1 --steps 3 --warmup 3 --no-e2e --no-cpu-baseline 2>/dev/null | fmt; }
for shp in 38,300,300 37,300,300 94,189,189; do
  echo "$shp one shape:  $(TB_BRICK_LAYERS=0 run1 $shp)"
  echo "$shp layered:    $(run1 $shp)"
  echo "$shp layered FORCE_HALO: $(TB_FORCE_HALO=1 run1 $shp)"
done
echo "c2 default: $(timeout 300 python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu-baseline 2>/dev/null | fmt)"
