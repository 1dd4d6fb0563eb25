#!/bin/bash
# wave quantisation on C2: 8000 tiles over 444 persistent CTAs = 19 rounds; fewer, fuller tiles (8x4x4 bricks) = 18 rounds
fmt() { grep '^{' | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); t=d['config'].get('tiling',{}); print(round(d['value']/1e9,2), 'us', round(d['roofline']['avg_launch_us'],1), 'tiles', t.get('n_tiles'), 'grid', t.get('grid'), 'smem', t.get('smem_bytes'))"; }
B="timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-e2e"
echo "5x5x5*20 (default):        $($B 2>/dev/null | fmt)"
echo "8x4x4*10,5x5x5*4 (7850):   $(TB_BRICK_SPEC='8x4x4*10,5x5x5*4' $B 2>/dev/null | fmt)"
echo "8x4x4*5,5x5x5*12 (7925):   $(TB_BRICK_SPEC='8x4x4*5,5x5x5*12' $B 2>/dev/null | fmt)"
echo "5x5x5*12,8x4x4*5 order:    $(TB_BRICK_SPEC='5x5x5*12,8x4x4*5' $B 2>/dev/null | fmt)"
echo "8x4x4*12,4x4x8*1 (8125):   $(TB_BRICK_SPEC='8x4x4*12,4x4x8*1' $B 2>/dev/null | fmt)"
echo "default grid 400 (20 rounds): $(TB_GRID=400 $B 2>/dev/null | fmt)"
