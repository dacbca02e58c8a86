#!/bin/bash
# phase timing on ONE GPU for the tile shapes of strong scaling (periodic self-neighbours: same code path)
mkdir -p gpurun_out
export B2C_STENCIL_DEBUG=32
for hw in "16384 16384" "16384 8192" "8192 8192" "8192 4096" "4096 4096"; do
  set -- $hw
  python bench/stencil_phases.py --th $1 --tw $2 2>/dev/null | grep '^{' | python -c "
import json,sys
d=json.loads(sys.stdin.read()); m=d['median_over_sweeps']
cells=d['tile_per_gpu'][0]*d['tile_per_gpu'][1]
print(d['tile_per_gpu'], m, 'Gcell/s at span', round(3*cells/m['grid_span_us']/1e3,1))"
done
for w in 1368 1026 684; do
  echo "B2C_TS_WIDTH=$w tile 8192x4096"
  B2C_TS_WIDTH=$w python bench/stencil_phases.py --th 8192 --tw 4096 2>/dev/null | grep '^{' | python -c "
import json,sys
d=json.loads(sys.stdin.read()); m=d['median_over_sweeps']
cells=d['tile_per_gpu'][0]*d['tile_per_gpu'][1]
print(m, 'Gcell/s at span', round(3*cells/m['grid_span_us']/1e3,1))"
done
