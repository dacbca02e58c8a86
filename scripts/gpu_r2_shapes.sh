#!/bin/bash
# 1 GPU: the fused T = 3 kernel on the tile shapes an 8-GPU strong-scaling run of 16384^2 can use
# (periodic self-neighbours: same code path as a rank with real neighbours)
mkdir -p gpurun_out
: > gpurun_out/tile_shapes.txt
for S in "8192 4096" "4096 8192" "2048 16384" "16384 2048" "16384 16384"; do
  set -- $S
  timeout 120 python bench.py --steps 60 --warmup 5 --th $1 --tw $2 --no-baseline --min-seconds 0.6 > gpurun_out/shape.json 2> gpurun_out/shape.err
  python - "$1" "$2" <<'PY' | tee -a gpurun_out/tile_shapes.txt
import json, sys
d=None
for l in open("gpurun_out/shape.json"):
    if l.startswith('{"metric"'): d=json.loads(l)
if d is None:
    print(sys.argv[1], sys.argv[2], "FAILED", open("gpurun_out/shape.err").read()[-500:])
else:
    print(f"tile {sys.argv[1]:>6} x {sys.argv[2]:>6}: {d['value']:8.1f} Gcell/s  {d['ms_per_step']*1e3:7.2f} us/step  e2e {d['e2e']['value']:8.1f}  "
          f"{d['config'].get('kernel','')[:40]}  {d['clocks']['sm_mhz']} MHz {d['clocks']['reasons']}")
PY
done
