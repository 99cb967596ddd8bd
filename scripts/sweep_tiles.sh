#!/bin/bash
# Sweeps tile length / warm-up of the tiled kernels on the C2 bench workload (one short bench run per point).
out=gpurun_out/sweep_tiles.log
: > $out
for L in 400 500 640 800 1000 1500; do
  for W in 64 128; do
    echo "L=$L W=$W" >> $out
    MARGIN_B200_TILE_LEN=$L MARGIN_B200_WARMUP=$W python bench.py --steps 5 --warmup 3 --no-cpu-baseline 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print(' ms=%.3f gcups=%.2f e2e_ms=%.3f k=%s'%(d['ms_per_step'],d['value'],d['e2e']['ms_per_step'],{k:round(v,3) for k,v in d['kernel_ms_per_step'].items()}))
" >> $out
  done
done
MARGIN_B200_TRACE=1 python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/trace.log 2>&1
