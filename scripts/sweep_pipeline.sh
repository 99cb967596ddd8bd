#!/bin/bash
# e2e (host buffers in, host lists out) of the C2 bench workload against the number of pipelined runs K and the
# SM share the runs' tiles are cut for.
out=gpurun_out/sweep_pipeline.log
: > $out
for cfg in "1 1 128" "3 1 128" "3 2 128" "3 3 128" "4 2 128" "4 4 128" "6 3 128" "6 6 128" "4 4 64" "6 6 64" "8 8 64"; do
  set -- $cfg
  echo "K=$1 FILL=$2 W=$3" >> $out
  MARGIN_B200_PIPELINE=$1 MARGIN_B200_PIPE_FILL=$2 MARGIN_B200_WARMUP=$3 python bench.py --steps 3 --warmup 3 --no-cpu-baseline 2>gpurun_out/pipe.err | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print(' ms=%.3f gcups=%.2f e2e_ms=%.3f e2e_gcups=%.2f'%(d['ms_per_step'],d['value'],d['e2e']['ms_per_step'],d['e2e']['value']))
" >> $out
done
