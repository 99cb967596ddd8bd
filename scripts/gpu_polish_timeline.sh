#!/bin/bash
# timestamped `-a info` log of margin polish on the synthetic 2 Mb contig (all host cores, all visible GPUs)
cd /root/repo; mkdir -p gpurun_out/tl
python - <<'PY' > gpurun_out/tl/dataset.log 2>&1
import subprocess, sys
sys.path.insert(0, '/root/repo')
from margin_b200 import synth
fa, sam, n, L = synth.write_polish_dataset('/tmp/tl', 42, 2000000, 60, 20000, synth.ONT, 0.01)
subprocess.run(['oracle/_ref/sam2bam', sam, '/tmp/tl/reads.bam'], check=True, stderr=subprocess.DEVNULL)
print(fa, n, L)
PY
P=oracle/_ref/data/params/polish/ont/r9.4/allParams.np.human.r94-g360.json
T=${1:-$(nproc)}
MARGIN_B200_TRACE=${TRACE:-} oracle/_ref/margin_b200cli polish /tmp/tl/reads.bam /tmp/tl/draft.fa $P -t $T -a info -o /tmp/tl/out 2>&1 | python scripts/timestamp_lines.py > gpurun_out/tl/timeline_t$T.log
grep -E " T03[ _]| margin_b200 " gpurun_out/tl/timeline_t$T.log | cut -c1-330 | head -60
tail -3 gpurun_out/tl/timeline_t$T.log
