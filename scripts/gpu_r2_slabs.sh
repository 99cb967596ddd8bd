#!/bin/bash
# observation slabs: binding suites + polish FASTA tests on the GPU, a timestamped single-chunk timeline, bench
mkdir -p gpurun_out/slabs
cd /root/repo
timeout 1500 python -m pytest tests/test_polish_e2e.py tests/test_reference_suites.py -q -x -m gpu > gpurun_out/slabs/polish.log 2>&1
echo "polish rc=$?" >> gpurun_out/slabs/polish.log
python - <<'PY' > gpurun_out/slabs/dataset.log 2>&1
import subprocess, sys
sys.path.insert(0, '/root/repo')
from margin_b200 import synth
fa, sam, n, L = synth.write_polish_dataset('/tmp', 42, 100000, 60, 20000, synth.ONT, 0.01)
subprocess.run(['oracle/_ref/sam2bam', sam, '/tmp/reads.bam'], check=True, stderr=subprocess.DEVNULL)
print(fa, n, L)
PY
P=oracle/_ref/data/params/polish/ont/r9.4/allParams.np.human.r94-g360.json
FA=$(head -1 gpurun_out/slabs/dataset.log | cut -d' ' -f1)
for slabs in 1 0; do
  MARGIN_B200_OBSERVATION_SLABS=$slabs oracle/_ref/margin_b200cli polish /tmp/reads.bam $FA $P -t 1 -a info -o /tmp/out_$slabs 2>&1 | python scripts/timestamp_lines.py > gpurun_out/slabs/timeline_slabs$slabs.log
  md5sum /tmp/out_$slabs.fa >> gpurun_out/slabs/timeline_slabs$slabs.log
done
timeout 900 python bench.py > gpurun_out/slabs/bench.json 2> gpurun_out/slabs/bench.err
tail -n 3 gpurun_out/slabs/polish.log
grep -o '"polish": {.*' gpurun_out/slabs/bench.json
