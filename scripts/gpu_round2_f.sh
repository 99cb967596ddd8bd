#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -q --timeout=2400 -p no:cacheprovider --durations=8 > gpurun_out/f_pytest.log 2>&1
echo "pytest exit $?" >> gpurun_out/f_pytest.log
tail -25 gpurun_out/f_pytest.log
for prec in f64 f32; do
  MARGIN_B200_PRECISION=$prec python scripts/ncu_target.py 2>&1 | tail -1 | cut -c1-600
done
d=oracle/_ref/data
( time MARGIN_B200_TRACE=1 oracle/_ref/margin_b200cli polish $d/HG002.r94g360.chr20_59M_100k.bam $d/hg38.chr20_59M_100k.fa $d/params/polish/ont/r9.4/allParams.np.human.r94-g360.json -t 2 -o /tmp/tr -a info ) > gpurun_out/f_polish_trace.log 2>&1
grep -E "trace|poa_realign|real" gpurun_out/f_polish_trace.log | cut -c1-330 | head -60
