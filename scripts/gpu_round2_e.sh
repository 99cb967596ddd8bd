#!/bin/bash
mkdir -p gpurun_out
export MARGIN_B200_PRECISION=f32
python scripts/ncu_target.py > gpurun_out/e_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:tile_kernel -s 4 -c 2 -f -o gpurun_out/e_f32_tiles python scripts/ncu_target.py > gpurun_out/e_ncu.log 2>&1
cat gpurun_out/e_plain.log | tail -3
tail -5 gpurun_out/e_ncu.log
# FASTA identity of margin polish with the FP32 tile kernels on the bundled region
d=oracle/_ref/data
( time oracle/_ref/margin_b200cli polish $d/HG002.r94g360.chr20_59M_100k.bam $d/hg38.chr20_59M_100k.fa $d/params/polish/ont/r9.4/allParams.np.human.r94-g360.json -t 2 -o /tmp/f32 -a info ) > gpurun_out/e_polish_f32.log 2>&1
md5sum /tmp/f32.fa | tee -a gpurun_out/e_polish_f32.log
grep "margin_b200 poa_realign" gpurun_out/e_polish_f32.log | cut -c1-200
