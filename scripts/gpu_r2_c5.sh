#!/bin/bash
# BASELINE configs[4] shape (diploid, 40x ONT-like, margin polish --diploid), dataset from scripts/make_c5_dataset.py.
#   bash scripts/gpu_r2_c5.sh slice   one GPU: unmodified reference and this library on the 1 Mb contig, digests compared
#   bash scripts/gpu_r2_c5.sh full    N GPUs: this library on 8 x 2.5 Mb
cd /root/repo
mode=${1:-slice}; O=gpurun_out/c5; mkdir -p $O /tmp/c5
P=oracle/_ref/data/params/polish/ont/r9.4/allParams.np.human.r94-g360.json
T=$(nproc)
run() { # name binary dataset-dir
  s=$(date +%s.%N)
  timeout ${TIMEOUT:-1500} oracle/_ref/$2 polish $3/reads.bam $3/draft.fa $P -t $T -2 -a info -o /tmp/c5/$1 2>&1 | python scripts/timestamp_lines.py > $O/$1.log
  e=$(date +%s.%N)
  mb=$(awk '{s+=$2} END {printf "%.3f", s/1e6}' $3/draft.fa.fai)
  awk -v a=$s -v b=$e -v mb=$mb -v n=$1 -v t=$T 'BEGIN{printf "%s: %.1f s wall, %s Mb, %.1f Mb/hour, -t %d\n", n, b-a, mb, mb/(b-a)*3600, t}' | tee -a $O/summary.txt
  for f in /tmp/c5/$1.fa.hap1 /tmp/c5/$1.fa.hap2; do echo "  $(md5sum < $f | cut -c1-32) $(basename $f)" | tee -a $O/summary.txt; done
  grep -c "margin_b200 poa_realign" $O/$1.log | sed 's/^/  poa_realign sweeps: /' | tee -a $O/summary.txt
}
nvidia-smi --query-gpu=name --format=csv,noheader | sort | uniq -c | tee -a $O/summary.txt
if [ $mode = slice ]; then
  run slice_reference margin_ref data_c5/slice
  run slice_b200 margin_b200cli data_c5/slice
else
  run full_b200 margin_b200cli data_c5
fi
