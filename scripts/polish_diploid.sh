#!/bin/bash
# `margin polish --diploid` on the bundled ONT region: the unmodified reference CLI against the CLI linked to
# libmargin_b200.so (bubble graphs of the phasing path batched by the binding).  Prints wall times and output digests.
#   bash scripts/polish_diploid.sh [threads] [out_dir]
R=oracle/_ref; D=$R/data; T=${1:-8}; O=${2:-/tmp/diploid}; mkdir -p $O
P=$D/params/polish/ont/r9.4/allParams.np.human.r94-g360.json
for arm in margin_ref margin_b200cli; do
  s=$(date +%s.%N)
  timeout ${TIMEOUT:-900} $R/$arm polish $D/HG002.r94g360.chr20_59M_100k.bam $D/hg38.chr20_59M_100k.fa $P -t $T -o $O/$arm -2 -a info > $O/$arm.log 2>&1
  rc=$?
  e=$(date +%s.%N)
  echo "$arm rc=$rc wall=$(awk -v a=$s -v b=$e 'BEGIN{printf "%.1f", b - a}') s"
  grep -E "margin_b200 bubble graph" $O/$arm.log | head -4
  for f in $O/$arm.fa.hap1 $O/$arm.fa.hap2 $O/$arm.chunks.csv; do echo "  $(md5sum < $f | cut -c1-32) $(basename $f)"; done
done
