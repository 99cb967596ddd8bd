#!/bin/bash
# Runs the patched margin CLI on the bundled region under several library settings and prints the stage log lines
# (score differences between settings point at a parity problem long before the FASTA changes).
R=oracle/_ref; D=$R/data
run() {
  echo "== $*"
  env "$@" $R/margin_b200cli polish $D/HG002.r94g360.chr20_59M_100k.bam $D/hg38.chr20_59M_100k.fa \
     $D/params/polish/ont/r9.4/allParams.np.human.r94-g360.json -t 2 -o /tmp/pm_out -a info 2>&1 | grep -E "T00 Took|rror" ; md5sum /tmp/pm_out.fa
}
run MARGIN_B200_MODE=tiles
run MARGIN_B200_MODE=tiles MARGIN_B200_PIPELINE=0
run MARGIN_B200_MODE=sequential
