R=oracle/_ref; D=$R/data; P=$D/params/polish/ont/r9.4/allParams.np.human.r94-g360.json
MARGIN_B200_TRACE=1 $R/margin_b200cli polish $D/HG002.r94g360.chr20_59M_100k.bam $D/hg38.chr20_59M_100k.fa $P -t 8 -o /tmp/dtr -2 -a info > /tmp/dtr.log 2>&1
echo "forward calls: $(grep -c 'fwd classify' /tmp/dtr.log)  posterior calls: $(grep -c 'band+sync' /tmp/dtr.log)"
grep -E "Took|margin_b200 (bubble|poa)|Finished|partition|filtered" /tmp/dtr.log | cut -c1-220 | head -60
