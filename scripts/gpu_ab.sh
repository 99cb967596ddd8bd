#!/bin/bash
# A/B of kernel builds on the C2 chunk: libraries under margin_b200/csrc/ab/ against the current one, both precisions.
for lib in "" $(ls margin_b200/csrc/ab/*.so 2>/dev/null); do
  for prec in f64 f32; do
    echo "== lib=${lib:-HEAD} $prec"
    MARGIN_B200_LIB=${lib:+$PWD/$lib} MARGIN_B200_PRECISION=$prec python scripts/ncu_target.py 2>&1 | tail -1 | sed -E "s/.*'guard_failures': [0-9]+\} //"
  done
done
