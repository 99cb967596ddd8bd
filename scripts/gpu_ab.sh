#!/bin/bash
mkdir -p gpurun_out
for lib in "" margin_b200/csrc/ab/lib_a26fe56.so margin_b200/csrc/ab/lib_dcdc04e.so; do
  for prec in f64 f32; do
    echo "== lib=${lib:-HEAD} $prec"
    MARGIN_B200_LIB=${lib:+$PWD/$lib} MARGIN_B200_PRECISION=$prec python scripts/ncu_target.py 2>&1 | tail -1 | sed -E "s/.*'guard_failures': [0-9]+\} //"
  done
done
