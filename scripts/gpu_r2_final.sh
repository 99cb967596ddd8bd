#!/bin/bash
# Round-2 evidence pass: the new test alone, the whole GPU suite, the bench lines, ncu launch list and full captures.
mkdir -p gpurun_out
python -m pytest tests/test_zz_no_alloc.py -m gpu -q -p no:cacheprovider > gpurun_out/r2_noalloc_alone.log 2>&1
echo "pytest exit $?" >> gpurun_out/r2_noalloc_alone.log
tail -3 gpurun_out/r2_noalloc_alone.log
python -m pytest tests -m gpu -q --timeout=2400 -p no:cacheprovider --durations=6 -s > gpurun_out/r2_pytest_gpu.log 2>&1
echo "pytest exit $?" >> gpurun_out/r2_pytest_gpu.log
grep -E "corpus f|passed|failed|pytest exit|FAILED|Error" gpurun_out/r2_pytest_gpu.log | cut -c1-300 | tail -12
python bench.py > gpurun_out/r2_bench.json 2> gpurun_out/r2_bench.err; tail -2 gpurun_out/r2_bench.err
python bench.py --impl reference --steps 2 --warmup 1 --polish-synth 0 > gpurun_out/r2_bench_reference.json 2> gpurun_out/r2_bench_reference.err
python bench.py --workload c1_replay --steps 5 --warmup 2 > gpurun_out/r2_bench_c1_replay.json 2> gpurun_out/r2_bench_c1_replay.err
python bench.py --steps 2 --warmup 1 --no-sustained --no-cpu-baseline --no-fast-mode --polish-synth 0 > gpurun_out/r2_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_launches.csv python bench.py --steps 2 --warmup 1 --no-sustained --no-cpu-baseline --no-fast-mode --polish-synth 0 > gpurun_out/r2_ncu_list.log 2>&1
for prec in f64 f32; do
  MARGIN_B200_PRECISION=$prec python scripts/ncu_target.py > gpurun_out/r2_plain_$prec.log 2>&1 &&
  MARGIN_B200_PRECISION=$prec ncu --set full --clock-control none --import-source on -k regex:tile_kernel -s 4 -c 2 -f -o gpurun_out/r2_${prec}_tiles python scripts/ncu_target.py > gpurun_out/r2_ncu_full_$prec.log 2>&1
done
ls -la gpurun_out | tail -20
