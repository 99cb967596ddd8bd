#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -q --timeout=2400 -p no:cacheprovider --durations=6 -s > gpurun_out/j_pytest.log 2>&1
echo "pytest exit $?" >> gpurun_out/j_pytest.log
grep -E "corpus f|tie:|passed|failed|pytest exit|FAILED|Error" gpurun_out/j_pytest.log | cut -c1-400 | tail -40
python bench.py > gpurun_out/j_bench_f64.json 2> gpurun_out/j_bench_f64.err; tail -2 gpurun_out/j_bench_f64.err
python bench.py --precision f32 --no-cpu-baseline --polish-synth 0 > gpurun_out/j_bench_f32.json 2> gpurun_out/j_bench_f32.err; tail -2 gpurun_out/j_bench_f32.err
