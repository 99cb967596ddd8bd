#!/bin/bash
# Round-2 evidence pass: the whole GPU suite, the bench lines (own arm, reference arm, C1 replay), one ncu launch list.
cd /root/repo; mkdir -p gpurun_out/final; O=gpurun_out/final
python -m pytest tests -m gpu -q --timeout=2400 -p no:cacheprovider --durations=6 -s > $O/pytest_gpu.log 2>&1
echo "pytest exit $?" >> $O/pytest_gpu.log
grep -E "corpus f|passed|failed|pytest exit|FAILED|Error" $O/pytest_gpu.log | cut -c1-300 | tail -12
python bench.py > $O/bench.json 2> $O/bench.err; tail -2 $O/bench.err
python bench.py --impl reference --steps 2 --warmup 1 --polish-synth 0 > $O/bench_reference.json 2> $O/bench_reference.err
python bench.py --workload c1_replay --steps 5 --warmup 2 > $O/bench_c1_replay.json 2> $O/bench_c1_replay.err
python bench.py --steps 2 --warmup 1 --no-sustained --no-cpu-baseline --no-fast-mode --polish-synth 0 > $O/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches.csv python bench.py --steps 2 --warmup 1 --no-sustained --no-cpu-baseline --no-fast-mode --polish-synth 0 > $O/ncu_list.log 2>&1
cut -c1-600 $O/bench.json
