#!/bin/bash
# GPU pass A of round 2: whole GPU test suite, default bench, C1 replay.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm --format=csv > gpurun_out/a_smi.txt 2>&1
python -m pytest tests -m gpu -q -x --timeout=2400 -p no:cacheprovider > gpurun_out/a_pytest.log 2>&1
echo "pytest exit $?" >> gpurun_out/a_pytest.log
tail -30 gpurun_out/a_pytest.log
python bench.py --steps 10 --warmup 3 > gpurun_out/a_bench.json 2> gpurun_out/a_bench.err
tail -c 3000 gpurun_out/a_bench.json
python bench.py --workload c1_replay --steps 5 --warmup 2 > gpurun_out/a_replay.json 2> gpurun_out/a_replay.err
tail -c 4000 gpurun_out/a_replay.json
tail -5 gpurun_out/a_replay.err
