#!/bin/bash
# packed symbols + anchor runs: parity, binding suites, polish FASTA, bench
mkdir -p gpurun_out/h2d
cd /root/repo
timeout 900 python -m pytest tests/test_gpu_parity.py -q -x -k "packed or compact or delta32 or errors or forward_only or random_batch" > gpurun_out/h2d/parity.log 2>&1
echo "parity rc=$?" >> gpurun_out/h2d/parity.log
timeout 1500 python -m pytest tests/test_polish_e2e.py tests/test_reference_suites.py -q -x -m gpu > gpurun_out/h2d/polish.log 2>&1
echo "polish rc=$?" >> gpurun_out/h2d/polish.log
timeout 900 python bench.py > gpurun_out/h2d/bench.json 2> gpurun_out/h2d/bench.err
tail -3 gpurun_out/h2d/parity.log gpurun_out/h2d/polish.log
cat gpurun_out/h2d/bench.json
