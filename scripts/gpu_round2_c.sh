#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py tests/test_gpu_fullsize.py -m gpu -q --timeout=2400 -p no:cacheprovider > gpurun_out/c_pytest.log 2>&1
echo "pytest exit $?" >> gpurun_out/c_pytest.log
tail -15 gpurun_out/c_pytest.log
python scripts/gpu_debug_c.py > gpurun_out/c_sweep.log 2>&1
cat gpurun_out/c_sweep.log | cut -c1-600
