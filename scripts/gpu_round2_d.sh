#!/bin/bash
mkdir -p gpurun_out
python scripts/gpu_f32_check.py > gpurun_out/d_f32.log 2>&1
cat gpurun_out/d_f32.log | cut -c1-400
