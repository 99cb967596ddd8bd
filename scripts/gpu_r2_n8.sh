#!/bin/bash
# eight GPUs: the sharded bench line (own arm only, polish/CPU/fast-mode objects skipped), then the 20 Mb diploid polish
cd /root/repo; mkdir -p gpurun_out/n8
N=$(nvidia-smi -L | wc -l)
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 \
  bench.py --gpus $N --no-cpu-baseline --no-fast-mode --polish-synth 0 > gpurun_out/n8/bench_n$N.json 2> gpurun_out/n8/bench_n$N.err
echo "bench rc=$?"; cut -c1-1500 gpurun_out/n8/bench_n$N.json
TIMEOUT=1200 bash scripts/gpu_r2_c5.sh full
