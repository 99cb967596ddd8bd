#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py tests/test_gpu_fullsize.py -m gpu -q --timeout=2400 -p no:cacheprovider > gpurun_out/g_pytest.log 2>&1
echo "pytest exit $?" >> gpurun_out/g_pytest.log
tail -6 gpurun_out/g_pytest.log
python bench.py > gpurun_out/g_bench_f64.json 2> gpurun_out/g_bench_f64.err; tail -3 gpurun_out/g_bench_f64.err
python bench.py --precision f32 --no-cpu-baseline > gpurun_out/g_bench_f32.json 2> gpurun_out/g_bench_f32.err; tail -3 gpurun_out/g_bench_f32.err
python bench.py --steps 2 --warmup 1 --no-sustained --no-cpu-baseline --polish-synth 0 > gpurun_out/g_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/g_launches.csv python bench.py --steps 2 --warmup 1 --no-sustained --no-cpu-baseline --polish-synth 0 > gpurun_out/g_ncu_list.log 2>&1
MARGIN_B200_PRECISION=f64 python scripts/ncu_target.py > gpurun_out/g_plain2.log 2>&1 &&
MARGIN_B200_PRECISION=f64 ncu --set full --clock-control none --import-source on -k regex:tile_kernel -s 4 -c 2 -f -o gpurun_out/g_f64_tiles python scripts/ncu_target.py > gpurun_out/g_ncu_full.log 2>&1
MARGIN_B200_PRECISION=f32 ncu --set full --clock-control none --import-source on -k regex:tile_kernel -s 4 -c 2 -f -o gpurun_out/g_f32_tiles python scripts/ncu_target.py >> gpurun_out/g_ncu_full.log 2>&1
python - <<'PY' > gpurun_out/g_polish100.log 2>&1
import os, subprocess, sys, tempfile, time
sys.path.insert(0, '.')
from margin_b200 import synth
ref = 'oracle/_ref'
with tempfile.TemporaryDirectory() as tmp:
    fa, sam, n, L = synth.write_polish_dataset(tmp, 42, 100000, 60, 20000, synth.ONT, 0.01)
    bam = os.path.join(tmp, 'reads.bam')
    subprocess.run([os.path.join(ref, 'sam2bam'), sam, bam], check=True, stderr=subprocess.DEVNULL)
    env = dict(os.environ, MARGIN_B200_FWD_STATS='1', MARGIN_B200_TRACE='1')
    t0 = time.time()
    r = subprocess.run([os.path.join(ref, 'margin_b200cli'), 'polish', bam, fa, os.path.join(ref, 'data/params/polish/ont/r9.4/allParams.np.human.r94-g360.json'), '-t', '1', '-o', os.path.join(tmp, 'o'), '-a', 'info'], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, env=env)
    print(r.stdout[-12000:])
    print('wall', time.time() - t0)
PY
grep -E "margin_b200 forward|margin_b200 poa|margin_b200 bubble|fwd classify|wall" gpurun_out/g_polish100.log | cut -c1-300
