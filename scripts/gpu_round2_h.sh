#!/bin/bash
mkdir -p gpurun_out
python - <<'PY' > gpurun_out/h_polish_f32.log 2>&1
import os, subprocess, sys, tempfile, time
sys.path.insert(0, '.')
from margin_b200 import synth
ref = 'oracle/_ref'
with tempfile.TemporaryDirectory() as tmp:
    fa, sam, n, L = synth.write_polish_dataset(tmp, 42, 300000, 60, 20000, synth.ONT, 0.01)
    bam = os.path.join(tmp, 'reads.bam')
    subprocess.run([os.path.join(ref, 'sam2bam'), sam, bam], check=True, stderr=subprocess.DEVNULL)
    env = dict(os.environ, MARGIN_B200_GUARD_STATS='1', MARGIN_B200_PRECISION='f32')
    t0 = time.time()
    r = subprocess.run([os.path.join(ref, 'margin_b200cli'), 'polish', bam, fa, os.path.join(ref, 'data/params/polish/ont/r9.4/allParams.np.human.r94-g360.json'), '-t', '3', '-o', os.path.join(tmp, 'o'), '-a', 'info'], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, env=env)
    print(r.stdout[-30000:])
    print('wall', time.time() - t0, 'rc', r.returncode)
PY
grep -E "guard\]|internal|rc " gpurun_out/h_polish_f32.log | cut -c1-260 | tail -40
