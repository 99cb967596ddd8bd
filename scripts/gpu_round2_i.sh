#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py -m gpu -q --timeout=1200 -p no:cacheprovider -k "forward or golden or errors or known" > gpurun_out/i_pytest.log 2>&1
echo "pytest exit $?" >> gpurun_out/i_pytest.log
tail -5 gpurun_out/i_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
python - <<'PY' > gpurun_out/i_polish100.log 2>&1
import os, subprocess, sys, tempfile, time, hashlib
sys.path.insert(0, '.')
from margin_b200 import synth
ref = 'oracle/_ref'
with tempfile.TemporaryDirectory() as tmp:
    fa, sam, n, L = synth.write_polish_dataset(tmp, 42, 100000, 60, 20000, synth.ONT, 0.01)
    bam = os.path.join(tmp, 'reads.bam')
    subprocess.run([os.path.join(ref, 'sam2bam'), sam, bam], check=True, stderr=subprocess.DEVNULL)
    for extra in ({}, {'MARGIN_B200_FWD_SLOW': '1'}):
        env = dict(os.environ, MARGIN_B200_FWD_STATS='1', MARGIN_B200_TRACE='1', **extra)
        t0 = time.time()
        r = subprocess.run([os.path.join(ref, 'margin_b200cli'), 'polish', bam, fa, os.path.join(ref, 'data/params/polish/ont/r9.4/allParams.np.human.r94-g360.json'), '-t', '1', '-o', os.path.join(tmp, 'o'), '-a', 'info'], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, env=env)
        print(r.stdout[-6000:])
        print('wall', time.time() - t0, extra, 'md5', hashlib.md5(open(os.path.join(tmp, 'o.fa'), 'rb').read()).hexdigest())
PY
grep -E "margin_b200 forward|margin_b200 bubble|fwd |wall" gpurun_out/i_polish100.log | cut -c1-300
