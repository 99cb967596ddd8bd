"""gprof flat profile of the host side of `margin polish` with the binding (synthetic 100 kb chunk, one thread).

  make -C oracle polish_pg && python scripts/profile_polish_gprof.py   # on a GPU box; writes gpurun_out/pg/gmon.out
  gprof -b -p oracle/_ref/margin_b200cli_pg gpurun_out/pg/gmon.out
"""
import os, subprocess, sys, tempfile, time
ROOT = os.getcwd()
sys.path.insert(0, ROOT)
from margin_b200 import synth
ref = os.path.join(ROOT, "oracle", "_ref")
out = os.path.join(ROOT, "gpurun_out", "pg"); os.makedirs(out, exist_ok=True)
with tempfile.TemporaryDirectory() as tmp:
    fa, sam, n, length = synth.write_polish_dataset(tmp, 42, 100000, 60, 20000)
    bam = os.path.join(tmp, "reads.bam")
    subprocess.run([os.path.join(ref, "sam2bam"), sam, bam], check=True)
    t = time.time()
    r = subprocess.run([os.path.join(ref, "margin_b200cli_pg"), "polish", bam, fa,
                        os.path.join(ref, "data/params/polish/ont/r9.4/allParams.np.human.r94-g360.json"),
                        "-t", "1", "-o", os.path.join(tmp, "out"), "-a", "info"], cwd=out,
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    print("rc", r.returncode, "wall %.2f" % (time.time() - t))
    open(os.path.join(out, "log.txt"), "w").write(r.stdout)
