"""Stage timings of the CLI linked against libmargin_b200.so on a synthetic contig (`-a info` log of margin polish).

  python scripts/polish_synth_info.py KB THREADS OUT_LOG
"""
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from margin_b200 import synth  # noqa: E402

if __name__ == "__main__":
    kb, threads, out_log = int(sys.argv[1]), sys.argv[2], sys.argv[3]
    ref = os.path.join(ROOT, "oracle", "_ref")
    with tempfile.TemporaryDirectory() as tmp:
        fa, sam, n, length = synth.write_polish_dataset(tmp, 42, kb * 1000, 60, 20000)
        bam = os.path.join(tmp, "reads.bam")
        subprocess.run([os.path.join(ref, "sam2bam"), sam, bam], check=True)
        t = time.time()
        r = subprocess.run([os.path.join(ref, "margin_b200cli"), "polish", bam, fa,
                            os.path.join(ref, "data/params/polish/ont/r9.4/allParams.np.human.r94-g360.json"),
                            "-t", threads, "-o", os.path.join(tmp, "out"), "-a", "info"],
                           stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
        print("rc", r.returncode, "wall %.2f s" % (time.time() - t))
        open(out_log, "w").write(r.stdout)
