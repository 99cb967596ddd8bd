"""BASELINE configs[4] shape at reduced length: a diploid multi-contig draft + 40x ONT-like reads as an indexed BAM
(margin_b200/synth.py: write_polish_dataset per contig in worker processes, merged here; oracle/_ref/sam2bam).

  python scripts/make_c5_dataset.py --contigs 8 --kb 2500 --depth 40 --out data_c5 [--workers 8]
"""
import argparse
import multiprocessing as mp
import os
import shutil
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from margin_b200 import synth  # noqa: E402


def one(job):
    k, tmp, seed, kb, depth = job
    d = os.path.join(tmp, "c%d" % k)
    fa, sam, n, length = synth.write_polish_dataset(d, seed + k, kb * 1000, depth, 20000, synth.ONT, 0.01,
                                                    contig_name="contig_%d" % (k + 1), het=synth.HET)
    return k, fa, sam, n, length


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--contigs", type=int, default=8)
    ap.add_argument("--kb", type=int, default=2500)
    ap.add_argument("--depth", type=int, default=40)
    ap.add_argument("--seed", type=int, default=42)
    ap.add_argument("--out", default="data_c5")
    ap.add_argument("--workers", type=int, default=8)
    a = ap.parse_args()
    t0 = time.time()
    os.makedirs(a.out, exist_ok=True)
    with tempfile.TemporaryDirectory() as tmp:
        with mp.Pool(a.workers) as pool:
            parts = sorted(pool.map(one, [(k, tmp, a.seed, a.kb, a.depth) for k in range(a.contigs)]))
        fa = os.path.join(a.out, "draft.fa")
        sam = os.path.join(tmp, "reads.sam")
        offset, reads, total = 0, 0, 0
        with open(fa, "w") as ffa, open(fa + ".fai", "w") as ffai, open(sam, "w") as fsam:
            fsam.write("@HD\tVN:1.6\tSO:coordinate\n")
            for k, pfa, psam, n, length in parts:
                fsam.write("@SQ\tSN:contig_%d\tLN:%d\n" % (k + 1, length))
            for k, pfa, psam, n, length in parts:
                name = "contig_%d" % (k + 1)
                text = open(pfa).read()
                ffa.write(text)
                ffai.write("%s\t%d\t%d\t80\t81\n" % (name, length, offset + len(name) + 2))
                offset += len(text)
                with open(psam) as fh:
                    for line in fh:
                        if not line.startswith("@"):
                            fsam.write(line)
                reads += n
                total += length
        subprocess.run([os.path.join(ROOT, "oracle", "_ref", "sam2bam"), sam, os.path.join(a.out, "reads.bam")], check=True)
    with open(os.path.join(a.out, "README.txt"), "w") as fh:
        fh.write("scripts/make_c5_dataset.py --contigs %d --kb %d --depth %d --seed %d: %d reads, %d bases of draft, %.0f s\n"
                 % (a.contigs, a.kb, a.depth, a.seed, reads, total, time.time() - t0))
    print(open(os.path.join(a.out, "README.txt")).read())
    os.system("ls -la %s" % a.out)
