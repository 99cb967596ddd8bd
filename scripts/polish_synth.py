"""`margin polish` on a synthetic multi-chunk contig: reference CLI vs the CLI linked against libmargin_b200.so
(bench.py: time_polish_synth).  Prints one JSON object.

  python scripts/polish_synth.py --kb 500 --threads 16 [--skip-reference] [--seed 42] [--diploid --depth 40]
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402

if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--kb", type=int, default=500)
    ap.add_argument("--threads", type=int, default=16)
    ap.add_argument("--seed", type=int, default=42)
    ap.add_argument("--skip-reference", action="store_true")
    ap.add_argument("--diploid", action="store_true", help="two haplotypes, margin polish --diploid (BASELINE configs[4] shape)")
    ap.add_argument("--depth", type=int, default=60)
    ap.add_argument("--platform", choices=["ont", "hifi"], default="ont", help="hifi: BASELINE configs[3] (use --depth 30)")
    a = ap.parse_args()
    print(json.dumps(bench.time_polish_synth(a.kb, min(bench.host_cores(), a.threads), a.seed, a.skip_reference, a.diploid, a.depth, a.platform)))
