"""Sustained end-to-end throughput of N contexts on one GPU (host buffers in/out), for the pipeline experiments."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from margin_b200 import abi  # noqa: E402


def main():
    contexts = int(sys.argv[1]) if len(sys.argv) > 1 else 2
    rounds = int(sys.argv[2]) if len(sys.argv) > 2 else 4
    pm, items = bench.build_items(bench.DEFAULT_WORKLOAD, 42)
    batch = abi.Batch(items, compact=True)
    bench.pin_batch(batch)
    from margin_b200.aligner import PairHmmAligner
    al = PairHmmAligner(pm.models, pm.p, device=0)
    al.upload(batch)
    st = al.posteriors_resident(True)
    al.close()
    r = bench.time_sustained(pm, batch, 0, 2.0 * st["band_cells"], contexts=contexts, rounds=rounds)
    print("contexts=%d K=%s: %.3f ms/chunk, %.2f GCUPS" % (contexts, os.environ.get("MARGIN_B200_PIPELINE", "default"),
                                                         r["ms_per_chunk"], r["value"]))


if __name__ == "__main__":
    main()
