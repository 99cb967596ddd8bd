"""Throughput of the forward-only batch (computeForwardProbability) on bubble-sized DPs."""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from margin_b200 import abi, model, synth  # noqa: E402
from margin_b200.aligner import PairHmmAligner  # noqa: E402


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 100000
    pm = model.load_flat_model(os.path.join(ROOT, "tests", "golden", "model_r94g360.bin"))
    max_len = int(sys.argv[2]) if len(sys.argv) > 2 else 43
    items = synth.bubble_workload(7, n, max_len=max_len)
    batch = abi.Batch(items)
    cells = int(sum((len(it[0]) + 1) * (len(it[1]) + 1) for it in items))
    al = PairHmmAligner(pm.models, pm.p, device=0)
    for _ in range(2):
        al.forward_batch(batch)
    t0 = time.perf_counter()
    reps = 5
    for _ in range(reps):
        al.forward_batch(batch)
    dt = (time.perf_counter() - t0) / reps
    al.upload(batch)
    al.set_kernel_timing(True)
    st = al.forward_resident()
    st = al.forward_resident()
    print({k: v for k, v in al.kernel_times().items() if v[1]})
    print("forward-only: %d DPs, %d cells: e2e %.3f ms (%.2f GCUPS), resident %.3f ms (%.2f GCUPS), launches %d" % (
        n, cells, dt * 1e3, cells / dt * 1e-9, st["gpu_ms"], cells / (st["gpu_ms"] * 1e-3) * 1e-9, st["kernel_launches"]))
    al.close()


if __name__ == "__main__":
    main()
