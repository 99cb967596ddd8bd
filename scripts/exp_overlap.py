"""Do host<->device copies of one context overlap the kernels of another?  Thread A: resident sweeps only.
Thread B: upload + download only.  Each alone, then together."""
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from margin_b200 import abi  # noqa: E402
from margin_b200.aligner import PairHmmAligner  # noqa: E402


def main():
    pm, items = bench.build_items(bench.DEFAULT_WORKLOAD, 42)
    batch = abi.Batch(items)
    bench.pin_batch(batch)
    a = PairHmmAligner(pm.models, pm.p, device=0)
    b = PairHmmAligner(pm.models, pm.p, device=0)
    for al in (a, b):
        al.upload(batch)
        al.posteriors_resident(True)
        al.download(batch.n)
    n = 10

    def compute():
        t0 = time.perf_counter()
        for _ in range(n):
            a.posteriors_resident(True)
        return (time.perf_counter() - t0) / n * 1e3

    def copies():
        t0 = time.perf_counter()
        for _ in range(n):
            b.upload(batch)
            b.download(batch.n)
        return (time.perf_counter() - t0) / n * 1e3

    print("alone: compute %.3f ms, upload+download %.3f ms" % (compute(), copies()))
    res = {}
    ta = threading.Thread(target=lambda: res.__setitem__("c", compute()))
    tb = threading.Thread(target=lambda: res.__setitem__("x", copies()))
    ta.start(); tb.start(); ta.join(); tb.join()
    print("together: compute %.3f ms, upload+download %.3f ms" % (res["c"], res["x"]))


if __name__ == "__main__":
    main()
