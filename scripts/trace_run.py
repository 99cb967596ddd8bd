"""Runs a few resident sweeps of the C2 workload with MARGIN_B200_TRACE=1 (host phase timings on stderr)."""
import os, sys, time
os.environ["MARGIN_B200_TRACE"] = "1"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
from margin_b200 import abi
from margin_b200.aligner import PairHmmAligner
pm, items = bench.build_items(bench.DEFAULT_WORKLOAD, 42)
batch = abi.Batch(items)
al = PairHmmAligner(pm.models, pm.p, device=0)
t = time.perf_counter(); al.upload(batch); print("upload ms", (time.perf_counter() - t) * 1e3)
for i in range(4):
    t = time.perf_counter(); st = al.posteriors_resident(True); print("resident ms", (time.perf_counter() - t) * 1e3, st["gpu_ms"])
t = time.perf_counter(); al.upload(batch); print("upload ms", (time.perf_counter() - t) * 1e3)
t = time.perf_counter(); res = al.download(batch.n); print("download ms", (time.perf_counter() - t) * 1e3)
t = time.perf_counter(); res = al.download(batch.n); print("download ms", (time.perf_counter() - t) * 1e3)
