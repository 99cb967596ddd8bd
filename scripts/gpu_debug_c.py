"""GPU debug pass: warm-up / gap-threshold sweep of the seam guard on the C1 replay batches."""
import os
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from margin_b200 import abi, replay  # noqa: E402
from margin_b200.aligner import PairHmmAligner  # noqa: E402

ref = os.path.join(ROOT, "oracle", "_ref")
with tempfile.TemporaryDirectory() as tmp:
    dumps = replay.dump_bundled_region(os.path.join(tmp, "d"), ref)
    loaded = [replay.load_dump(p) for p in dumps]
    for env in ({}, {"MARGIN_B200_WARMUP": "192"}, {"MARGIN_B200_WARMUP": "256"}, {"MARGIN_B200_GAP_MIN": "8"},
                {"MARGIN_B200_GAP_MIN": "8", "MARGIN_B200_WARMUP": "192"}, {"MARGIN_B200_GUARD_TOL": "2e-8"},
                {"MARGIN_B200_GUARD": "0"}):
        for k in ("MARGIN_B200_WARMUP", "MARGIN_B200_GAP_MIN", "MARGIN_B200_GUARD_TOL", "MARGIN_B200_GUARD"):
            os.environ.pop(k, None)
        os.environ.update(env)
        for path, (items, models, params, want) in zip(dumps, loaded):
            batch = abi.Batch(items, compact=True)
            al = PairHmmAligner(models, params, device=0)
            al.set_pair_format(abi.PAIRS_PACKED64)
            al.set_kernel_timing(True)
            al.upload(batch)
            for _ in range(2):
                st = al.posteriors_resident(bool(want))
            ms = []
            for _ in range(4):
                st = al.posteriors_resident(bool(want))
                ms.append(st["gpu_ms"])
            kt = {k: round(v[0], 3) for k, v in al.kernel_times().items() if v[1]}
            print("%-40s %s: %.2f ms (min %.2f) tiles %d wide %d seq %d seams %d failed %d launches %d  kernels %s" % (
                env, os.path.basename(path), sum(ms) / len(ms), min(ms), st["tiles"], st["wide_cells"], st["sequential_cells"],
                st["guard_seams"], st["guard_failures"], st["kernel_launches"], kt), flush=True)
            al.close()
