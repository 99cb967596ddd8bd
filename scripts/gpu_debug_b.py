"""GPU debug pass: seam statistics and host-side phase traces on the C1 replay batches and on the synthetic C2 chunk."""
import os
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from margin_b200 import abi, model, replay, synth  # noqa: E402
from margin_b200.aligner import PairHmmAligner  # noqa: E402

os.environ["MARGIN_B200_GUARD_STATS"] = "1"
os.environ["MARGIN_B200_TRACE"] = "1"


def run(name, items, models, params, want):
    batch = abi.Batch(items, compact=True)
    al = PairHmmAligner(models, params, device=0)
    al.set_pair_format(abi.PAIRS_PACKED64)
    al.upload(batch)
    for k in range(3):
        sys.stderr.write("---- %s resident pass %d\n" % (name, k))
        sys.stderr.flush()
        st = al.posteriors_resident(bool(want))
    sys.stderr.write("%s stats: %s\n" % (name, st))
    for k in range(3):
        sys.stderr.write("---- %s host call %d\n" % (name, k))
        sys.stderr.flush()
        t0 = time.perf_counter()
        al.posteriors_host(batch, bool(want))
        sys.stderr.write("host call wall %.2f ms\n" % ((time.perf_counter() - t0) * 1e3))
    al.close()


pm = model.load_flat_model(os.path.join(ROOT, "tests", "golden", "model_r94g360.bin"))
items = synth.workload("c2_ont_100kb_60x_20kb", pm, seed=42)
run("c2", items, pm.models, pm.p, 1)
ref = os.path.join(ROOT, "oracle", "_ref")
with tempfile.TemporaryDirectory() as tmp:
    os.environ.pop("MARGIN_B200_TRACE")
    os.environ.pop("MARGIN_B200_GUARD_STATS")
    dumps = replay.dump_bundled_region(os.path.join(tmp, "d"), ref)
    os.environ["MARGIN_B200_GUARD_STATS"] = "1"
    os.environ["MARGIN_B200_TRACE"] = "1"
    for path in dumps:
        it, models, params, want = replay.load_dump(path)
        run(os.path.basename(path), it, models, params, want)
