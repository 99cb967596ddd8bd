"""Small program for ncu: three resident sweeps of the C2 chunk (precision from MARGIN_B200_PRECISION)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from margin_b200 import abi, model, synth  # noqa: E402
from margin_b200.aligner import PairHmmAligner  # noqa: E402

pm = model.load_flat_model(os.path.join(ROOT, "tests", "golden", "model_r94g360.bin"))
items = synth.workload("c2_ont_100kb_60x_20kb", pm, seed=42)
batch = abi.Batch(items, compact=True)
al = PairHmmAligner(pm.models, pm.p, device=0)
al.set_kernel_timing(True)
al.upload(batch)
for _ in range(3):
    st = al.posteriors_resident(True)
print(al.arithmetic(), st, {k: round(v[0], 3) for k, v in al.kernel_times().items() if v[1]})
al.close()
