"""GPU pass: FP32 fast mode against the oracle (max |dp|, tuple-set differences) and against FP64 on the C2 chunk."""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import gen  # noqa: E402
from margin_b200 import abi, model, synth  # noqa: E402
from margin_b200.aligner import PairHmmAligner  # noqa: E402
from oracle.oracle_py import Oracle  # noqa: E402


HIST = np.zeros(8, np.int64)


def diff_lists(got, want):
    """(tuples in want, missing, extra, max |dw| on common cells)."""
    n = miss = extra = 0
    worst = 0
    for l in range(3):
        kg = got[l]["x"].astype(np.int64) * (1 << 32) + got[l]["y"].astype(np.int64) + (1 << 31)
        kw = want[l]["x"].astype(np.int64) * (1 << 32) + want[l]["y"].astype(np.int64) + (1 << 31)
        n += len(kw)
        common, ig, iw = np.intersect1d(kg, kw, return_indices=True)
        miss += len(kw) - len(common)
        extra += len(kg) - len(common)
        if len(common):
            worst = max(worst, int(np.abs(got[l]["w"][ig].astype(np.int64) - want[l]["w"][iw].astype(np.int64)).max()))
        # order of the common part must be the same
        if len(common) == len(kg) == len(kw) and not np.array_equal(kg, kw):
            extra += 10 ** 6
        if len(common):
            dw = np.abs(got[l]["w"][ig].astype(np.int64) - want[l]["w"][iw].astype(np.int64))
            HIST[:] += np.bincount(np.minimum(np.floor(np.log10(np.maximum(dw, 1))).astype(int), 7), minlength=8)
    return n, miss, extra, worst


pm = model.load_flat_model(os.path.join(ROOT, "tests", "golden", "model_r94g360.bin"))
orc = Oracle()
rng = np.random.default_rng(99)
cases = {
    "random (short tracebacks)": ([gen.random_case(rng, int(rng.integers(0, 1500))) for _ in range(64)], pm.p.copy(min_diags_between_traceback=150)),
    "long, shipped parameters": ([gen.gapped_anchor_case(rng, 9000, 0, n_gaps=0) for _ in range(6)], pm.p),
    "corridors + low complexity": ([gen.long_indel_case(rng, 4000, 1500, 611, True), gen.long_indel_case(rng, 4000, 1500, 300, False),
                                    gen.low_complexity_case(rng, 3000, 2), gen.low_complexity_case(rng, 3000, 3),
                                    gen.gapped_anchor_case(rng, 6000, 110)], pm.p),
}
for name, (items, p) in cases.items():
    refs = [orc.posteriors(pm.models[it[3]], p, it[0], it[1], it[2], it[4], it[5], 1) for it in items]
    for prec in (abi.F64, abi.F32):
        for env in ({}, {"MARGIN_B200_TILE_LEN": "200"}):
            os.environ.pop("MARGIN_B200_TILE_LEN", None)
            os.environ.update(env)
            al = PairHmmAligner(pm.models, p, device=0)
            al.set_precision(prec)
            got = al.posteriors_batch(items, True)
            tot = [0, 0, 0, 0]
            for i in range(len(items)):
                n, m, e, w = diff_lists(got[i], refs[i])
                tot[0] += n
                tot[1] += m
                tot[2] += e
                tot[3] = max(tot[3], w)
            print("%-28s %s %-26s: %d tuples, %d missing, %d extra, max |dw| %d (1e-7 units); tiles %d wide %d seq %d seams %d failed %d" % (
                name, al.arithmetic(), env, tot[0], tot[1], tot[2], tot[3], got.stats["tiles"], got.stats["wide_cells"],
                got.stats["sequential_cells"], got.stats["guard_seams"], got.stats["guard_failures"]), flush=True)
            al.close()
os.environ.pop("MARGIN_B200_TILE_LEN", None)

# ---- C2 chunk: FP32 against FP64 at full size, and kernel times
items = synth.workload("c2_ont_100kb_60x_20kb", pm, seed=42)
batch = abi.Batch(items, compact=True)
res = {}
for prec in (abi.F64, abi.F32):
    al = PairHmmAligner(pm.models, pm.p, device=0)
    al.set_precision(prec)
    al.set_kernel_timing(True)
    al.upload(batch)
    for _ in range(3):
        st = al.posteriors_resident(True)
    ms = []
    for _ in range(8):
        st = al.posteriors_resident(True)
        ms.append(st["gpu_ms"])
    kt = {k: round(v[0], 3) for k, v in al.kernel_times().items() if v[1]}
    print("C2 %s: %.3f ms/step (min %.3f) = %.1f GCUPS; kernels %s; tiles %d seams %d failed %d" % (
        al.arithmetic(), sum(ms) / len(ms), min(ms), 2 * st["band_cells"] / min(ms) * 1e-6, kt, st["tiles"], st["guard_seams"],
        st["guard_failures"]), flush=True)
    al.set_kernel_timing(False)
    got = al.posteriors_batch(batch, True)
    res[prec] = [got[i] for i in range(len(items))]
    al.close()
tot = [0, 0, 0, 0]
HIST[:] = 0
for a, b in zip(res[abi.F32], res[abi.F64]):
    n, m, e, w = diff_lists(a, b)
    tot[0] += n
    tot[1] += m
    tot[2] += e
    tot[3] = max(tot[3], w)
print("C2 f32 vs f64: %d tuples, %d missing, %d extra, max |dw| %d (1e-7 units)" % tuple(tot))
print("   |dw| histogram by decade (1e-7 units: <10, <100, <1e3, <1e4, ...):", HIST.tolist())
