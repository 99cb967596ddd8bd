"""Full-size checks on BASELINE.json configs[1] (synthetic 100 kb contig, 60x 20 kb ONT-like reads: 300 alignments,
~50 M band cells), where the CPU oracle would take minutes: size-independent properties of the posteriors and the
identity of the two execution modes of the library.

Properties of a pair-HMM posterior (impl/pairwiseAligner.c:598-681): every complete path emits read symbol y exactly
once, in state match or gapY, so  sum_x P(match x,y) + sum_x P(gapY x,y) = 1  for every y, and likewise
sum_y P(match x,y) + sum_y P(gapX x,y) = 1  for every x inside the band.  The reference drops posteriors below the
threshold (0.01), so the emitted weights of a row / column sum to at most 1 (+ the error of the cubic logAdd,
SURVEY.md D6: up to 5.5e-4 per operation, systematic) and, on anchored data, to almost 1."""
import os

import numpy as np
import pytest

from margin_b200 import model, synth
from margin_b200.aligner import PairHmmAligner

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PAIR_PROBABILITY_ONE = 10000000  # PAIR_ALIGNMENT_PROB_1, inc/pairwiseAligner.h


@pytest.fixture(scope="module")
def chunk():
    pm = model.load_flat_model(os.path.join(ROOT, "tests", "golden", "model_r94g360.bin"))
    return pm, synth.workload("c2_ont_100kb_60x_20kb", pm, seed=7)


def _run(pm, items, monkeypatch, mode):
    monkeypatch.setenv("MARGIN_B200_MODE", mode)
    al = PairHmmAligner(pm.models, pm.p, 0)
    res = al.posteriors_batch(items, True)
    lists = [res[i] for i in range(len(items))]
    stats = dict(res.stats)
    al.close()
    return lists, stats


def test_row_and_column_sums_at_full_size(chunk, monkeypatch):
    pm, items = chunk
    lists, stats = _run(pm, items, monkeypatch, "tiles")
    assert stats["band_cells"] > 40e6 and len(lists) >= 290
    worst_hi, mean_rows, mean_cols = 0.0, [], []
    for (sx, sy, anc, slot, rl, rr), (m, gx, gy) in zip(items, lists):
        lx, ly = len(sx), len(sy)
        for l in (m, gx, gy):
            assert l["w"].min(initial=PAIR_PROBABILITY_ONE) >= 0.01 * PAIR_PROBABILITY_ONE - 1
            assert l["w"].max(initial=0) <= PAIR_PROBABILITY_ONE
        assert m["x"].min() >= 0 and m["x"].max() < lx and m["y"].min() >= 0 and m["y"].max() < ly
        rows = np.bincount(m["y"], weights=m["w"], minlength=ly) + np.bincount(gy["y"], weights=gy["w"], minlength=ly)
        cols = np.bincount(m["x"], weights=m["w"], minlength=lx) + np.bincount(gx["x"], weights=gx["w"], minlength=lx)
        worst_hi = max(worst_hi, rows.max() / PAIR_PROBABILITY_ONE, cols.max() / PAIR_PROBABILITY_ONE)
        mean_rows.append(rows.mean() / PAIR_PROBABILITY_ONE)
        mean_cols.append(cols.mean() / PAIR_PROBABILITY_ONE)
    assert worst_hi <= 1.0 + 5e-3, worst_hi          # never more than one (up to the logAdd approximation)
    assert min(mean_rows) > 0.85 and min(mean_cols) > 0.85, (min(mean_rows), min(mean_cols))


def test_tiled_and_sequential_modes_agree_at_full_size(chunk, monkeypatch):
    """The tiled product path against the sequential per-alignment sweep (operation-for-operation the reference) on
    the whole chunk: identical tuple coordinates in identical order, weights within one unit of 1e-7."""
    pm, items = chunk
    tiled, _ = _run(pm, items, monkeypatch, "tiles")
    seq, _ = _run(pm, items, monkeypatch, "sequential")
    ntot = ndiff = 0
    for a, b in zip(tiled, seq):
        for l in range(3):
            assert len(a[l]) == len(b[l])
            assert np.array_equal(a[l]["x"], b[l]["x"]) and np.array_equal(a[l]["y"], b[l]["y"])
            dw = np.abs(a[l]["w"].astype(np.int64) - b[l]["w"].astype(np.int64))
            assert dw.max(initial=0) <= 1
            ndiff += int((dw > 0).sum())
            ntot += len(dw)
    assert ntot > 8e6
    assert ndiff <= ntot // 20
