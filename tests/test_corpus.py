"""The parity corpus (SURVEY.md §8 C-4 i): 10 240 fixed-seed alignments — short random problems with every edge case,
wide spots, long indels, tandem repeats, alignments of more than 5000 diagonals at the shipped parameters — against the
digests the UNMODIFIED reference produced for them (tests/golden/make_corpus.py -> corpus_digests.npz: per list the tuple
count, an order-sensitive hash of the (x, y) sequence, the weight sum).  Inputs are regenerated from the case index."""
import os
from concurrent.futures import ThreadPoolExecutor

import numpy as np
import pytest

import gen
from margin_b200 import abi

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _digests():
    z = np.load(os.path.join(GOLDEN, "corpus_digests.npz"))
    assert int(z["size"][0]) == gen.CORPUS_SIZE
    return z["count"], z["xy_hash"], z["weight_sum"]


def _check(i, lists, want, tol_per_tuple):
    count, xy_hash, weight_sum = want
    for l in range(3):
        n, h, w = gen.list_digest(lists[l])
        assert n == count[i, l], "case %d list %d: %d tuples, the reference has %d" % (i, l, n, count[i, l])
        assert np.uint64(h) == xy_hash[i, l], "case %d list %d: coordinates or their order differ" % (i, l)
        assert abs(w - int(weight_sum[i, l])) <= tol_per_tuple * n, "case %d list %d: weight sum off by %d over %d tuples" % (
            i, l, w - int(weight_sum[i, l]), n)


def test_oracle_on_a_corpus_sample(oracle, golden_model):
    """CPU: the restatement reproduces the reference's digests exactly on every 16th case (640 alignments)."""
    pm = golden_model("r94g360")
    want = _digests()

    def one(i):
        it, ps = gen.corpus_case(i)
        md, dyn = gen.CORPUS_PARAM_SETS[ps]
        p = pm.p.copy(min_diags_between_traceback=md, dynamic_anchor_expansion=dyn)
        return i, oracle.posteriors(pm.models[it[3]], p, it[0], it[1], it[2], it[4], it[5], 1)
    with ThreadPoolExecutor(os.cpu_count() or 1) as pool:
        for i, lists in pool.map(one, range(0, gen.CORPUS_SIZE, 16)):
            _check(i, lists, want, 0)


@pytest.mark.gpu
@pytest.mark.parametrize("precision", ["f64", "f32"])
def test_whole_corpus_on_the_gpu(golden_model, precision):
    """All 10 240 cases through the C ABI.  FP64: identical tuple sets in identical order, weights within one unit of
    1e-7 per tuple.  FP32 fast mode: the sets may differ at threshold ties (listed and bounded), weights within 1e-4."""
    from margin_b200.aligner import PairHmmAligner
    pm = golden_model("r94g360")
    want = _digests()
    by_ps = {}
    for i in range(gen.CORPUS_SIZE):
        it, ps = gen.corpus_case(i)
        by_ps.setdefault(ps, []).append((i, it))
    ties, n_tuples, stats = [], 0, {}
    for ps, cases in sorted(by_ps.items()):
        md, dyn = gen.CORPUS_PARAM_SETS[ps]
        al = PairHmmAligner(pm.models, pm.p.copy(min_diags_between_traceback=md, dynamic_anchor_expansion=dyn), device=0)
        al.set_precision(abi.F32 if precision == "f32" else abi.F64)
        for lo in range(0, len(cases), 1024):
            part = cases[lo:lo + 1024]
            res = al.posteriors_batch([it for _, it in part], True)
            for k in ("wide_cells", "sequential_cells", "guard_seams", "guard_failures", "band_cells", "tiles"):
                stats[k] = stats.get(k, 0) + int(res.stats[k])
            for (i, _), lists in zip(part, res.lists):
                n_tuples += sum(len(x) for x in lists)
                if precision == "f64":
                    _check(i, lists, want, 1)
                else:
                    try:
                        _check(i, lists, want, 1000)
                    except AssertionError as e:
                        ties.append(str(e))
        al.close()
    print("corpus %s: %d tuples, %s, %d cases with threshold ties" % (precision, n_tuples, stats, len(ties)))
    for t in ties[:20]:
        print("  tie:", t)
    assert n_tuples > 2.5e7
    # FP32: a posterior within ~1e-5 of the 0.01 threshold can fall on the other side (north_star: "threshold-tied cells,
    # which are listed"); 28.9 M tuples, so a few dozen cases at most
    assert len(ties) <= 60, ties[:5]
