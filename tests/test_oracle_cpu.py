"""CPU suite: pins oracle/pairhmm_oracle.c against (a) the committed golden vectors made by the reference,
(b) the reference itself where oracle/_ref is built, (c) the reference's known-answer tests."""
import os

import numpy as np
import pytest

import gen
from common import compare_lists, load_golden_cases
from margin_b200 import abi, model, symbols

REF_PARAMS = {
    "r94g360": "/root/reference/params/polish/ont/r9.4/allParams.np.human.r94-g360.json",
    "hifi": "/root/reference/params/polish/pacbio/hifi/allParams.hifi.json",
    "no_rle": "/root/reference/params/polish/misc/allParams.no_rle.json",
}


@pytest.mark.parametrize("name", ["r94g360", "r94g344", "hifi", "no_rle"])
def test_oracle_matches_golden_vectors(oracle, golden_model, name):
    pm = golden_model(name)
    for c in load_golden_cases(name):
        p = pm.p.copy(min_diags_between_traceback=c["min_diags"], dynamic_anchor_expansion=c["dynamic"])
        got = oracle.posteriors(pm.models[c["slot"]], p, c["sx"], c["sy"], c["a"], c["rl"], c["rr"], 1)
        assert compare_lists(got, c["lists"], tol_w=0) == 0
        assert oracle.forward(pm.models[c["slot"]], p, c["sx"][:60], c["sy"][:60], (), c["rl"], c["rr"]) == c["fwd"]
        kl, kr = oracle.band(c["a"], len(c["sx"]), len(c["sy"]), c["dynamic"], p.diagonal_expansion)
        assert np.array_equal(kl, c["kl"]) and np.array_equal(kr, c["kr"])


@pytest.mark.parametrize("name", sorted(REF_PARAMS))
def test_model_mirror_matches_reference_parser(reference, golden_model, name):
    """margin_b200.model (our JSON mirror) == the reference's parser, bit for bit; so are the golden files."""
    path = REF_PARAMS[name]
    if not os.path.exists(path):
        pytest.skip("reference params not present")
    ref = reference.load_params(path)
    pm = model.load_params(path)
    gm = golden_model(name)
    for s in (0, 1):
        assert pm.models[s].same_as(ref["models"][s])
        assert gm.models[s].same_as(ref["models"][s])
    assert bytes(pm.p) == bytes(ref["p"]) == bytes(gm.p)
    assert pm.use_repeat_counts == ref["use_repeat_counts"] and pm.use_rle == ref["use_rle"]
    # the flattened emission tables reproduce the reference's emission callbacks
    rng = np.random.default_rng(3)
    for _ in range(300):
        s = int(rng.integers(0, 2))
        cx, cy = (int(gen.random_symbols(rng, 1, pm.use_repeat_counts, 0.1)[0]) for _ in range(2))
        want = reference.emission_match(ref["sms"][s], cx, cy)
        m = pm.models[s]
        if m.rle:
            xb, yb = cx & 255, cy & 255
            sub = -2.772588722 if (xb >= 4 or yb >= 4) else m.match[xb * 4 + yb]
            got = sub + 2.3025 * m.rsm[((0 if xb >= 4 else xb) * 51 + (cx >> 8)) * 51 + (cy >> 8)]
        else:
            got = -2.772588722 if (cx >= 4 or cy >= 4) else m.match[cx * 4 + cy]
        assert got == want


def test_oracle_vs_reference_random(oracle, reference, golden_model):
    pm = golden_model("r94g360")
    sms = [reference.sm_from_model(pm.models[s]) for s in (0, 1)]
    rng = np.random.default_rng(11)
    for it in range(120):
        sx, sy, a, slot, rl, rr = gen.random_case(rng, int(rng.integers(0, 500)))
        p = pm.p.copy(min_diags_between_traceback=int(rng.choice([5000, 60, 100, 333])),
                      dynamic_anchor_expansion=int(rng.random() < 0.7))
        wi = int(rng.random() < 0.8)
        got = oracle.posteriors(pm.models[slot], p, sx, sy, a, rl, rr, wi)
        want = reference.posteriors(sms[slot], p, sx, sy, a, rl, rr, wi)
        assert compare_lists(got, want, tol_w=0) == 0
        assert oracle.forward(pm.models[slot], p, sx[:50], sy[:50], (), rl, rr) == \
            reference.forward(sms[slot], p, sx[:50], sy[:50], (), rl, rr)
        kl, kr = oracle.band(a, len(sx), len(sy), p.dynamic_anchor_expansion, p.diagonal_expansion)
        rkl, rkr = reference.band(a, len(sx), len(sy), p.dynamic_anchor_expansion, p.diagonal_expansion)
        assert np.array_equal(kl, rkl) and np.array_equal(kr, rkr)


def test_logadd(oracle, reference):
    """test_logAdd (tests/pairwiseAlignerTest.c:129-139) + bit equality with the reference."""
    rng = np.random.default_rng(5)
    for _ in range(2000):
        x, y = float(rng.uniform(-20, 5)), float(rng.uniform(-20, 5))
        assert oracle.logadd(x, y) == reference.logadd(x, y)
        assert abs(oracle.logadd(x, y) - np.logaddexp(x, y)) < 1e-3
    ninf = float("-inf")
    assert oracle.logadd(ninf, 1.5) == 1.5 and oracle.logadd(1.5, ninf) == 1.5 and oracle.logadd(ninf, ninf) == ninf


def test_bands_known_answer(oracle):
    """test_bands (tests/pairwiseAlignerTest.c:64-127): anchors (1,0),(2,1),(3,3), lX=6, lY=5, expansion 2."""
    anchors = [(1, 0, 2), (2, 1, 2), (3, 3, 2)]
    kl, kr = oracle.band(anchors, 6, 5, 0, 2)
    ref_walk = [(0, 0, 0), (1, -1, 1), (2, -2, 2), (3, -1, 3), (4, -2, 4), (5, -1, 3), (6, -2, 4), (7, -3, 3),
                (8, -2, 2), (9, -1, 3), (10, 0, 2), (11, 1, 1)]
    got = [(d, int(kl[d]), int(kr[d])) for d in range(12)]
    assert got == ref_walk


def test_diagonal_dp_known_answer(oracle):
    """test_diagonalDPCalculations (tests/pairwiseAlignerTest.c:257-342): AGCG vs AGTTCG, default nucleotide model,
    no banding constraints, threshold 0.2 => exactly the pairs (0,0),(1,1),(2,4),(3,5)."""
    m = model.default_nucleotide_model()
    sx = symbols.bases_to_symbols("AGCG")
    sy = symbols.bases_to_symbols("AGTTCG")
    p = abi.Params.defaults().copy(threshold=0.2)
    lists = oracle.posteriors(m, p, sx, sy, np.zeros((0, 3), np.int32), 0, 0, 0)
    pairs = sorted((int(t["x"]), int(t["y"])) for t in lists[0])
    assert pairs == [(0, 0), (1, 1), (2, 4), (3, 5)]


def test_split_points_known_answer(oracle, reference):
    """test_getSplitPoints (tests/pairwiseAlignerTest.c:489-559)."""
    ms = 2000 * 2000
    assert oracle.split_points([], 3000, 1000, ms, 0, 0).tolist() == [[0, 0, 3000, 1000]]
    lx, ly = 20000, 25000
    assert oracle.split_points([], lx, ly, ms, 1, 1).tolist() == []
    assert oracle.split_points([], lx, ly, ms, 1, 0).tolist() == [[18000, 23000, lx, ly]]
    assert oracle.split_points([], lx, ly, ms, 0, 1).tolist() == [[0, 0, 2000, 2000]]
    assert oracle.split_points([], lx, ly, ms, 0, 0).tolist() == [[0, 0, 2000, 2000], [18000, 23000, lx, ly]]
    anchors = [(2000, 2000, 0), (4002, 4001, 0), (5000, 5000, 0), (8000, 6000, 0), (9000, 9000, 0), (10000, 14000, 0),
               (15000, 15000, 0), (16000, 16000, 0)]
    assert oracle.split_points(anchors, lx, ly, ms, 0, 0).tolist() == [
        [0, 0, 3001, 3001], [3002, 3001, 9500, 11001], [9501, 12000, 12001, 14500], [13000, 14501, 18000, 18001],
        [18001, 23000, 20000, 25000]]
    rng = np.random.default_rng(8)
    for _ in range(200):
        sx, sy, a, slot, rl, rr = gen.random_case(rng, int(rng.integers(1, 300)), anchor_keep=0.05)
        lim = int(rng.choice([100, 400, 2500]))
        assert np.array_equal(oracle.split_points(a, len(sx), len(sy), lim, rl, rr),
                              reference.split_points(a, len(sx), len(sy), lim, rl, rr))
