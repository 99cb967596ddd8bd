"""GPU parity tests proper: everything goes through the C ABI (libmargin_b200.so) on cuda:0 and is compared with
the oracle, the committed golden vectors of the reference, and (where oracle/_ref travelled) the reference itself."""
import os

import numpy as np
import pytest

import gen
from common import compare_lists, load_golden_cases
from margin_b200 import abi, model, symbols
from margin_b200.aligner import AlignerError, PairHmmAligner

pytestmark = pytest.mark.gpu

# Execution modes of the library (margin_b200/csrc/pairhmm_tiles.cuh): the default tiled mode, the tiled mode with
# tiles much shorter than the test sequences (many tile boundaries), and the sequential per-alignment sweep.
# Sequential mode is operation-for-operation the reference: weights are bit-exact except where device exp() and glibc
# exp() round differently.  Tiled mode shifts F and B by a constant per tile, which changes last-ulp rounding: a
# weight floor(p * 1e7) may then differ by ONE unit (1e-7 in probability), typically at cells with p == 1 up to
# rounding; the (x, y) sets and their order stay identical.
MODES = {
    "tiles": dict(MARGIN_B200_MODE="tiles"),
    "tiles_short": dict(MARGIN_B200_MODE="tiles", MARGIN_B200_TILE_LEN="60", MARGIN_B200_WARMUP="128"),
    "sequential": dict(MARGIN_B200_MODE="sequential"),
}


@pytest.fixture(params=list(MODES))
def mode(request, monkeypatch):
    for k in ("MARGIN_B200_MODE", "MARGIN_B200_TILE_LEN", "MARGIN_B200_WARMUP"):
        monkeypatch.delenv(k, raising=False)
    for k, v in MODES[request.param].items():
        monkeypatch.setenv(k, v)
    return request.param


def _allowed_diffs(mode, ntot):
    return max(2, ntot // 100000) if mode == "sequential" else max(2, ntot // 20)


def _mk_batch(rng, n, max_len, rle=True, **kw):
    return [gen.random_case(rng, int(rng.integers(0, max_len)), rle=rle, **kw) for _ in range(n)]


@pytest.mark.parametrize("name", ["r94g360", "r94g344", "hifi", "no_rle"])
def test_golden_vectors_of_the_reference(golden_model, name, mode):
    pm = golden_model(name)
    al = PairHmmAligner(pm.models, pm.p, 0)
    ndiff = ntot = 0
    for c in load_golden_cases(name):
        al.set_params(pm.p.copy(min_diags_between_traceback=c["min_diags"], dynamic_anchor_expansion=c["dynamic"]))
        item = (c["sx"], c["sy"], c["a"], c["slot"], c["rl"], c["rr"])
        got = al.posteriors_batch([item], True)[0]
        ndiff += compare_lists(got, c["lists"], tol_w=1)
        ntot += sum(len(x) for x in got)
        fwd = al.forward_batch([(c["sx"][:60], c["sy"][:60], np.zeros((0, 3), np.int32), c["slot"], c["rl"], c["rr"])])
        assert fwd[0] == c["fwd"]  # FP64 forward sweep is bit-exact
        kl, kr = al.bands([item])[0]
        assert np.array_equal(kl, c["kl"]) and np.array_equal(kr, c["kr"])
    assert ndiff <= _allowed_diffs(mode, ntot)
    al.close()


@pytest.mark.parametrize("force_group", ["", "8", "32"])
def test_random_batch_vs_oracle(oracle, golden_model, force_group, monkeypatch, mode):
    """Mixed batch: empty and 1-symbol sequences, no anchors (full matrix => wide diagonals), ragged ends, N symbols,
    clipped run lengths, both strands, several tracebacks.  MARGIN_B200_FORCE_GROUP exercises the wide-diagonal
    fallback (diagonals wider than the lane group) and the 32-lane path on every job."""
    if force_group:
        monkeypatch.setenv("MARGIN_B200_FORCE_GROUP", force_group)
    pm = golden_model("r94g360")
    rng = np.random.default_rng(17)
    items = _mk_batch(rng, 96, 420)
    items[0] = (np.zeros(0, np.uint16), np.zeros(0, np.uint16), np.zeros((0, 3), np.int32), 0, 0, 0)
    items[1] = (items[1][0][:1], np.zeros(0, np.uint16), np.zeros((0, 3), np.int32), 1, 0, 0)
    items[2] = (np.zeros(0, np.uint16), items[2][1][:3], np.zeros((0, 3), np.int32), 0, 1, 1)
    for md, dyn, wi in ((5000, 1, 1), (90, 1, 1), (64, 0, 1), (130, 1, 0)):
        p = pm.p.copy(min_diags_between_traceback=md, dynamic_anchor_expansion=dyn)
        al = PairHmmAligner(pm.models, p, 0)
        res = al.posteriors_batch(items, bool(wi))
        ndiff = ntot = 0
        for i, it in enumerate(items):
            want = oracle.posteriors(pm.models[it[3]], p, it[0], it[1], it[2], it[4], it[5], wi)
            ndiff += compare_lists(res[i], want, tol_w=1)
            ntot += sum(len(x) for x in want)
        assert ndiff <= _allowed_diffs(mode, ntot)
        assert res.stats["kernel_launches"] > 0 and res.stats["band_cells"] > 0
        al.close()


def test_long_alignment_many_tracebacks(oracle, golden_model, mode):
    pm = golden_model("r94g360")
    rng = np.random.default_rng(23)
    items = [gen.random_case(rng, 9000, anchor_keep=0.9, expansions=(4,), ragged=False) for _ in range(3)]
    al = PairHmmAligner(pm.models, pm.p, 0)  # shipped parameters: traceback every >= 5000 diagonals
    res = al.posteriors_batch(items, True)
    assert res.stats["segments"] >= 3 * 3
    for i, it in enumerate(items):
        want = oracle.posteriors(pm.models[it[3]], pm.p, it[0], it[1], it[2], it[4], it[5], 1)
        compare_lists(res[i], want, tol_w=1)
    al.close()


def test_forward_only_batch(oracle, golden_model):
    """computeForwardProbability on bubble-graph-sized problems (median 11x11) and a few larger ones: bit-exact."""
    pm = golden_model("r94g360")
    rng = np.random.default_rng(29)
    items = []
    for i in range(400):
        sx = gen.random_symbols(rng, int(rng.integers(0, 44)), True, 0.02)
        sy = gen.random_symbols(rng, int(rng.integers(0, 44)), True, 0.02)
        items.append((sx, sy, np.zeros((0, 3), np.int32), int(rng.integers(0, 2)), int(rng.random() < 0.2),
                      int(rng.random() < 0.2)))
    al = PairHmmAligner(pm.models, pm.p, 0)
    got = al.forward_batch(items)
    for i, it in enumerate(items):
        want = oracle.forward(pm.models[it[3]], pm.p, it[0], it[1], (), it[4], it[5])
        assert got[i] == want, (i, got[i], want)
    al.close()


def test_known_answers_on_gpu():
    """test_diagonalDPCalculations (tests/pairwiseAlignerTest.c:257-342) and test_bands (:64-127) through the GPU."""
    m = model.default_nucleotide_model()
    p = abi.Params.defaults().copy(threshold=0.2)
    al = PairHmmAligner((m, m), p, 0)
    sx, sy = symbols.bases_to_symbols("AGCG"), symbols.bases_to_symbols("AGTTCG")
    pairs = al.get_aligned_pairs(0, sx, sy)
    assert sorted((int(t["x"]), int(t["y"])) for t in pairs) == [(0, 0), (1, 1), (2, 4), (3, 5)]
    al.set_params(abi.Params.defaults().copy(diagonal_expansion=2, dynamic_anchor_expansion=0))
    kl, kr = al.bands([(np.zeros(6, np.uint16), np.zeros(5, np.uint16), [(1, 0, 2), (2, 1, 2), (3, 3, 2)], 0, 0, 0)])[0]
    assert list(zip(kl.tolist(), kr.tolist())) == [(0, 0), (-1, 1), (-2, 2), (-1, 3), (-2, 4), (-1, 3), (-2, 4), (-3, 3),
                                                    (-2, 2), (-1, 3), (0, 2), (1, 1)]
    al.close()


def test_large_gap_split_vs_oracle(oracle, golden_model, mode):
    """Multi-block alignments (the unpatched reference is broken here; the oracle implements the
    intended semantics, validated in the survey against the one-line-patched reference)."""
    pm = golden_model("r94g360")
    rng = np.random.default_rng(31)
    items = [gen.random_case(rng, int(rng.integers(150, 500)), anchor_keep=0.04) for _ in range(40)]
    p = pm.p.copy(split_matrix_bigger_than_this=60 * 60, min_diags_between_traceback=100)
    al = PairHmmAligner(pm.models, p, 0)
    res = al.posteriors_batch(items, True)
    assert res.stats["jobs"] > len(items)
    for i, it in enumerate(items):
        want = oracle.posteriors(pm.models[it[3]], p, it[0], it[1], it[2], it[4], it[5], 1)
        compare_lists(res[i], want, tol_w=1)
    al.close()


def test_reference_itself_on_this_box(reference, golden_model, mode):
    """Where oracle/_ref travelled with the snapshot: GPU vs the unmodified reference, same process."""
    pm = golden_model("r94g360")
    sms = [reference.sm_from_model(pm.models[s]) for s in (0, 1)]
    rng = np.random.default_rng(37)
    items = _mk_batch(rng, 32, 600)
    p = pm.p.copy(min_diags_between_traceback=200)
    al = PairHmmAligner(pm.models, p, 0)
    res = al.posteriors_batch(items, True)
    for i, it in enumerate(items):
        want = reference.posteriors(sms[it[3]], p, it[0], it[1], it[2], it[4], it[5], 1)
        compare_lists(res[i], want, tol_w=1)
    al.close()


def test_errors_are_loud(golden_model):
    pm = golden_model("r94g360")
    al = PairHmmAligner(pm.models, pm.p, 0)
    sx = gen.random_symbols(np.random.default_rng(1), 20)
    with pytest.raises(AlignerError):  # anchors must increase strictly (asserts at pairwiseAligner.c:150-158)
        al.posteriors_batch([(sx, sx, [(5, 5, 4), (4, 6, 4)], 0, 0, 0)])
    with pytest.raises(AlignerError):  # odd expansion
        al.posteriors_batch([(sx, sx, [(5, 5, 3)], 0, 0, 0)])
    with pytest.raises(AlignerError):  # traceBackDiagonals + 1 < minDiagsBetweenTraceBack (:718)
        al.set_params(pm.p.copy(min_diags_between_traceback=30))
    assert len(al.posteriors_batch([])) == 0
    bad = sx.copy()
    bad[3] = (51 << 8) | 1  # a run length the repeat-count model cannot hold (reported by the sweep)
    with pytest.raises(AlignerError, match="run length"):
        al.posteriors_batch([(bad, sx, [(5, 5, 4)], 0, 0, 0)])
    with pytest.raises(AlignerError, match="run length"):
        al.forward_batch([(sx, bad, np.zeros((0, 3), np.int32), 0, 0, 0)])
    assert len(al.posteriors_batch([(sx, sx, [(5, 5, 4)], 0, 0, 0)])) == 1  # the context is usable afterwards
    al.close()


@pytest.mark.gpu
@pytest.mark.parametrize("k", [2, 3, 5])
def test_pipelined_batch_matches_the_plain_call(oracle, golden_model, monkeypatch, k):
    """margin_b200_posteriors_batch cut into k overlapping runs (sub-contexts, own streams) returns the same lists in
    the same contiguous layout as the unpipelined call, and both match the oracle."""
    from margin_b200.aligner import PairHmmAligner
    pm = golden_model("r94g360")
    p = pm.p.copy(min_diags_between_traceback=300)
    rng = np.random.default_rng(77)
    items = [gen.random_case(rng, int(rng.integers(1, 1500))) for _ in range(23)]
    items.insert(4, (np.zeros(0, np.uint16), items[0][1], np.zeros((0, 3), np.int32), 0, 0, 0))  # an empty sequence
    monkeypatch.setenv("MARGIN_B200_PIPELINE", "1")
    plain = PairHmmAligner(pm.models, p, device=0)
    want = plain.posteriors_batch(items, True)
    plain.close()
    monkeypatch.setenv("MARGIN_B200_PIPELINE", str(k))
    monkeypatch.setenv("MARGIN_B200_PIPELINE_MIN_BYTES", "0")
    monkeypatch.setenv("MARGIN_B200_PIPELINE_FORCE", "1")  # other tests' contexts may be alive on the device
    al = PairHmmAligner(pm.models, p, device=0)
    for _ in range(2):  # the second call reuses the pinned lists sized by the first
        got = al.posteriors_batch(items, True)
        assert len(got) == len(items)
        for i in range(len(items)):
            compare_lists(got[i], want[i], tol_w=1)  # tile cuts differ between the two calls
        assert got.stats["band_cells"] == want.stats["band_cells"]
    for i in (0, 4, 11, 23):
        it = items[i]
        ref = oracle.posteriors(pm.models[it[3]], p, it[0], it[1], it[2], it[4], it[5], 1)
        compare_lists(got[i], ref, tol_w=1)
    al.close()


@pytest.mark.gpu
@pytest.mark.parametrize("pipeline", ["1", "3"])
def test_compact_anchor_words_and_packed_tuples(golden_model, monkeypatch, pipeline):
    """The compact boundary formats (4-byte anchor steps in, 8-byte tuple words out; include/margin_b200.h) give
    exactly the lists of the plain formats, through the plain and the pipelined call."""
    pm = golden_model("r94g360")
    p = pm.p.copy(min_diags_between_traceback=300, dynamic_anchor_expansion=1)
    rng = np.random.default_rng(5)
    items = [gen.random_case(rng, int(rng.integers(1, 1200))) for _ in range(17)]
    monkeypatch.setenv("MARGIN_B200_PIPELINE", pipeline)
    monkeypatch.setenv("MARGIN_B200_PIPELINE_MIN_BYTES", "0")
    monkeypatch.setenv("MARGIN_B200_PIPELINE_FORCE", "1")
    al = PairHmmAligner(pm.models, p, device=0)
    want = al.posteriors_batch(abi.Batch(items), True)
    compact = abi.Batch(items, compact=True)
    assert compact.compact and compact.h2d_bytes < abi.Batch(items).h2d_bytes
    al.set_pair_format(abi.PAIRS_PACKED64)
    got = al.posteriors_batch(compact, True)
    for i in range(len(items)):
        compare_lists(got[i], want[i], tol_w=0)
    # an anchor step that does not fit a word: the host keeps the triples
    far = (items[0][0], items[0][1], np.array([[0, 0, 4]], np.int32), 0, 0, 0)
    big = np.zeros(6000, np.uint16)
    wide = abi.Batch([(big, big, np.array([[10, 10, 2], [5000, 5000, 2]], np.int32), 0, 0, 0), far], compact=True)
    assert not wide.compact
    # gap lists carry -1 coordinates through the packed words
    assert min(int(g[l]["x"].min(initial=0)) for g in got.lists for l in (1, 2)) >= -1
    al.close()


def test_pack_helpers_roundtrip():
    words = np.array([0, (1 << 24) | 5, 10000000 | (123457 << 24) | (7 << 44)], np.uint64)
    r = abi.unpack_pairs(words)
    assert r["w"].tolist() == [0, 5, 10000000] and r["x"].tolist() == [-1, 0, 123456] and r["y"].tolist() == [-1, -1, 6]
    a = np.array([[0, 0, 4], [3, 1, 0], [7, 9, 510]], np.int32)
    d = abi.pack_anchors_delta(a, np.array([0, 1, 3]))
    assert d.tolist() == [1 | (1 << 12) | (2 << 24), 4 | (2 << 12), 4 | (8 << 12) | (255 << 24)]


@pytest.mark.gpu
@pytest.mark.parametrize("tile_len", ["0", "200"])
def test_long_indel_corridors_keep_tile_parity(oracle, golden_model, monkeypatch, tile_len):
    """A long deletion or insertion between dense anchors makes the band a corridor in which the DP forgets nothing:
    a tile edge inside or just after it must take its warm-up from before the corridor (band_chunk_kernel reports
    the anchor-free stretches, build_tiles keeps the warm-up out of them).  Found on the bundled ONT region (a
    611-base deletion); a fixed warm-up of 128 diagonals emitted ~500 wrong tuples there."""
    from margin_b200.aligner import PairHmmAligner
    pm = golden_model("r94g360")
    rng = np.random.default_rng(611)
    items = []
    for length, deletion in ((611, True), (611, False), (150, True), (40, False), (1500, True)):
        items.append(gen.long_indel_case(rng, 5000, int(rng.integers(1500, 3000)), length, deletion))
    two = gen.long_indel_case(rng, 6000, 2000, 300, True)  # two corridors close to each other
    sx, sy, a = two[0], two[1], two[2]
    keep = ~((a[:, 0] >= 2400) & (a[:, 0] < 2650))  # drop the anchors of 250 columns: a second anchor-free stretch
    items.append((sx, sy, a[keep], 0, 0, 0))
    if tile_len != "0":
        monkeypatch.setenv("MARGIN_B200_TILE_LEN", tile_len)  # many tile edges inside every corridor
    monkeypatch.setenv("MARGIN_B200_PIPELINE", "1")
    al = PairHmmAligner(pm.models, pm.p, device=0)
    got = al.posteriors_batch(items, True)
    assert got.stats["tiles"] > len(items)
    for i, it in enumerate(items):
        ref = oracle.posteriors(pm.models[it[3]], pm.p, it[0], it[1], it[2], it[4], it[5], 1)
        compare_lists(got[i], ref, tol_w=1)  # every seam is guarded: one unit of 1e-7, as everywhere else
    al.close()


@pytest.mark.gpu
def test_wide_spots_cost_a_tile_not_the_read(oracle, golden_model, monkeypatch):
    """Per-tile classification: a long read whose band opens wider than 32 cells in a few places (anchors missing over
    60-200 columns) keeps its narrow stretches on the lane-group tile kernels; only the tiles that contain a wide
    diagonal take the general path (private F), and no job falls back to the sequential sweep.  Strict parity."""
    pm = golden_model("r94g360")
    rng = np.random.default_rng(108)
    items = [gen.gapped_anchor_case(rng, 6000, gap) for gap in (60, 110, 200)]
    items.append(gen.gapped_anchor_case(rng, 5000, 0, n_gaps=0))  # all narrow
    sx = gen.random_symbols(rng, 300)
    sy, _ = gen.mutate(rng, sx)
    items.append((sx, sy, np.zeros((0, 3), np.int32), 0, 0, 0))  # no anchors at all: one wide tile per traceback
    monkeypatch.setenv("MARGIN_B200_PIPELINE", "1")
    al = PairHmmAligner(pm.models, pm.p, device=0)
    got = al.posteriors_batch(items, True)
    st = got.stats
    assert st["wide_cells"] > 0 and st["sequential_cells"] == 0 and st["tiles"] > 20, st
    assert st["wide_cells"] < 0.8 * st["band_cells"], st
    assert st["guard_seams"] > 0 and st["guard_failures"] == 0, st
    for i, it in enumerate(items):
        ref = oracle.posteriors(pm.models[it[3]], pm.p, it[0], it[1], it[2], it[4], it[5], 1)
        compare_lists(got[i], ref, tol_w=1)
    al.close()


@pytest.mark.gpu
def test_seam_guard_catches_an_unconverged_warmup(oracle, golden_model, monkeypatch):
    """With a warm-up far too short for the recurrences to forget their flat start (8 diagonals) the tile edges disagree:
    the guard must flag them, the tiles concerned run again with a longer warm-up (or their jobs sequentially) inside the
    call, and the lists are the oracle's.
    With the guard off the same setting produces wrong weights (which is what the guard is for)."""
    pm = golden_model("r94g360")
    rng = np.random.default_rng(4242)
    items = [gen.random_case(rng, 2500, anchor_keep=0.9, expansions=(4,), ragged=False) for _ in range(4)]
    refs = [oracle.posteriors(pm.models[it[3]], pm.p, it[0], it[1], it[2], it[4], it[5], 1) for it in items]
    monkeypatch.setenv("MARGIN_B200_PIPELINE", "1")
    monkeypatch.setenv("MARGIN_B200_TILE_LEN", "200")
    monkeypatch.setenv("MARGIN_B200_WARMUP", "8")
    al = PairHmmAligner(pm.models, pm.p, device=0)
    got = al.posteriors_batch(items, True)
    assert got.stats["guard_failures"] > 0, got.stats  # repaired in place (longer warm-up) or swept sequentially
    for i in range(len(items)):
        compare_lists(got[i], refs[i], tol_w=1)
    al.close()
    monkeypatch.setenv("MARGIN_B200_GUARD", "0")
    al = PairHmmAligner(pm.models, pm.p, device=0)
    bad = al.posteriors_batch(items, True)
    assert bad.stats["guard_seams"] == 0
    worst = 0
    for i in range(len(items)):
        for l in range(3):
            if len(bad[i][l]) != len(refs[i][l]) or not np.array_equal(bad[i][l]["x"], refs[i][l]["x"]):
                worst = 10 ** 9
            else:
                worst = max(worst, int(np.abs(bad[i][l]["w"].astype(np.int64) - refs[i][l]["w"]).max(initial=0)))
    assert worst > 100, "an 8-diagonal warm-up should not have converged (worst weight error %d)" % worst
    al.close()


@pytest.mark.gpu
@pytest.mark.parametrize("tile_len", ["0", "200"])
def test_adversarial_seams_stay_exact(oracle, golden_model, monkeypatch, tile_len):
    """Inputs chosen against the forgetting assumption: anchor-free corridors of 16-40 diagonals (below the
    32-diagonal threshold of the corridor logic), tandem repeats with sparse wide anchors, bands just under 32 cells.
    Whatever the tiler does, the lists must be the oracle's to one unit of 1e-7: either the seams converge or the guard
    sends the job to the sequential sweep."""
    pm = golden_model("r94g360")
    rng = np.random.default_rng(1632)
    items = []
    for length in (16, 20, 24, 28, 31, 33, 36, 40):
        items.append(gen.long_indel_case(rng, 3000, int(rng.integers(800, 2200)), length, bool(length % 8 == 0)))
    for period in (1, 2, 3, 5):
        items.append(gen.low_complexity_case(rng, 3000, period))
    items.append(gen.random_case(rng, 3000, anchor_keep=0.5, expansions=(28, 30), ragged=False))  # bands just under 32
    if tile_len != "0":
        monkeypatch.setenv("MARGIN_B200_TILE_LEN", tile_len)
    monkeypatch.setenv("MARGIN_B200_PIPELINE", "1")
    al = PairHmmAligner(pm.models, pm.p, device=0)
    got = al.posteriors_batch(items, True)
    print("adversarial seams (tile_len %s): %s" % (tile_len, got.stats))
    for i, it in enumerate(items):
        ref = oracle.posteriors(pm.models[it[3]], pm.p, it[0], it[1], it[2], it[4], it[5], 1)
        compare_lists(got[i], ref, tol_w=1)
    al.close()


@pytest.mark.gpu
def test_delta32_tuples_decode_to_the_plain_lists(golden_model, monkeypatch):
    """MARGIN_B200_PAIRS_DELTA32 (4-byte step records + escapes, include/margin_b200.h) decodes to exactly the lists of
    the plain format: narrow tiles, wide tiles, sequential jobs, split alignments, empty lists, -1 coordinates of the gap
    lists, match-only calls."""
    pm = golden_model("r94g360")
    rng = np.random.default_rng(32)
    items = [gen.random_case(rng, int(rng.integers(0, 1500))) for _ in range(40)]
    items += [gen.gapped_anchor_case(rng, 4000, 90), gen.long_indel_case(rng, 3000, 1200, 400, True)]
    items.insert(3, (np.zeros(0, np.uint16), np.zeros(0, np.uint16), np.zeros((0, 3), np.int32), 0, 0, 0))
    for md, split, wi in ((300, 3000 * 3000, 1), (5000, 80 * 80, 1), (150, 3000 * 3000, 0)):
        p = pm.p.copy(min_diags_between_traceback=md, split_matrix_bigger_than_this=split)
        al = PairHmmAligner(pm.models, p, device=0)
        want = al.posteriors_batch(items, bool(wi))
        al.set_pair_format(abi.PAIRS_DELTA32)
        got = al.posteriors_batch(items, bool(wi))
        assert sum(len(x) for g in got.lists for x in g) == sum(len(x) for g in want.lists for x in g) > 0
        for i in range(len(items)):
            compare_lists(got[i], want[i], tol_w=0)
        al.close()


@pytest.mark.gpu
def test_packed_symbols_and_anchor_runs_give_the_plain_results(golden_model):
    """The one-byte symbols and the run-length anchors (margin_b200_batch.packed_x/packed_y/anchors_runs,
    include/margin_b200.h) are expanded on the device to exactly the arrays of the plain call: identical lists,
    identical forward totals, in both tuple formats; out-of-range anchors inside a run are still reported."""
    pm = golden_model("r94g360")
    rng = np.random.default_rng(77)
    items = [gen.random_case(rng, int(rng.integers(1, 1500))) for _ in range(25)]
    items += [gen.gapped_anchor_case(rng, 4000, 90), gen.long_indel_case(rng, 3000, 1200, 400, True)]
    # one anchor per column over a long identical stretch: runs longer than a word's 255 followers, expansion changes
    sx = gen.random_symbols(rng, 1400, True, 0.0)
    dense = np.array([[i, i, 2 if i < 900 else 4] for i in range(3, 1390)], np.int32)
    items.append((sx, sx.copy(), dense, 1, 0, 0))
    items.insert(2, (np.zeros(0, np.uint16), np.zeros(0, np.uint16), np.zeros((0, 3), np.int32), 0, 0, 0))
    p = pm.p.copy(min_diags_between_traceback=300)
    al = PairHmmAligner(pm.models, p, device=0)
    plain, compact = abi.Batch(items), abi.Batch(items, compact=True)
    assert compact.packed and compact.runs and compact.h2d_bytes < 0.5 * plain.h2d_bytes
    assert len(compact.anchors_runs) < 0.5 * len(compact.anchors)
    for fmt in (abi.PAIRS_TRIPLES, abi.PAIRS_DELTA32):
        al.set_pair_format(fmt)
        want = al.posteriors_batch(plain, True)
        got = al.posteriors_batch(compact, True)
        assert sum(len(x) for g in want.lists for x in g) > 0
        for i in range(len(items)):
            compare_lists(got[i], want[i], tol_w=0)
    small = [(it[0][:40], it[1][:40], np.zeros((0, 3), np.int32), it[3], it[4], it[5]) for it in items]
    big = [(it[0], it[1], np.zeros((0, 3), np.int32), it[3], 0, 0) for it in items[:6]]
    for fw in (small, small + big):
        a, b = al.forward_batch(abi.Batch(fw)), al.forward_batch(abi.Batch(fw, compact=True))
        assert np.array_equal(np.asarray(a), np.asarray(b))
    # what does not fit stays on the wider forms
    far = abi.Batch([(np.zeros(3000, np.uint16), np.zeros(3000, np.uint16), np.array([[10, 10, 2], [2000, 2100, 2]], np.int32),
                      0, 0, 0)], compact=True)
    assert far.compact and not far.runs and far.packed
    assert len(al.posteriors_batch(far, False)) == 1
    # a run that walks off the end of a sequence
    off = abi.Batch([(sx[:100], sx[:100], np.array([[i, i, 2] for i in range(90, 104)], np.int32), 0, 0, 0)], compact=True)
    assert off.runs
    with pytest.raises(AlignerError):
        al.posteriors_batch(off, False)
    assert len(al.posteriors_batch(plain, False)) == len(items)  # the context is usable afterwards
    al.close()


@pytest.mark.gpu
def test_pinned_blocks_are_recycled_not_returned_to_the_driver():
    """margin_b200_pinned_free hands a block to the process-wide cache (cudaFreeHost would synchronise the device under
    every other host thread's feet): the next request of the same size class gets the same memory."""
    lib = abi.load_library()
    a = lib.margin_b200_pinned_alloc(3 << 20)  # classes are 2^k x {1, 1.25, 1.5, 1.75}; 64 bytes of header
    assert a and a % 64 == 0
    np.ctypeslib.as_array((abi.C.c_uint8 * (3 << 20)).from_address(a))[:] = 7  # writable over its whole length
    lib.margin_b200_pinned_free(a)
    again = []  # earlier tests may have left blocks of this class in the cache: a comes back once those are handed out
    for _ in range(64):
        again.append(lib.margin_b200_pinned_alloc((3 << 20) + 4096))  # same size class (3 MB + header -> 3.5 MB)
        if again[-1] == a:
            break
    assert a in again and len(set(again)) == len(again)
    for p in again:
        lib.margin_b200_pinned_free(p)
    z = lib.margin_b200_pinned_alloc(0)  # a zero-byte request still returns memory
    assert z
    lib.margin_b200_pinned_free(z)
