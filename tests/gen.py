"""Seeded random alignment problems shared by the parity tests (own xorshift-free numpy RNG)."""
import numpy as np


def random_symbols(rng, n, rle=True, n_frac=0.02, max_rl=51):
    b = rng.integers(0, 4, n).astype(np.uint16)
    if n_frac > 0:
        b[rng.random(n) < n_frac] = 4
    if rle:
        rl = np.minimum(rng.geometric(0.55, n), max_rl - 1).astype(np.uint16)
        if n > 3:
            rl[rng.integers(0, n)] = max_rl - 1  # exercise the clipped maximum
        b = (rl << 8) | b
    return b


def mutate(rng, sx, sub=0.05, ins=0.04, dele=0.04, rle=True):
    """Returns (sy, anchors) where anchors are the true match columns (x, y)."""
    sy, anchors = [], []
    for x, s in enumerate(sx):
        r = rng.random()
        if r < dele:
            continue
        if r < dele + ins:
            sy.append(int(random_symbols(rng, 1, rle, 0.0)[0]))
        if rng.random() < sub:
            s2 = int(random_symbols(rng, 1, rle, 0.0)[0])
        else:
            s2 = int(s)
            if rle and rng.random() < 0.15:
                rlv = max(1, min(50, (s2 >> 8) + int(rng.integers(-1, 2))))
                s2 = (rlv << 8) | (s2 & 255)
        anchors.append((x, len(sy)))
        sy.append(s2)
    return np.array(sy, dtype=np.uint16), anchors


def random_case(rng, lx, rle=True, anchor_keep=0.7, expansions=(0, 2, 4, 10), ragged=True, n_frac=0.02):
    sx = random_symbols(rng, lx, rle, n_frac)
    sy, true = mutate(rng, sx, rle=rle)
    anchors = [(x, y, int(rng.choice(expansions))) for (x, y) in true if rng.random() < anchor_keep]
    if rng.random() < 0.15:
        anchors = []
    rl = int(ragged and rng.random() < 0.3)
    rr = int(ragged and rng.random() < 0.3)
    slot = int(rng.random() < 0.5)
    return sx, sy, np.array(anchors, dtype=np.int32).reshape(-1, 3), slot, rl, rr


def long_indel_case(rng, lx, at, length, deletion=True, expansion=4, rle=True):
    """Dense anchors (one per match column, as the BAM walk of htsIntegration.c:722-729 emits them) around one long
    deletion (reference bases missing from the read) or insertion: between the two anchors flanking it the band is a
    corridor of constant width along one sequence, `length` anti-diagonals long."""
    sx = random_symbols(rng, lx, rle, 0.0)
    if deletion:
        left, right = sx[:at], sx[at + length:]
        sy0 = np.concatenate([left, right])
        xs = np.concatenate([np.arange(at), np.arange(at + length, lx)])
    else:
        sy0 = np.concatenate([sx[:at], random_symbols(rng, length, rle, 0.0), sx[at:]])
        xs = np.arange(lx)
    ys = np.arange(len(sy0)) if deletion else np.concatenate([np.arange(at), np.arange(at + length, at + length + lx - at)])
    sy = sy0.copy()
    flip = rng.random(len(sy)) < 0.05  # substitutions keep their anchors (CIGAR M covers mismatches)
    sy[flip] = random_symbols(rng, int(flip.sum()), rle, 0.0)
    anchors = np.stack([xs, ys, np.full(len(xs), expansion)], axis=1).astype(np.int32)
    return sx, sy.astype(np.uint16), anchors, 0, 0, 0


def low_complexity_case(rng, lx, period=2, anchor_keep=0.15, expansion=10, rle=True):
    """A tandem repeat (unit of `period` symbols) with a few unique islands: inside the repeat every shift by a multiple
    of the period aligns equally well, the posterior spreads over several diagonals and the recurrences forget their
    start slowly.  Sparse anchors with a wide expansion keep the band open (round-2 style anchors)."""
    unit = random_symbols(rng, period, rle, 0.0)
    sx = np.tile(unit, lx // period + 1)[:lx].copy()
    for _ in range(max(1, lx // 700)):  # unique islands
        at = int(rng.integers(0, max(1, lx - 40)))
        sx[at:at + 30] = random_symbols(rng, len(sx[at:at + 30]), rle, 0.0)
    sy, true = mutate(rng, sx, sub=0.03, ins=0.03, dele=0.03, rle=rle)
    anchors = [(x, y, expansion) for (x, y) in true if rng.random() < anchor_keep]
    return sx, sy, np.array(anchors, dtype=np.int32).reshape(-1, 3), 0, 0, 0


def gapped_anchor_case(rng, lx, gap_cols, n_gaps=3, expansion=4, rle=True):
    """Dense anchors (one per match column) with `n_gaps` stretches of `gap_cols` columns whose anchors are dropped: the
    band opens into a diamond about gap_cols wide there (a wide tile when gap_cols > 32) and is narrow elsewhere."""
    sx = random_symbols(rng, lx, rle, 0.0)
    sy, true = mutate(rng, sx, sub=0.04, ins=0.03, dele=0.03, rle=rle)
    a = np.array([(x, y, expansion) for (x, y) in true], dtype=np.int32).reshape(-1, 3)
    keep = np.ones(len(a), bool)
    for _ in range(n_gaps):
        at = int(rng.integers(lx // 10, max(lx // 10 + 1, lx - gap_cols - lx // 10)))
        keep &= ~((a[:, 0] >= at) & (a[:, 0] < at + gap_cols))
    return sx, sy, a[keep], int(rng.random() < 0.5), 0, 0


# ------------------------------------------------------------------------------------------ the parity corpus
#
# >= 10 k fixed-seed alignments.  Case i is a pure function of i (its own Generator), so the
# corpus needs no input files: tests/golden/make_corpus.py runs the UNMODIFIED reference on every case here and commits
# only digests (tests/golden/corpus_digests.npz); the GPU test regenerates the inputs on the box.

CORPUS_SIZE = 10240
CORPUS_PARAM_SETS = ((5000, 1), (100, 1), (150, 1), (5000, 0), (130, 0))  # (minDiagsBetweenTraceBack, dynamicAnchorExpansion)


def corpus_case(i):
    """-> (item, param_set index).  Mix: short random problems with every edge the generator knows (empty sequences, no
    anchors, ragged ends, N, clipped run lengths), wide spots, long indels, tandem repeats, and alignments of more
    than 5000 diagonals at the shipped parameters (several tracebacks)."""
    rng = np.random.default_rng([20260105, i])
    kind = i % 20
    if kind < 13:
        item = random_case(rng, int(rng.integers(0, 700)))
        ps = int(rng.integers(0, len(CORPUS_PARAM_SETS)))
    elif kind < 15:
        item = gapped_anchor_case(rng, int(rng.integers(1200, 2600)), int(rng.integers(36, 120)), n_gaps=int(rng.integers(1, 4)))
        ps = int(rng.choice([0, 1]))
    elif kind == 15:
        item = long_indel_case(rng, int(rng.integers(1500, 3000)), int(rng.integers(300, 1000)), int(rng.integers(12, 700)),
                               bool(rng.random() < 0.5))
        ps = 0
    elif kind == 16:
        item = low_complexity_case(rng, int(rng.integers(800, 2400)), int(rng.integers(1, 6)))
        ps = int(rng.choice([0, 2]))
    elif kind == 17:
        item = gapped_anchor_case(rng, int(rng.integers(4500, 7000)), 0, n_gaps=0)  # > 5000 diagonals: several tracebacks
        ps = 0
    elif kind == 18:
        lx, ly = int(rng.integers(0, 4)), int(rng.integers(0, 4))
        item = (random_symbols(rng, lx), random_symbols(rng, ly), np.zeros((0, 3), np.int32), int(rng.random() < 0.5),
                int(rng.random() < 0.5), int(rng.random() < 0.5))
        ps = int(rng.integers(0, len(CORPUS_PARAM_SETS)))
    else:
        sx = random_symbols(rng, int(rng.integers(20, 260)))
        sy, _ = mutate(rng, sx)
        item = (sx, sy, np.zeros((0, 3), np.int32), int(rng.random() < 0.5), int(rng.random() < 0.3), int(rng.random() < 0.3))
        ps = int(rng.integers(0, len(CORPUS_PARAM_SETS)))
    return item, ps


def list_digest(lst):
    """(count, order-sensitive hash of the (x, y) sequence, sum of weights) of one tuple list."""
    n = len(lst)
    if n == 0:
        return 0, 0, 0
    x = lst["x"].astype(np.int64).astype(np.uint64)
    y = lst["y"].astype(np.int64).astype(np.uint64)
    k = np.arange(1, n + 1, dtype=np.uint64)
    with np.errstate(over="ignore"):
        h = ((x * np.uint64(1000003) + y + np.uint64(7)) * (k * np.uint64(0x9E3779B97F4A7C15) + np.uint64(1))).sum(dtype=np.uint64)
    return n, int(h), int(lst["w"].astype(np.int64).sum())
