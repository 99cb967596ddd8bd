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
