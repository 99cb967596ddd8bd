"""Synthetic workloads of BASELINE.json's configs, as alignment batches for the hot path.

A chunk is what ``poa_realign`` (impl/poa.c:668-716) sees: one draft (reference) sequence, a pile of reads with
their BAM-derived anchors, everything run-length encoded.  The generator follows the reference's input side:

* error model: a parametrised restatement of ``evolveSequence`` (impl/randomSequences.c:49-72) — per-base
  substitution / insertion / deletion rates, with a fraction of the indels placed in homopolymers (run length +-1)
  so that RLE behaves like ONT data;
* anchors: one ``(refPos, readPos, diagonalExpansion)`` per match column of the true edit path, exactly what
  ``convertToReadsAndAlignmentsWithFiltered`` emits per CIGAR match (impl/htsIntegration.c:722-729);
* ``runLengthEncodeAlignment`` (impl/rle.c:232-253), the crop + anchor shift of
  ``getAlignedPairsWithIndelsCroppingReference`` (impl/poa.c:612-666) and the Symbol strings
  (impl/stateMachine.c:133-146) come from :mod:`margin_b200.symbols`.

Everything is seeded (numpy Generator), so the same batch can be rebuilt on any box.
"""
import numpy as np

from . import symbols

ONT = dict(sub=0.03, ins=0.03, dele=0.04, hp_frac=0.6)   # ~10 % error
HIFI = dict(sub=0.002, ins=0.004, dele=0.004, hp_frac=0.6)  # ~1 % error


def _hp_positions(seq):
    """Positions that sit in a homopolymer run of length >= 2."""
    if len(seq) < 2:
        return np.zeros(0, np.int64)
    same = seq[1:] == seq[:-1]
    m = np.zeros(len(seq), bool)
    m[1:] |= same
    m[:-1] |= same
    return np.flatnonzero(m)


def mutate(rng, seq, sub, ins, dele, hp_frac):
    """Returns (out, pos) where pos[t] is the index in ``out`` of truth base t, or -1 if it was deleted."""
    n = len(seq)
    seq = np.asarray(seq, dtype=np.uint16)
    hp = _hp_positions(seq)

    def pick(rate):
        k = int(rng.binomial(n, rate)) if n else 0
        k_hp = int(round(k * hp_frac)) if len(hp) else 0
        a = hp[rng.integers(0, len(hp), k_hp)] if k_hp else np.zeros(0, np.int64)
        b = rng.integers(0, n, k - k_hp) if k - k_hp > 0 else np.zeros(0, np.int64)
        return a, b

    keep = np.ones(n, bool)
    d_hp, d_any = pick(dele)
    keep[d_hp] = False
    keep[d_any] = False
    base = seq.copy()
    s_mask = rng.random(n) < sub
    base[s_mask] = (base[s_mask] + rng.integers(1, 4, int(s_mask.sum()))) % 4
    # insertions before position t: homopolymer ones repeat the base at t, the others are random
    i_hp, i_any = pick(ins)
    n_ins = np.zeros(n, np.int64)
    np.add.at(n_ins, i_hp, 1)
    np.add.at(n_ins, i_any, 1)
    ins_is_hp = np.zeros(n, bool)
    ins_is_hp[i_hp] = True
    out_count = n_ins + keep
    starts = np.concatenate([[0], np.cumsum(out_count)])
    total = int(starts[-1])
    out = np.zeros(total, np.uint16)
    pos = np.where(keep, starts[:-1] + n_ins, -1)
    out[pos[keep]] = base[keep]
    # fill the inserted bases
    t_ins = np.repeat(np.arange(n), n_ins)
    if len(t_ins):
        off = np.arange(len(t_ins)) - np.repeat(np.cumsum(n_ins) - n_ins, n_ins)
        idx = starts[t_ins] + off
        rnd = rng.integers(0, 4, len(t_ins)).astype(np.uint16)
        out[idx] = np.where(ins_is_hp[t_ins], seq[t_ins], rnd)
    return out, pos


def _rle_anchors_fast(x2, y2, e):
    """runLengthEncodeAlignment's greedy filter (impl/rle.c:232-253) for monotone x2, y2."""
    xs, ys = x2.tolist(), y2.tolist()
    keep = []
    x = y = -1
    for i in range(len(xs)):
        if xs[i] > x and ys[i] > y:
            keep.append(i)
            x, y = xs[i], ys[i]
    k = np.array(keep, dtype=np.int64)
    out = np.empty((len(k), 3), np.int32)
    out[:, 0] = x2[k]
    out[:, 1] = y2[k]
    out[:, 2] = e
    return out


class Chunk:
    """One polishing chunk: RLE draft + reads + RLE-space anchors, ready to be turned into a Batch."""

    def __init__(self, draft_rle, reads):
        self.draft_rle = draft_rle
        self.reads = reads  # list of (RleString, forward_strand, anchors int32[k,3] in RLE coordinates)

    def items(self, polish_model):
        """The per-read aligner inputs poa_realign builds (impl/poa.c:682-704)."""
        out = []
        for rle, fwd, anc in self.reads:
            item, _ = symbols.cropped_alignment_item(self.draft_rle, rle, fwd, anc, polish_model)
            out.append(item)
        return out


def make_chunk(seed, contig_len=100000, depth=60, read_len=20000, err=ONT, draft_err=0.01, expansion=4):
    """BASELINE.json configs[1] with the defaults: 100 kb contig, 60x, 20 kb ONT-like reads at ~10 % error."""
    rng = np.random.default_rng(seed)
    truth = rng.integers(0, 4, contig_len).astype(np.uint16)
    d3 = draft_err / 3.0
    draft, dpos = mutate(rng, truth, d3, d3, d3, 0.6)
    draft_rle = symbols.RleString(draft)
    dmap = draft_rle.non_rle_to_rle_map()
    n_reads = max(1, int(round(depth * contig_len / float(read_len))))
    reads = []
    for _ in range(n_reads):
        rl = min(read_len, contig_len)
        s = int(rng.integers(0, contig_len - rl + 1))
        read, rpos = mutate(rng, truth[s:s + rl], **err)
        fwd = bool(rng.random() < 0.5)
        both = (rpos >= 0) & (dpos[s:s + rl] >= 0)
        ax = dpos[s:s + rl][both]
        ay = rpos[both]
        rrle = symbols.RleString(read)
        if len(rrle) == 0 or len(ax) == 0:
            continue
        rmap = rrle.non_rle_to_rle_map()
        anc = _rle_anchors_fast(dmap[ax], rmap[ay], expansion)
        reads.append((rrle, fwd, anc))
    return Chunk(draft_rle, reads)


def workload(name, polish_model, seed=42, scale=1.0):
    """Named workloads -> list of Batch items.  ``scale`` < 1 shrinks depth (bounded CPU samples, tests)."""
    if name == "c2_ont_100kb_60x_20kb":
        ch = make_chunk(seed, 100000, 60 * scale, 20000, ONT, 0.01, int(polish_model.p.diagonal_expansion))
    elif name == "c4_hifi_100kb_30x_15kb":
        ch = make_chunk(seed, 100000, 30 * scale, 15000, HIFI, 0.001, int(polish_model.p.diagonal_expansion))
    elif name == "tiny":
        ch = make_chunk(seed, 4000, 8 * scale, 1500, ONT, 0.01, int(polish_model.p.diagonal_expansion))
    else:
        raise ValueError("unknown workload %r" % name)
    return ch.items(polish_model)


def bubble_workload(seed, n, rle=True, max_len=43):
    """Forward-only DPs of the size poa_polish issues (lX, lY median 11, p95 21, max 43)."""
    rng = np.random.default_rng(seed)
    items = []
    lens = np.clip(rng.geometric(1.0 / 11.0, n), 1, max_len)
    for i in range(n):
        lx = int(lens[i])
        ly = int(max(1, min(max_len, lx + rng.integers(-2, 3))))
        sx = rng.integers(0, 4, lx).astype(np.uint16)
        sy = rng.integers(0, 4, ly).astype(np.uint16)
        if rle:
            sx |= (np.minimum(rng.geometric(0.7, lx), 50).astype(np.uint16) << 8)
            sy |= (np.minimum(rng.geometric(0.7, ly), 50).astype(np.uint16) << 8)
        items.append((sx, sy, np.zeros((0, 3), np.int32), int(rng.random() < 0.5), 0, 0))
    return items


# ------------------------------------------------------------------------------------- `margin polish` datasets
#
# SURVEY.md §8 D-2 "BAM mode": the same contig / read / error model as make_chunk, written as the files the
# reference CLI reads (polish.c:87: <BAM> <FASTA> <PARAMS>): a draft FASTA and a coordinate-sorted SAM whose CIGARs
# are the true edit paths, so that convertToReadsAndAlignmentsWithFiltered (impl/htsIntegration.c:557-891) derives
# the per-match anchors itself.  oracle/_ref/sam2bam turns the SAM into an indexed BAM (no samtools on the box).

_BASES = np.frombuffer(b"ACGT", dtype=np.uint8)


def _cigar(ax, ay, read_len):
    """CIGAR of a read from its match columns (ax: draft positions, ay: read positions, both strictly increasing):
    leading / trailing unaligned read bases are soft-clipped, gaps between match columns are D then I."""
    dx = np.diff(ax) - 1
    dy = np.diff(ay) - 1
    brk = np.flatnonzero((dx > 0) | (dy > 0))  # a gap sits between match column i and i+1
    parts = []
    if ay[0] > 0:
        parts.append("%dS" % ay[0])
    prev = 0
    for i, gx, gy in zip(brk.tolist(), dx[brk].tolist(), dy[brk].tolist()):
        parts.append("%dM" % (i + 1 - prev))
        if gx > 0:
            parts.append("%dD" % gx)
        if gy > 0:
            parts.append("%dI" % gy)
        prev = i + 1
    parts.append("%dM" % (len(ax) - prev))
    tail = read_len - 1 - int(ay[-1])
    if tail > 0:
        parts.append("%dS" % tail)
    return "".join(parts)


HET = dict(sub=0.001, ins=0.0001, dele=0.0001, hp_frac=0.6)  # second haplotype of BASELINE configs[4] (SURVEY D-2)


def write_polish_dataset(out_dir, seed=42, contig_len=100000, depth=60, read_len=20000, err=ONT, draft_err=0.01,
                         contig_name="contig_1", het=None, low_mapq_frac=0.0):
    """Writes <out_dir>/draft.fa and <out_dir>/reads.sam; returns (fasta_path, sam_path, n_reads, draft_len).
    ``het`` (e.g. HET): a second haplotype = the truth mutated at these rates; every read is drawn from one of the
    two haplotypes with probability 1/2 (the diploid workload of `margin polish --diploid`).
    ``low_mapq_frac``: this share of the reads gets MAPQ 1, below filterAlignmentsWithMapQBelowThisThreshold (5): margin
    sets them aside as filtered reads and, in a diploid run, assigns them to haplotypes afterwards
    (bubbleGraph_partitionFilteredReads, polish.c:776-821)."""
    import os
    rng = np.random.default_rng(seed)
    truth = rng.integers(0, 4, contig_len).astype(np.uint16)
    d3 = draft_err / 3.0
    draft, dpos = mutate(rng, truth, d3, d3, d3, 0.6)
    hap_b, bpos = mutate(rng, truth, **het) if het else (None, None)
    n_reads = max(1, int(round(depth * contig_len / float(read_len))))
    recs = []
    for k in range(n_reads):
        rl = min(read_len, contig_len)
        s = int(rng.integers(0, contig_len - rl + 1))
        if het and rng.random() < 0.5:
            # a read of the second haplotype: mutate its stretch of hap_b, then carry read positions back to the
            # truth coordinates of the window through bpos
            valid = bpos[s:s + rl] >= 0
            hb = bpos[s:s + rl][valid]
            if len(hb) < 2:
                continue
            read, rpos_b = mutate(rng, hap_b[hb[0]:hb[-1] + 1], **err)
            rpos = np.full(rl, -1, np.int64)
            rpos[valid] = rpos_b[hb - hb[0]]
        else:
            read, rpos = mutate(rng, truth[s:s + rl], **err)
        fwd = bool(rng.random() < 0.5)
        both = (rpos >= 0) & (dpos[s:s + rl] >= 0)
        ax = dpos[s:s + rl][both]
        ay = rpos[both]
        if len(ax) < 2 or len(read) == 0:
            continue
        seq = _BASES[read].tobytes().decode()
        # a reverse-strand read is stored reverse-complemented in a BAM, i.e. in reference orientation: the same
        # text, flag 16 (margin picks the reverse-strand state machine from the flag, htsIntegration.c:845-848)
        mapq = 1 if rng.random() < low_mapq_frac else 60
        recs.append((int(ax[0]), "read_%06d" % k, 0 if fwd else 16, _cigar(ax, ay, len(read)), seq, mapq))
    recs.sort(key=lambda r: (r[0], r[1]))
    os.makedirs(out_dir, exist_ok=True)
    fa = os.path.join(out_dir, "draft.fa")
    with open(fa, "w") as fh:
        text = _BASES[draft].tobytes().decode()
        fh.write(">%s\n" % contig_name)
        for i in range(0, len(text), 80):
            fh.write(text[i:i + 80] + "\n")
    with open(fa + ".fai", "w") as fh:  # written here: margin's chunk threads would otherwise race to build it
        fh.write("%s\t%d\t%d\t80\t81\n" % (contig_name, len(draft), len(contig_name) + 2))
    sam = os.path.join(out_dir, "reads.sam")
    with open(sam, "w") as fh:
        fh.write("@HD\tVN:1.6\tSO:coordinate\n@SQ\tSN:%s\tLN:%d\n" % (contig_name, len(draft)))
        for pos, name, flag, cigar, seq, mapq in recs:
            fh.write("%s\t%d\t%s\t%d\t%d\t%s\t*\t0\t0\t%s\t*\n" % (name, flag, contig_name, pos + 1, mapq, cigar, seq))
    return fa, sam, len(recs), len(draft)
