"""Symbol strings, run-length encoding and the caller-side contract of the aligner.

Vectorised (numpy) restatements of the small O(n) host preparations that sit
immediately before the hot path in the reference:

* ``convertNucleotideCharToSymbol`` / ``symbol_addRepeatCount`` /
  ``rleString_constructSymbolString``             impl/stateMachine.c:25-42, :116-146
* ``rleString_construct``                          impl/rle.c:7-40
* ``runLengthEncodeAlignment``                     impl/rle.c:232-253
* ``getAlignedPairsWithIndelsCroppingReference``   impl/poa.c:612-666 (crop + anchor shift)
"""
import numpy as np

_LUT = np.full(256, 4, dtype=np.uint16)
for _i, _c in enumerate("ACGT"):
    _LUT[ord(_c)] = _i
    _LUT[ord(_c.lower())] = _i


def bases_to_symbols(seq):
    """str/bytes of ACGTN -> uint16 symbols 0..4 (impl/stateMachine.c:25-42)."""
    if isinstance(seq, str):
        seq = seq.encode()
    return _LUT[np.frombuffer(seq, dtype=np.uint8)]


def symbols_to_bases(sym):
    return "".join("ACGTN"[min(int(s) & 255, 4)] for s in sym)


class RleString:
    """rleString_construct (impl/rle.c:7-40) on a symbol array (0..4)."""

    def __init__(self, expanded):
        expanded = np.asarray(expanded, dtype=np.uint16)
        n = len(expanded)
        if n == 0:
            self.bases = np.zeros(0, np.uint16)
            self.counts = np.zeros(0, np.int64)
            self.starts = np.zeros(0, np.int64)
        else:
            brk = np.flatnonzero(np.concatenate([[True], expanded[1:] != expanded[:-1]]))
            self.starts = brk.astype(np.int64)
            self.bases = expanded[brk]
            self.counts = np.diff(np.concatenate([brk, [n]])).astype(np.int64)
        self.non_rle_length = n

    def __len__(self):
        return len(self.bases)

    def non_rle_to_rle_map(self):
        """rleString_getNonRleToRleCoordinateMap: expanded position -> run index."""
        return np.repeat(np.arange(len(self.bases), dtype=np.int64), self.counts)

    def symbol_string(self, start, length, include_repeat_counts, max_rl_exclusive):
        """rleString_constructSymbolString (impl/stateMachine.c:133-146)."""
        s = self.bases[start:start + length].astype(np.uint16)
        if include_repeat_counts:
            rl = np.minimum(self.counts[start:start + length], max_rl_exclusive - 1).astype(np.uint16)
            s = (rl << 8) | s
        return np.ascontiguousarray(s)


def run_length_encode_alignment(anchors, map_x, map_y):
    """runLengthEncodeAlignment (impl/rle.c:232-253): anchors int[k,3] in expanded coordinates."""
    anchors = np.asarray(anchors, dtype=np.int64).reshape(-1, 3)
    if len(anchors) == 0:
        return np.zeros((0, 3), np.int32)
    x2 = map_x[anchors[:, 0]]
    y2 = map_y[anchors[:, 1]]
    # keep a pair only if it advances both coordinates past the last KEPT pair (greedy, sequential)
    out = []
    x = y = -1
    # fast path: runs of strictly increasing pairs dominate; fall back to the scalar rule per element
    for i in range(len(anchors)):
        if x2[i] > x and y2[i] > y:
            out.append((x2[i], y2[i], anchors[i, 2]))
            x, y = x2[i], y2[i]
    return np.array(out, dtype=np.int32).reshape(-1, 3)


def crop_reference(ref_len, read_len, anchors):
    """Cropping coordinates of impl/poa.c:621-633.  Returns (first_ref_position, end_ref_position)."""
    if len(anchors) > 0:
        first = int(anchors[0][0]) - int(anchors[0][1])
        first = max(first, 0)
        end = 1 + int(anchors[-1][0]) + (read_len - int(anchors[-1][1]))
        end = min(end, ref_len)
    else:
        first, end = 0, ref_len
    return first, end


def cropped_alignment_item(ref_rle, read_rle, forward_strand, anchors, polish_model):
    """Everything impl/poa.c:637-653 does before calling the aligner for one read.

    Returns (item, first_ref_position) where item is a Batch item
    (sX, sY, shifted anchors, model slot, ragged_left=0, ragged_right=0).
    """
    anchors = np.asarray(anchors, dtype=np.int32).reshape(-1, 3)
    first, end = crop_reference(len(ref_rle), len(read_rle), anchors)
    max_rl = polish_model.max_repeat_count_exclusive
    sx = ref_rle.symbol_string(first, end - first, polish_model.use_repeat_counts, max_rl)
    sy = read_rle.symbol_string(0, len(read_rle), polish_model.use_repeat_counts, max_rl)
    shifted = anchors.copy()
    shifted[:, 0] -= first
    return (sx, sy, shifted, 0 if forward_strand else 1, 0, 0), first
