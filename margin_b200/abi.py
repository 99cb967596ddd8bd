"""ctypes mirror of include/margin_b200.h and the loader of the CUDA library.

There is no CPU fallback in this package: :func:`load_library` raises if
``libmargin_b200.so`` has not been built (``python -c 'import __graft_entry__ as g; g.build()'``)
and the library itself fails every compute call when no CUDA device is usable.
"""
import ctypes as C
import os

import numpy as np

MAX_RL = 51  # MAXIMUM_REPEAT_LENGTH, inc/margin.h:133
PROB_1 = 10000000  # PAIR_ALIGNMENT_PROB_1, inc/pairwiseAligner.h:26

T_MATCH_CONTINUE, T_MATCH_FROM_GAP_X, T_MATCH_FROM_GAP_Y = 0, 1, 2
T_GAP_OPEN_X, T_GAP_EXTEND_X, T_GAP_SWITCH_TO_X = 3, 4, 5
T_GAP_OPEN_Y, T_GAP_EXTEND_Y, T_GAP_SWITCH_TO_Y = 6, 7, 8


class Model(C.Structure):
    """margin_b200_model: one flattened StateMachine3 + emissions (one strand)."""

    _fields_ = [
        ("t", C.c_double * 9),
        ("match", C.c_double * 16),
        ("gap_x", C.c_double * 4),
        ("gap_y", C.c_double * 4),
        ("rle", C.c_int32),
        ("max_rl", C.c_int32),
        ("rsm", C.c_double * (4 * MAX_RL * MAX_RL)),
    ]

    def same_as(self, other):
        return bytes(self) == bytes(other)


class Params(C.Structure):
    """margin_b200_params: PairwiseAlignmentParameters (inc/pairwiseAligner.h:28-39)."""

    _fields_ = [
        ("threshold", C.c_double),
        ("min_diags_between_traceback", C.c_int64),
        ("traceback_diagonals", C.c_int64),
        ("diagonal_expansion", C.c_int64),
        ("split_matrix_bigger_than_this", C.c_int64),
        ("dynamic_anchor_expansion", C.c_int32),
        ("reserved", C.c_int32),
    ]

    @classmethod
    def defaults(cls):
        """pairwiseAlignmentBandingParameters_construct (impl/pairwiseAligner.c:1048-1060)."""
        return cls(0.01, 1000, 40, 20, 3000 * 3000, 0, 0)

    def copy(self, **kw):
        p = Params.from_buffer_copy(bytes(self))
        for k, v in kw.items():
            setattr(p, k, v)
        return p


class BatchStruct(C.Structure):
    _fields_ = [
        ("n", C.c_int64),
        ("sym_x", C.c_void_p),
        ("off_x", C.c_void_p),
        ("sym_y", C.c_void_p),
        ("off_y", C.c_void_p),
        ("anchors", C.c_void_p),
        ("off_a", C.c_void_p),
        ("model_slot", C.c_void_p),
        ("ragged_left", C.c_void_p),
        ("ragged_right", C.c_void_p),
        ("anchors_delta", C.c_void_p),
        ("packed_x", C.c_void_p),
        ("packed_y", C.c_void_p),
        ("anchors_runs", C.c_void_p),
        ("off_r", C.c_void_p),
    ]


PAIR_DTYPE = np.dtype([("w", "<i4"), ("x", "<i4"), ("y", "<i4")])
PAIRS_TRIPLES, PAIRS_PACKED64, PAIRS_DELTA32 = 0, 1, 2  # MARGIN_B200_PAIRS_*
F64, F32 = 0, 1  # MARGIN_B200_F64 / MARGIN_B200_F32 (margin_b200_set_precision)


def unpack_pairs(words):
    """MARGIN_B200_PAIRS_PACKED64 words -> (w, x, y) records (margin_b200_unpack_pair)."""
    words = np.asarray(words, dtype=np.uint64)
    out = np.empty(len(words), dtype=PAIR_DTYPE)
    out["w"] = (words & np.uint64(0xffffff)).astype(np.int32)
    out["x"] = ((words >> np.uint64(24)) & np.uint64(0xfffff)).astype(np.int32) - 1
    out["y"] = (words >> np.uint64(44)).astype(np.int32) - 1
    return out


def unpack_delta32(records, escapes, off, esc_off):
    """MARGIN_B200_PAIRS_DELTA32 records of ONE list of a whole batch -> (w, x, y) records (margin_b200_delta_next,
    vectorised): `off` / `esc_off` are the per-alignment offsets into `records` / `escapes`.  Every alignment's first
    tuple is an escape, so one running sum over the batch decodes all alignments at once."""
    rec = np.asarray(records, dtype=np.uint32)
    out = np.empty(len(rec), dtype=PAIR_DTYPE)
    out["w"] = (rec & np.uint32(0xffffff)).astype(np.int32)
    if len(rec) == 0:
        return out
    code = (rec >> np.uint32(24)).astype(np.int64)
    is_esc = code == 0xFF
    dx = np.where(is_esc, 0, (code & 15) - 8)
    dy = np.where(is_esc, 0, (code >> 4) - ((code & 15) - 8))
    esc = np.asarray(escapes, dtype=np.int32).reshape(-1, 2).astype(np.int64)
    k = np.cumsum(is_esc) - 1  # index of the governing escape
    assert is_esc[0] and int(is_esc.sum()) == len(esc), "escape stream does not match the records"
    # offsets of the escapes are consistent with the alignments' first tuples
    first = np.asarray(off[:-1])[np.diff(np.asarray(off)) > 0]
    assert is_esc[first].all() and np.array_equal(k[first], np.asarray(esc_off[:-1])[np.diff(np.asarray(off)) > 0])
    cx, cy = np.cumsum(dx), np.cumsum(dy)
    at = np.flatnonzero(is_esc)
    out["x"] = (esc[k, 0] + cx - cx[at][k]).astype(np.int32)
    out["y"] = (esc[k, 1] + cy - cy[at][k]).astype(np.int32)
    return out


def pack_anchors_delta(anchors, off_a):
    """(x, y, expansion) triples -> the compact words of margin_b200_batch.anchors_delta (margin_b200_pack_anchor),
    or None when some anchor does not fit (step >= 4096, expansion >= 512 or odd)."""
    a = np.asarray(anchors, dtype=np.int64).reshape(-1, 3)
    if len(a) == 0:
        return np.zeros(1, np.uint32)
    prev = np.empty((len(a), 2), np.int64)
    prev[1:] = a[:-1, :2]
    first = np.asarray(off_a[:-1], dtype=np.int64)
    first = first[first < len(a)]
    prev[first] = -1
    d = a[:, :2] - prev
    e = a[:, 2]
    if (d < 1).any() or (d > 4095).any() or (e < 0).any() or (e > 510).any() or (e & 1).any():
        return None
    return (d[:, 0] | (d[:, 1] << 12) | ((e // 2) << 24)).astype(np.uint32)


class ResultStruct(C.Structure):
    _fields_ = [
        ("n", C.c_int64),
        ("pairs", C.c_void_p * 3),
        ("off", C.c_void_p * 3),
        ("band_cells", C.c_int64),
        ("diagonals", C.c_int64),
        ("jobs", C.c_int64),
        ("segments", C.c_int64),
        ("gpu_ms", C.c_float),
        ("kernel_launches", C.c_int32),
        ("tiles", C.c_int64),
        ("pair_format", C.c_int32),
        ("reserved", C.c_int32),
        ("wide_cells", C.c_int64),
        ("sequential_cells", C.c_int64),
        ("guard_seams", C.c_int32),
        ("guard_failures", C.c_int32),
        ("escapes", C.c_void_p * 3),
        ("escape_off", C.c_void_p * 3),
    ]


class Batch:
    """Host-side packing of a list of alignments into the flat margin_b200_batch layout.

    Each item: (sym_x uint16[], sym_y uint16[], anchors int32[k,3], model_slot, ragged_left, ragged_right).
    """

    def __init__(self, items, compact=False):
        n = len(items)
        self.n = n
        self.compact = False
        lx = np.array([len(it[0]) for it in items], dtype=np.int64)
        ly = np.array([len(it[1]) for it in items], dtype=np.int64)
        la = np.array([len(it[2]) for it in items], dtype=np.int64)
        self.off_x = np.concatenate([[0], np.cumsum(lx)]).astype(np.int64)
        self.off_y = np.concatenate([[0], np.cumsum(ly)]).astype(np.int64)
        self.off_a = np.concatenate([[0], np.cumsum(la)]).astype(np.int64)
        cat = lambda arrs, dt, shape: (np.ascontiguousarray(np.concatenate(arrs).astype(dt)) if n and sum(
            len(a) for a in arrs) else np.zeros(shape, dtype=dt))
        self.sym_x = cat([np.asarray(it[0], dtype=np.uint16) for it in items], np.uint16, (1,))
        self.sym_y = cat([np.asarray(it[1], dtype=np.uint16) for it in items], np.uint16, (1,))
        self.anchors = cat([np.asarray(it[2], dtype=np.int32).reshape(-1, 3) for it in items], np.int32, (1, 3))
        self.model_slot = np.array([it[3] for it in items], dtype=np.uint8) if n else np.zeros(1, np.uint8)
        self.ragged_left = np.array([it[4] for it in items], dtype=np.uint8) if n else np.zeros(1, np.uint8)
        self.ragged_right = np.array([it[5] for it in items], dtype=np.uint8) if n else np.zeros(1, np.uint8)
        self.struct = BatchStruct(
            n, self.sym_x.ctypes.data, self.off_x.ctypes.data, self.sym_y.ctypes.data, self.off_y.ctypes.data,
            self.anchors.ctypes.data, self.off_a.ctypes.data, self.model_slot.ctypes.data,
            self.ragged_left.ctypes.data, self.ragged_right.ctypes.data, None, None, None, None, None)
        self.runs = False
        self.packed = False
        if compact:
            self.use_compact_anchors()
            self.use_packed_symbols()
            self.use_anchor_runs()

    def use_compact_anchors(self):
        """Switches the batch to the 4-byte anchor words (what a C host emits while walking its anchor lists); keeps
        the triples when some anchor does not fit.  Returns whether the compact form is in use."""
        d = pack_anchors_delta(self.anchors, self.off_a) if self.n else None
        if d is None:
            return False
        self.anchors_delta = np.ascontiguousarray(d)
        self.struct.anchors_delta = self.anchors_delta.ctypes.data  # takes precedence; the triples stay for the oracle
        self.compact = True
        return True

    def use_packed_symbols(self):
        """One byte per symbol (margin_b200_pack_symbol); keeps the uint16 arrays when a run length does not fit."""
        def pack(a):
            b = np.minimum(a & 255, 4).astype(np.uint16)
            r = a >> 8
            return None if (r >= MAX_RL).any() else (b * MAX_RL + r).astype(np.uint8)
        px, py = pack(self.sym_x), pack(self.sym_y)
        if px is None or py is None:
            return False
        self.packed_x, self.packed_y = np.ascontiguousarray(px), np.ascontiguousarray(py)
        self.struct.packed_x, self.struct.packed_y = self.packed_x.ctypes.data, self.packed_y.ctypes.data
        self.packed = True
        return True

    def use_anchor_runs(self):
        """The anchors as runs (margin_b200_pack_anchor_run): an anchor at step (+1, +1) from its predecessor with the same
        expansion extends the predecessor's run (up to 255 followers).  Keeps the other forms when an anchor does not fit."""
        a = np.asarray(self.anchors, dtype=np.int64).reshape(-1, 3)
        na = int(self.off_a[-1])
        if na == 0 or self.n == 0:
            return False
        a = a[:na]
        prev = np.empty((na, 3), np.int64)
        prev[1:] = a[:-1]
        first = np.asarray(self.off_a[:-1], dtype=np.int64)
        first = first[np.diff(self.off_a) > 0]
        prev[first, :2] = -1
        prev[first, 2] = -2  # never equal to an expansion: the first anchor of an alignment starts a run
        d = a[:, :2] - prev[:, :2]
        e = a[:, 2]
        if (d < 1).any() or (d > 1023).any() or (e < 0).any() or (e > 30).any() or (e & 1).any():
            return False
        follow = (d[:, 0] == 1) & (d[:, 1] == 1) & (e == prev[:, 2])
        # position inside a maximal stretch of followers; every 256th follower starts a new run
        idx = np.arange(na)
        start = np.maximum.accumulate(np.where(~follow, idx, 0))
        follow &= ((idx - start) % 256) != 0
        heads = np.flatnonzero(~follow)
        run = np.diff(np.append(heads, na)) - 1
        words = (run | (d[heads, 0] << 8) | (d[heads, 1] << 18) | ((e[heads] // 2) << 28)).astype(np.uint32)
        self.anchors_runs = np.ascontiguousarray(words)
        self.off_r = np.searchsorted(heads, np.asarray(self.off_a, dtype=np.int64)).astype(np.int64)
        self.struct.anchors_runs, self.struct.off_r = self.anchors_runs.ctypes.data, self.off_r.ctypes.data
        self.runs = True
        return True

    def to_pinned(self):
        """Moves the arrays that cross PCIe into page-locked memory from margin_b200_pinned_alloc — where the margin-side
        binding assembles its batches (compat/pairwiseAligner_b200.c: b200_pinned_grow) — and re-points the struct."""
        lib = load_library()
        self._pinned = []

        def move(a):
            p = lib.margin_b200_pinned_alloc(max(a.nbytes, 1))
            if not p:
                raise RuntimeError(lib.margin_b200_last_error().decode())
            self._pinned.append(p)
            out = np.frombuffer((C.c_char * max(a.nbytes, 1)).from_address(p), dtype=a.dtype, count=a.size).reshape(a.shape)
            out[...] = a
            return out
        if self.packed:
            self.packed_x, self.packed_y = move(self.packed_x), move(self.packed_y)
            self.struct.packed_x, self.struct.packed_y = self.packed_x.ctypes.data, self.packed_y.ctypes.data
        else:
            self.sym_x, self.sym_y = move(self.sym_x), move(self.sym_y)
            self.struct.sym_x, self.struct.sym_y = self.sym_x.ctypes.data, self.sym_y.ctypes.data
        if self.runs:
            self.anchors_runs = move(self.anchors_runs)
            self.struct.anchors_runs = self.anchors_runs.ctypes.data
        elif self.compact:
            self.anchors_delta = move(self.anchors_delta)
            self.struct.anchors_delta = self.anchors_delta.ctypes.data
        else:
            self.anchors = move(self.anchors)
            self.struct.anchors = self.anchors.ctypes.data
        return self

    def release_pinned(self):
        lib = load_library()
        for p in getattr(self, "_pinned", []):
            lib.margin_b200_pinned_free(p)
        self._pinned = []

    def item(self, i):
        return (self.sym_x[self.off_x[i]:self.off_x[i + 1]], self.sym_y[self.off_y[i]:self.off_y[i + 1]],
                self.anchors[self.off_a[i]:self.off_a[i + 1]], int(self.model_slot[i]), int(self.ragged_left[i]),
                int(self.ragged_right[i]))

    @property
    def h2d_bytes(self):
        anchors = (self.anchors_runs.nbytes + self.off_r.nbytes if self.runs
                   else (self.anchors_delta.nbytes if self.compact else self.anchors.nbytes))
        syms = self.packed_x.nbytes + self.packed_y.nbytes if self.packed else self.sym_x.nbytes + self.sym_y.nbytes
        return int(syms + anchors + 3 * self.off_x.nbytes + 3 * self.n)


_LIB = None
LIB_PATH = os.environ.get("MARGIN_B200_LIB") or os.path.join(os.path.dirname(os.path.abspath(__file__)), "csrc",
                                                           "libmargin_b200.so")  # env: A/B builds of the kernels


def load_library():
    """dlopen the CUDA library.  Raises (never falls back) when it is missing."""
    global _LIB
    if _LIB is not None:
        return _LIB
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            "margin_b200: %s is not built; run __graft_entry__.build() (needs nvcc). "
            "There is no CPU fallback." % LIB_PATH)
    lib = C.CDLL(LIB_PATH)
    vp, i32, i64 = C.c_void_p, C.c_int, C.c_int64
    lib.margin_b200_create.argtypes = [C.POINTER(vp), i32]
    lib.margin_b200_create.restype = i32
    lib.margin_b200_destroy.argtypes = [vp]
    lib.margin_b200_destroy.restype = None
    lib.margin_b200_last_error.argtypes = []
    lib.margin_b200_last_error.restype = C.c_char_p
    lib.margin_b200_device_count.argtypes = []
    lib.margin_b200_device_count.restype = i32
    lib.margin_b200_prewarm.argtypes = []
    lib.margin_b200_prewarm.restype = i32
    lib.margin_b200_prewarm_devices.argtypes = [i32]
    lib.margin_b200_prewarm_devices.restype = i32
    lib.margin_b200_version.argtypes = []
    lib.margin_b200_version.restype = i32
    lib.margin_b200_set_model.argtypes = [vp, i32, C.POINTER(Model)]
    lib.margin_b200_set_model.restype = i32
    lib.margin_b200_set_params.argtypes = [vp, C.POINTER(Params)]
    lib.margin_b200_set_params.restype = i32
    lib.margin_b200_set_kernel_timing.argtypes = [vp, i32]
    lib.margin_b200_set_kernel_timing.restype = i32
    lib.margin_b200_set_pair_format.argtypes = [vp, i32]
    lib.margin_b200_set_pair_format.restype = i32
    lib.margin_b200_pinned_alloc.argtypes = [C.c_size_t]
    lib.margin_b200_pinned_alloc.restype = vp
    lib.margin_b200_pinned_free.argtypes = [vp]
    lib.margin_b200_pinned_free.restype = None
    lib.margin_b200_set_precision.argtypes = [vp, i32]
    lib.margin_b200_set_precision.restype = i32
    lib.margin_b200_get_precision.argtypes = [vp]
    lib.margin_b200_get_precision.restype = i32
    lib.margin_b200_posteriors_batch.argtypes = [vp, C.POINTER(BatchStruct), i32, C.POINTER(ResultStruct)]
    lib.margin_b200_posteriors_batch.restype = i32
    lib.margin_b200_upload_batch.argtypes = [vp, C.POINTER(BatchStruct)]
    lib.margin_b200_upload_batch.restype = i32
    lib.margin_b200_posteriors_resident.argtypes = [vp, i32, C.POINTER(ResultStruct)]
    lib.margin_b200_posteriors_resident.restype = i32
    lib.margin_b200_download_result.argtypes = [vp, C.POINTER(ResultStruct)]
    lib.margin_b200_download_result.restype = i32
    lib.margin_b200_forward_batch.argtypes = [vp, C.POINTER(BatchStruct), vp, C.POINTER(ResultStruct)]
    lib.margin_b200_forward_batch.restype = i32
    lib.margin_b200_forward_tiny_noalloc.argtypes = [vp, i64, vp, vp, i64, vp, vp, i64, vp, vp]
    lib.margin_b200_forward_tiny_noalloc.restype = i32
    lib.margin_b200_forward_resident.argtypes = [vp, C.POINTER(ResultStruct)]
    lib.margin_b200_forward_resident.restype = i32
    lib.margin_b200_bands.argtypes = [vp, C.POINTER(BatchStruct), vp, vp]
    lib.margin_b200_bands.restype = i32
    lib.margin_b200_last_kernel_ms.argtypes = [vp, vp, vp, i32]
    lib.margin_b200_last_kernel_ms.restype = i32
    _LIB = lib
    return lib


EXPORTED_SYMBOLS = [
    "margin_b200_create", "margin_b200_destroy", "margin_b200_last_error", "margin_b200_version",
    "margin_b200_set_model", "margin_b200_set_params", "margin_b200_set_pair_format", "margin_b200_set_precision",
    "margin_b200_get_precision", "margin_b200_pinned_alloc", "margin_b200_pinned_free", "margin_b200_set_kernel_timing", "margin_b200_posteriors_batch",
    "margin_b200_upload_batch", "margin_b200_posteriors_resident", "margin_b200_download_result",
    "margin_b200_forward_batch", "margin_b200_forward_resident", "margin_b200_forward_tiny_noalloc", "margin_b200_bands",
    "margin_b200_last_kernel_ms", "margin_b200_device_count", "margin_b200_prewarm", "margin_b200_prewarm_devices",
]

KERNEL_CLASSES = ["band", "schedule", "forward", "backward", "scan", "gather", "offsets", "final", "guard"]
