"""Host-side mirror of the reference's aligner interface (inc/pairwiseAligner.h) on top of the C ABI.

Function names, argument meaning and error behaviour follow the reference:

* :meth:`PairHmmAligner.get_aligned_pairs_with_indels_using_anchors`  <- getAlignedPairsWithIndelsUsingAnchors (:81-84)
* :meth:`PairHmmAligner.get_aligned_pairs_using_anchors`              <- getAlignedPairsUsingAnchors (:77-79)
* :meth:`PairHmmAligner.get_aligned_pairs` / ``_with_indels``          <- getAlignedPairs / getAlignedPairsWithIndels (:62-72)
* :meth:`PairHmmAligner.compute_forward_probability`                   <- computeForwardProbability (:54-57)
* :meth:`PairHmmAligner.posteriors_batch` / :meth:`forward_batch`      <- the batched entry points

Lists come back as numpy structured arrays (w, x, y) in the reference's stList order.  Errors of the
library (bad anchors -> PAIRWISE_ALIGNMENT_EXCEPTION, missing device, ...) raise :class:`AlignerError`.
All computation happens in libmargin_b200.so on the GPU; nothing here computes.
"""
import ctypes as C

import numpy as np

from . import abi


class AlignerError(RuntimeError):
    pass


class BatchResult:
    def __init__(self, lists, stats):
        self.lists = lists  # per alignment: [match, gapX, gapY] structured arrays
        self.stats = stats

    def __len__(self):
        return len(self.lists)

    def __getitem__(self, i):
        return self.lists[i]


def _stats(res):
    return dict(band_cells=res.band_cells, diagonals=res.diagonals, jobs=res.jobs, segments=res.segments,
                gpu_ms=res.gpu_ms, kernel_launches=res.kernel_launches, tiles=res.tiles, wide_cells=res.wide_cells,
                sequential_cells=res.sequential_cells, guard_seams=res.guard_seams, guard_failures=res.guard_failures)


class PairHmmAligner:
    def __init__(self, models, params, device=0):
        self.lib = abi.load_library()
        h = C.c_void_p()
        self._h = None
        self._check(self.lib.margin_b200_create(C.byref(h), device))
        self._h = h
        self.device = device
        self.set_models(models)
        self.set_params(params)

    # ------------------------------------------------------------------ plumbing
    def _check(self, rc):
        if rc != 0:
            raise AlignerError(self.lib.margin_b200_last_error().decode())

    def close(self):
        if self._h is not None:
            self.lib.margin_b200_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_models(self, models):
        self.models = models
        for slot in (0, 1):
            self._check(self.lib.margin_b200_set_model(self._h, slot, C.byref(models[slot])))

    def set_params(self, params):
        self.params = params.copy()
        self._check(self.lib.margin_b200_set_params(self._h, C.byref(self.params)))

    def set_kernel_timing(self, on):
        """CUDA events around every launch (kernel_times()); off by default."""
        self._check(self.lib.margin_b200_set_kernel_timing(self._h, int(bool(on))))

    def set_pair_format(self, fmt):
        """abi.PAIRS_TRIPLES (12-byte records), abi.PAIRS_PACKED64 (8-byte words) or abi.PAIRS_DELTA32 (4-byte step
        records + escapes; include/margin_b200.h)."""
        self._check(self.lib.margin_b200_set_pair_format(self._h, int(fmt)))

    # ------------------------------------------------------------- batched calls
    def _collect(self, res, n, want_lists=True):
        lists = []
        if want_lists:
            arrs, offs = [], []
            for l in range(3):
                off = np.ctypeslib.as_array(C.cast(res.off[l], C.POINTER(C.c_int64)), shape=(n + 1,)).copy()
                tot = int(off[n])
                if tot and res.pair_format == abi.PAIRS_DELTA32:
                    rec = np.frombuffer((C.c_char * (tot * 4)).from_address(res.pairs[l]), dtype=np.uint32)
                    eoff = np.ctypeslib.as_array(C.cast(res.escape_off[l], C.POINTER(C.c_int64)), shape=(n + 1,)).copy()
                    ne = int(eoff[n])
                    esc = (np.frombuffer((C.c_char * (ne * 8)).from_address(res.escapes[l]), dtype=np.int32) if ne
                           else np.zeros(0, np.int32))
                    arrs.append(abi.unpack_delta32(rec, esc, off, eoff))
                elif tot and res.pair_format == abi.PAIRS_PACKED64:
                    buf = (C.c_char * (tot * 8)).from_address(res.pairs[l])
                    arrs.append(abi.unpack_pairs(np.frombuffer(buf, dtype=np.uint64)))
                elif tot:
                    buf = (C.c_char * (tot * 12)).from_address(res.pairs[l])
                    arrs.append(np.frombuffer(buf, dtype=abi.PAIR_DTYPE).copy())
                else:
                    arrs.append(np.zeros(0, dtype=abi.PAIR_DTYPE))
                offs.append(off)
            for i in range(n):
                lists.append([arrs[l][offs[l][i]:offs[l][i + 1]] for l in range(3)])
        return BatchResult(lists, _stats(res))

    def posteriors_batch(self, batch, want_indels=True):
        if not isinstance(batch, abi.Batch):
            batch = abi.Batch(batch)
        res = abi.ResultStruct()
        self._check(self.lib.margin_b200_posteriors_batch(self._h, C.byref(batch.struct), int(want_indels), C.byref(res)))
        return self._collect(res, batch.n)

    def posteriors_host(self, batch, want_indels=True):
        """margin_b200_posteriors_batch with host buffers in and the tuple lists left in the context's pinned host
        memory (no Python-side copies): what a C caller does.  Returns the counters and the number of tuples."""
        res = abi.ResultStruct()
        self._check(self.lib.margin_b200_posteriors_batch(self._h, C.byref(batch.struct), int(want_indels), C.byref(res)))
        out = _stats(res)
        n = batch.n
        out["tuples"] = sum(int(C.cast(res.off[l], C.POINTER(C.c_int64))[n]) for l in range(3)) if n else 0
        out["tuple_bytes"] = {abi.PAIRS_PACKED64: 8, abi.PAIRS_DELTA32: 4}.get(res.pair_format, 12)
        out["escapes"] = (sum(int(C.cast(res.escape_off[l], C.POINTER(C.c_int64))[n]) for l in range(3))
                          if n and res.pair_format == abi.PAIRS_DELTA32 else 0)
        return out

    def set_precision(self, precision):
        """abi.F64 (the reference's arithmetic) or abi.F32 (renormalised FP32 tile kernels, |dp| <= 1e-4)."""
        self._check(self.lib.margin_b200_set_precision(self._h, int(precision)))

    def arithmetic(self):
        """Arithmetic type of the lane-group tile kernels."""
        return "f32" if self.lib.margin_b200_get_precision(self._h) == abi.F32 else "f64"

    def f_bytes_per_cell(self):
        """Bytes of forward state per band cell that cross HBM in each direction (3 doubles, or one float4)."""
        return 16.0 if self.arithmetic() == "f32" else 24.0

    def launches_per_step(self, kernel_class):
        """Launches of a kernel class per call: one per lane-group width (8/16/32) that has jobs."""
        t = self.kernel_times()
        return max(1, t.get(kernel_class, (0.0, 1))[1])

    def upload(self, batch):
        self._check(self.lib.margin_b200_upload_batch(self._h, C.byref(batch.struct)))

    def posteriors_resident(self, want_indels=True):
        res = abi.ResultStruct()
        self._check(self.lib.margin_b200_posteriors_resident(self._h, int(want_indels), C.byref(res)))
        return _stats(res)

    def download(self, n):
        res = abi.ResultStruct()
        self._check(self.lib.margin_b200_download_result(self._h, C.byref(res)))
        return res

    def kernel_times(self):
        """{class: (ms, launches)} of the last call (CUDA events around every launch)."""
        n = len(abi.KERNEL_CLASSES)
        ms = np.zeros(n, np.float32)
        ln = np.zeros(n, np.int32)
        self._check(self.lib.margin_b200_last_kernel_ms(self._h, ms.ctypes.data, ln.ctypes.data, n))
        return {k: (float(ms[i]), int(ln[i])) for i, k in enumerate(abi.KERNEL_CLASSES)}

    def forward_batch(self, batch):
        if not isinstance(batch, abi.Batch):
            batch = abi.Batch(batch)
        out = np.zeros(max(batch.n, 1), np.float64)
        res = abi.ResultStruct()
        self._check(self.lib.margin_b200_forward_batch(self._h, C.byref(batch.struct), out.ctypes.data, C.byref(res)))
        self.last_stats = _stats(res)
        return out[:batch.n]

    def forward_tiny_noalloc(self, n, d_sym_x, d_off_x, total_x, d_sym_y, d_off_y, total_y, d_model_slot, d_out):
        """margin_b200_forward_tiny_noalloc: device pointers (ints) in and out, fixed scratch inside the library, no
        allocation (include/margin_b200.h)."""
        self._check(self.lib.margin_b200_forward_tiny_noalloc(self._h, int(n), d_sym_x, d_off_x, int(total_x), d_sym_y, d_off_y,
                                                              int(total_y), d_model_slot, d_out))

    def forward_resident(self):
        res = abi.ResultStruct()
        self._check(self.lib.margin_b200_forward_resident(self._h, C.byref(res)))
        return _stats(res)

    def bands(self, batch):
        if not isinstance(batch, abi.Batch):
            batch = abi.Batch(batch)
        nd = [int(batch.off_x[i + 1] - batch.off_x[i] + batch.off_y[i + 1] - batch.off_y[i] + 1) for i in range(batch.n)]
        tot = sum(nd)
        kl = np.zeros(max(tot, 1), np.int32)
        kr = np.zeros(max(tot, 1), np.int32)
        self._check(self.lib.margin_b200_bands(self._h, C.byref(batch.struct), kl.ctypes.data, kr.ctypes.data))
        out, o = [], 0
        for k in nd:
            out.append((kl[o:o + k].copy(), kr[o:o + k].copy()))
            o += k
        return out

    # ---------------------------------------------- the reference's per-call API
    def get_aligned_pairs_with_indels_using_anchors(self, slot, sx, sy, anchors, ragged_left=False, ragged_right=False):
        """Returns (alignedPairs, gapXPairs, gapYPairs)."""
        r = self.posteriors_batch([(sx, sy, anchors, slot, int(ragged_left), int(ragged_right))], True)
        return tuple(r[0])

    def get_aligned_pairs_using_anchors(self, slot, sx, sy, anchors, ragged_left=False, ragged_right=False):
        r = self.posteriors_batch([(sx, sy, anchors, slot, int(ragged_left), int(ragged_right))], False)
        return r[0][0]

    def get_aligned_pairs(self, slot, sx, sy, ragged_left=False, ragged_right=False):
        return self.get_aligned_pairs_using_anchors(slot, sx, sy, np.zeros((0, 3), np.int32), ragged_left, ragged_right)

    def get_aligned_pairs_with_indels(self, slot, sx, sy, ragged_left=False, ragged_right=False):
        return self.get_aligned_pairs_with_indels_using_anchors(slot, sx, sy, np.zeros((0, 3), np.int32), ragged_left,
                                                                ragged_right)

    def compute_forward_probability(self, slot, sx, sy, anchors=(), ragged_left=False, ragged_right=False):
        a = np.asarray(anchors, dtype=np.int32).reshape(-1, 3)
        return float(self.forward_batch([(sx, sy, a, slot, int(ragged_left), int(ragged_right))])[0])
