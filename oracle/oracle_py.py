"""ctypes doorway to the parity checkers.  TEST INFRASTRUCTURE ONLY.

* :class:`Oracle`    — oracle/liboracle.so (our C restatement, pairhmm_oracle.c)
* :class:`Reference` — oracle/_ref/libmargin_ref.so (the unmodified reference + ref_harness.c)

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
import this module.  Nothing under margin_b200/ does.
"""
import ctypes as C
import os
import subprocess

import numpy as np

from margin_b200 import abi

HERE = os.path.dirname(os.path.abspath(__file__))
ORACLE_SO = os.path.join(HERE, "liboracle.so")
REF_SO = os.path.join(HERE, "_ref", "libmargin_ref.so")


def build(ref=True):
    """Compile the checkers (gcc).  The reference part only when /root/reference exists."""
    subprocess.check_call(["make", "-s", "-C", HERE, "oracle"])
    if ref and os.path.isdir("/root/reference"):
        subprocess.check_call(["make", "-s", "-j8", "-C", HERE, "ref"])
        # the reference CLI, with and without compat/pairwiseAligner_b200.c, + the bundled region (end-to-end test)
        subprocess.check_call(["make", "-s", "-j8", "-C", HERE, "polish"])


class Lists(C.Structure):
    _fields_ = [("pairs", C.c_void_p * 3), ("count", C.c_int64 * 3), ("cells", C.c_int64),
                ("diagonals", C.c_int64), ("tracebacks", C.c_int64), ("max_width", C.c_int64),
                ("blocks", C.c_int64)]

    def to_numpy(self):
        out = []
        for l in range(3):
            n = self.count[l]
            if n:
                buf = (C.c_char * (n * 12)).from_address(self.pairs[l])
                out.append(np.frombuffer(buf, dtype=abi.PAIR_DTYPE).copy())
            else:
                out.append(np.zeros(0, dtype=abi.PAIR_DTYPE))
        return out


def _u16(a):
    return np.ascontiguousarray(a, dtype=np.uint16)


def _anc(a):
    return np.ascontiguousarray(np.asarray(a, dtype=np.int32).reshape(-1, 3))


class Oracle:
    def __init__(self):
        if not os.path.exists(ORACLE_SO):
            build(ref=False)
        lib = C.CDLL(ORACLE_SO)
        vp, i64, i32, dbl = C.c_void_p, C.c_int64, C.c_int, C.c_double
        lib.mb_oracle_logadd.argtypes = [dbl, dbl]
        lib.mb_oracle_logadd.restype = dbl
        lib.mb_oracle_band.argtypes = [vp, i64, i64, i64, i32, i64, vp, vp]
        lib.mb_oracle_band.restype = None
        lib.mb_oracle_split_points.argtypes = [vp, i64, i64, i64, i64, i32, i32, vp]
        lib.mb_oracle_split_points.restype = i64
        lib.mb_oracle_posteriors.argtypes = [C.POINTER(abi.Model), C.POINTER(abi.Params), vp, i64, vp, i64, vp, i64,
                                             i32, i32, i32, C.POINTER(Lists)]
        lib.mb_oracle_posteriors.restype = i32
        lib.mb_oracle_free_lists.argtypes = [C.POINTER(Lists)]
        lib.mb_oracle_free_lists.restype = None
        lib.mb_oracle_forward.argtypes = [C.POINTER(abi.Model), C.POINTER(abi.Params), vp, i64, vp, i64, vp, i64, i32,
                                          i32, C.POINTER(i64)]
        lib.mb_oracle_forward.restype = dbl
        lib.mb_oracle_posteriors_batch.argtypes = [vp, C.POINTER(abi.Params), C.POINTER(abi.BatchStruct), i32, vp, i32]
        lib.mb_oracle_posteriors_batch.restype = i32
        lib.mb_oracle_forward_batch.argtypes = [vp, C.POINTER(abi.Params), C.POINTER(abi.BatchStruct), vp,
                                                C.POINTER(i64), i32]
        lib.mb_oracle_forward_batch.restype = i32
        self.lib = lib

    def logadd(self, x, y):
        return self.lib.mb_oracle_logadd(x, y)

    def band(self, anchors, lx, ly, dynamic, expansion):
        a = _anc(anchors)
        kl = np.zeros(lx + ly + 1, np.int64)
        kr = np.zeros(lx + ly + 1, np.int64)
        self.lib.mb_oracle_band(a.ctypes.data, len(a), lx, ly, int(dynamic), expansion, kl.ctypes.data,
                                kr.ctypes.data)
        return kl, kr

    def split_points(self, anchors, lx, ly, limit, ragged_left, ragged_right):
        a = _anc(anchors)
        sp = np.zeros(4 * (len(a) + 2), np.int64)
        n = self.lib.mb_oracle_split_points(a.ctypes.data, len(a), lx, ly, limit, int(ragged_left),
                                            int(ragged_right), sp.ctypes.data)
        return sp[:4 * n].reshape(-1, 4)

    def posteriors(self, model, params, sx, sy, anchors, ragged_left=0, ragged_right=0, want_indels=1,
                   with_stats=False):
        sx, sy, a = _u16(sx), _u16(sy), _anc(anchors)
        res = Lists()
        self.lib.mb_oracle_posteriors(C.byref(model), C.byref(params), sx.ctypes.data, len(sx), sy.ctypes.data,
                                      len(sy), a.ctypes.data, len(a), int(ragged_left), int(ragged_right),
                                      int(want_indels), C.byref(res))
        out = res.to_numpy()
        stats = dict(cells=res.cells, diagonals=res.diagonals, tracebacks=res.tracebacks, max_width=res.max_width,
                     blocks=res.blocks)
        self.lib.mb_oracle_free_lists(C.byref(res))
        return (out, stats) if with_stats else out

    def forward(self, model, params, sx, sy, anchors=(), ragged_left=0, ragged_right=0):
        sx, sy, a = _u16(sx), _u16(sy), _anc(anchors)
        return self.lib.mb_oracle_forward(C.byref(model), C.byref(params), sx.ctypes.data, len(sx), sy.ctypes.data,
                                          len(sy), a.ctypes.data, len(a), int(ragged_left), int(ragged_right), None)

    def posteriors_batch(self, models, params, batch, want_indels=1, threads=1, keep=True):
        """Returns (list of [3 arrays] per alignment or None, dict of totals)."""
        marr = (abi.Model * 2)(models[0], models[1])
        res = (Lists * max(batch.n, 1))()
        self.lib.mb_oracle_posteriors_batch(C.addressof(marr), C.byref(params), C.byref(batch.struct),
                                            int(want_indels), C.addressof(res), threads)
        out = [] if keep else None
        tot = dict(cells=0, diagonals=0, tracebacks=0, tuples=0)
        for i in range(batch.n):
            if keep:
                out.append(res[i].to_numpy())
            tot["cells"] += res[i].cells
            tot["diagonals"] += res[i].diagonals
            tot["tracebacks"] += res[i].tracebacks
            tot["tuples"] += sum(res[i].count)
            self.lib.mb_oracle_free_lists(C.byref(res[i]))
        return out, tot

    def forward_batch(self, models, params, batch, threads=1):
        marr = (abi.Model * 2)(models[0], models[1])
        out = np.zeros(max(batch.n, 1), np.float64)
        cells = C.c_int64(0)
        self.lib.mb_oracle_forward_batch(C.addressof(marr), C.byref(params), C.byref(batch.struct), out.ctypes.data,
                                         C.byref(cells), threads)
        return out[:batch.n], cells.value


class Reference:
    """The unmodified reference (compiled from /root/reference by oracle/Makefile)."""

    @staticmethod
    def available():
        return os.path.exists(REF_SO)

    def __init__(self):
        lib = C.CDLL(REF_SO)
        vp, i64, i32, dbl = C.c_void_p, C.c_int64, C.c_int, C.c_double
        lib.ref_params_load.argtypes = [C.c_char_p]
        lib.ref_params_load.restype = vp
        lib.ref_params_state_machine.argtypes = [vp, i32]
        lib.ref_params_state_machine.restype = vp
        lib.ref_params_alignment.argtypes = [vp, C.POINTER(abi.Params), C.POINTER(i32), C.POINTER(i32),
                                             C.POINTER(i32)]
        lib.ref_params_alignment.restype = None
        lib.ref_default_nucleotide_state_machine.argtypes = []
        lib.ref_default_nucleotide_state_machine.restype = vp
        lib.ref_flatten_state_machine.argtypes = [vp, i32, C.POINTER(abi.Model)]
        lib.ref_flatten_state_machine.restype = i32
        lib.ref_emission_match.argtypes = [vp, i32, i32]
        lib.ref_emission_match.restype = dbl
        lib.ref_emission_gap_x.argtypes = [vp, i32]
        lib.ref_emission_gap_x.restype = dbl
        lib.ref_emission_gap_y.argtypes = [vp, i32]
        lib.ref_emission_gap_y.restype = dbl
        lib.ref_state_machine_from_model.argtypes = [C.POINTER(abi.Model)]
        lib.ref_state_machine_from_model.restype = vp
        lib.ref_logadd.argtypes = [dbl, dbl]
        lib.ref_logadd.restype = dbl
        lib.ref_band.argtypes = [vp, i64, i64, i64, i32, i64, vp, vp]
        lib.ref_band.restype = None
        lib.ref_split_points.argtypes = [vp, i64, i64, i64, i64, i32, i32, vp]
        lib.ref_split_points.restype = i64
        lib.ref_posteriors.argtypes = [vp, C.POINTER(abi.Params), vp, i64, vp, i64, vp, i64, i32, i32, i32,
                                       C.POINTER(Lists)]
        lib.ref_posteriors.restype = i32
        lib.ref_free_lists.argtypes = [C.POINTER(Lists)]
        lib.ref_free_lists.restype = None
        lib.ref_forward.argtypes = [vp, C.POINTER(abi.Params), vp, i64, vp, i64, vp, i64, i32, i32]
        lib.ref_forward.restype = dbl
        lib.ref_posteriors_batch.argtypes = [vp, C.POINTER(abi.Params), C.POINTER(abi.BatchStruct), i32, vp, i32]
        lib.ref_posteriors_batch.restype = i32
        lib.ref_forward_batch.argtypes = [vp, C.POINTER(abi.Params), C.POINTER(abi.BatchStruct), vp, i32]
        lib.ref_forward_batch.restype = i32
        self.lib = lib

    # -- models -------------------------------------------------------------
    def load_params(self, path):
        """Returns dict(sms=(fwd, rev) StateMachine*, models=(Model, Model), p=Params, flags...)."""
        h = self.lib.ref_params_load(path.encode())
        p = abi.Params()
        a, b, c = C.c_int(), C.c_int(), C.c_int()
        self.lib.ref_params_alignment(h, C.byref(p), C.byref(a), C.byref(b), C.byref(c))
        sms = (self.lib.ref_params_state_machine(h, 0), self.lib.ref_params_state_machine(h, 1))
        models = (self.flatten(sms[0], b.value), self.flatten(sms[1], b.value))
        return dict(handle=h, sms=sms, models=models, p=p, use_rle=bool(a.value), use_repeat_counts=bool(b.value),
                    max_rl=c.value)

    def default_nucleotide_sm(self):
        return self.lib.ref_default_nucleotide_state_machine()

    def flatten(self, sm, rle):
        m = abi.Model()
        if self.lib.ref_flatten_state_machine(sm, int(rle), C.byref(m)) != 0:
            raise RuntimeError("ref_flatten_state_machine failed")
        return m

    def sm_from_model(self, model):
        """Reference StateMachine built from a flattened model (no params JSON needed)."""
        return self.lib.ref_state_machine_from_model(C.byref(model))

    def emission_match(self, sm, cx, cy):
        return self.lib.ref_emission_match(sm, cx, cy)

    def emission_gap_x(self, sm, cx):
        return self.lib.ref_emission_gap_x(sm, cx)

    def emission_gap_y(self, sm, cy):
        return self.lib.ref_emission_gap_y(sm, cy)

    # -- functions ----------------------------------------------------------
    def logadd(self, x, y):
        return self.lib.ref_logadd(x, y)

    def band(self, anchors, lx, ly, dynamic, expansion):
        a = _anc(anchors)
        kl = np.zeros(lx + ly + 1, np.int64)
        kr = np.zeros(lx + ly + 1, np.int64)
        self.lib.ref_band(a.ctypes.data, len(a), lx, ly, int(dynamic), expansion, kl.ctypes.data, kr.ctypes.data)
        return kl, kr

    def split_points(self, anchors, lx, ly, limit, ragged_left, ragged_right):
        a = _anc(anchors)
        sp = np.zeros(4 * (len(a) + 2), np.int64)
        n = self.lib.ref_split_points(a.ctypes.data, len(a), lx, ly, limit, int(ragged_left), int(ragged_right),
                                      sp.ctypes.data)
        return sp[:4 * n].reshape(-1, 4)

    def posteriors(self, sm, params, sx, sy, anchors, ragged_left=0, ragged_right=0, want_indels=1):
        sx, sy, a = _u16(sx), _u16(sy), _anc(anchors)
        res = Lists()
        self.lib.ref_posteriors(sm, C.byref(params), sx.ctypes.data, len(sx), sy.ctypes.data, len(sy), a.ctypes.data,
                                len(a), int(ragged_left), int(ragged_right), int(want_indels), C.byref(res))
        out = res.to_numpy()
        self.lib.ref_free_lists(C.byref(res))
        return out

    def forward(self, sm, params, sx, sy, anchors=(), ragged_left=0, ragged_right=0):
        sx, sy, a = _u16(sx), _u16(sy), _anc(anchors)
        return self.lib.ref_forward(sm, C.byref(params), sx.ctypes.data, len(sx), sy.ctypes.data, len(sy),
                                    a.ctypes.data, len(a), int(ragged_left), int(ragged_right))

    def posteriors_batch(self, sms, params, batch, want_indels=1, threads=1):
        arr = (C.c_void_p * 2)(sms[0], sms[1])
        counts = np.zeros(3, np.int64)
        self.lib.ref_posteriors_batch(C.addressof(arr), C.byref(params), C.byref(batch.struct), int(want_indels),
                                      counts.ctypes.data, threads)
        return counts

    def forward_batch(self, sms, params, batch, threads=1):
        arr = (C.c_void_p * 2)(sms[0], sms[1])
        out = np.zeros(max(batch.n, 1), np.float64)
        self.lib.ref_forward_batch(C.addressof(arr), C.byref(params), C.byref(batch.struct), out.ctypes.data, threads)
        return out[:batch.n]
