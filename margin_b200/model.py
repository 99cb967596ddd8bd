"""Host-side mirror of the reference's parameter loading for the pair-HMM path.

Reads a margin params JSON (with ``include`` chaining) and produces the two
flattened state machines (forward-strand / reverse-strand read) and the
alignment parameters that the CUDA library consumes.  Follows, step by step:

* ``params_readParams2`` include chaining           impl/parser.c:556-620
* ``polishParams_jsonParse``                         impl/parser.c:292-488
* ``hmm_jsonParse`` / ``hmm_getStateMachine``        impl/stateMachine.c:206-268, :690-703
* ``stateMachine3_loadSymmetric/Asymmetric``         impl/stateMachine.c:646-682
* ``emissions_construct`` (log of the 16+4+4 table)  impl/stateMachine.c:481-491
* ``nucleotideEmissions_reverseComplement``          impl/stateMachine.c:457-473
* ``repeatSubMatrix_jsonParse`` / ``_setLogProb``    impl/parser.c:198-251, impl/repeatSubMatrix.c:11-35
* ``polishParams_finishParsing`` (RLE emissions)     impl/parser.c:495-525

tests/test_model.py checks the result bit-for-bit against the reference parser
(through oracle/_ref) for every params file the reference ships for this path.
"""
import json
import math
import os

from . import abi

_BASES = "ACGT"


def _log(x):
    return math.log(x) if x > 0.0 else -math.inf


class PolishModel:
    """What the hot path needs out of ``PolishParams``."""

    def __init__(self):
        self.hmm = None  # dict with type, transitions, emissions
        self.use_rle = True  # polishParams_constructEmpty default (impl/parser.c:258)
        self.use_repeat_counts = False
        self.rsm = None  # 5 x 51 x 51 flat list, index (strand ? b : 3-b) resolved at flatten time
        self.p = abi.Params.defaults()
        self.models = None

    # ---------------------------------------------------------------- parsing
    def _parse_polish(self, d):
        for k, v in d.items():
            if k == "hmmForwardStrandReadGivenReference":
                self.hmm = v
            elif k == "useRunLengthEncoding":
                self.use_rle = bool(v)
            elif k == "useRepeatCountsInAlignment":
                self.use_repeat_counts = bool(v)
            elif k == "pairwiseAlignmentParameters":
                for kk, vv in v.items():
                    if kk == "threshold":
                        self.p.threshold = float(vv)
                    elif kk == "minDiagsBetweenTraceBack":
                        self.p.min_diags_between_traceback = int(vv)
                    elif kk == "traceBackDiagonals":
                        self.p.traceback_diagonals = int(vv)
                    elif kk == "diagonalExpansion":
                        self.p.diagonal_expansion = int(vv)
                    elif kk == "splitMatrixBiggerThanThis":
                        self.p.split_matrix_bigger_than_this = int(vv)
                    elif kk == "dynamicAnchorExpansion":
                        self.p.dynamic_anchor_expansion = int(bool(vv))
            elif k == "repeatCountSubstitutionMatrix":
                if self.rsm is None:
                    self.rsm = [0.0] * (5 * abi.MAX_RL * abi.MAX_RL)  # st_calloc, repeatSubMatrix.c:62
                for kk, vv in v.items():
                    if kk in ("baseLogRepeatCounts_AT", "baseLogRepeatCounts_GC"):
                        continue
                    if len(kk) != 31 or kk[28] not in _BASES or kk[30] != "F":
                        raise ValueError("Unrecognised key in repeat sub matrix json: %s" % kk)
                    b = _BASES.index(kk[28])
                    sq = abi.MAX_RL * abi.MAX_RL
                    if len(vv) != sq:
                        raise ValueError("repeat matrix %s has %d entries" % (kk, len(vv)))
                    self.rsm[b * sq:(b + 1) * sq] = [float(z) for z in vv]

    def _read(self, path):
        with open(path) as fh:
            d = json.load(fh)
        for k, v in d.items():
            if k == "include":
                nested = v if v.startswith("/") else os.path.join(os.path.dirname(path), v)
                self._read(nested)
            elif k == "polish":
                self._parse_polish(v)

    # ------------------------------------------------------------- flattening
    def _flatten(self, strand_forward):
        h = self.hmm
        t = [float(z) for z in h["transitions"]]
        e = [float(z) for z in h["emissions"]]
        tr = lambda a, b: t[a * 3 + b]
        m = abi.Model()
        M, X, Y = 0, 1, 2
        if int(h["type"]) == 2:  # threeState: symmetric load
            m.t[abi.T_MATCH_CONTINUE] = _log(tr(M, M))
            v = _log((tr(X, M) + tr(Y, M)) / 2.0)
            m.t[abi.T_MATCH_FROM_GAP_X] = v
            m.t[abi.T_MATCH_FROM_GAP_Y] = v
            v = _log((tr(M, X) + tr(M, Y)) / 2.0)
            m.t[abi.T_GAP_OPEN_X] = v
            m.t[abi.T_GAP_OPEN_Y] = v
            v = _log((tr(X, X) + tr(Y, Y)) / 2.0)
            m.t[abi.T_GAP_EXTEND_X] = v
            m.t[abi.T_GAP_EXTEND_Y] = v
            v = _log((tr(Y, X) + tr(X, Y)) / 2.0)
            m.t[abi.T_GAP_SWITCH_TO_X] = v
            m.t[abi.T_GAP_SWITCH_TO_Y] = v
        elif int(h["type"]) == 3:  # threeStateAsymmetric
            m.t[abi.T_MATCH_CONTINUE] = _log(tr(M, M))
            m.t[abi.T_MATCH_FROM_GAP_X] = _log(tr(X, M))
            m.t[abi.T_MATCH_FROM_GAP_Y] = _log(tr(Y, M))
            m.t[abi.T_GAP_OPEN_X] = _log(tr(M, X))
            m.t[abi.T_GAP_OPEN_Y] = _log(tr(M, Y))
            m.t[abi.T_GAP_EXTEND_X] = _log(tr(X, X))
            m.t[abi.T_GAP_EXTEND_Y] = _log(tr(Y, Y))
            m.t[abi.T_GAP_SWITCH_TO_X] = _log(tr(Y, X))
            m.t[abi.T_GAP_SWITCH_TO_Y] = _log(tr(X, Y))
        else:
            raise ValueError("Unrecognised state type: %s" % h["type"])
        if int(h["emissionsType"]) != 0 or len(e) != 24:
            raise ValueError("Load from hmm: unrecognized emission type")
        match = [_log(z) for z in e[0:16]]
        gx = [_log(z) for z in e[16:20]]
        gy = [_log(z) for z in e[20:24]]
        if not strand_forward:  # nucleotideEmissions_reverseComplement
            match = [match[(3 - i) * 4 + (3 - j)] for i in range(4) for j in range(4)]
            gx = [gx[3 - i] for i in range(4)]
            gy = [gy[3 - i] for i in range(4)]
        for i in range(16):
            m.match[i] = match[i]
        for i in range(4):
            m.gap_x[i] = gx[i]
            m.gap_y[i] = gy[i]
        m.max_rl = abi.MAX_RL
        m.rle = 1 if self.use_repeat_counts else 0
        if m.rle:
            sq = abi.MAX_RL * abi.MAX_RL
            for b in range(4):
                src = b if strand_forward else 3 - b
                m.rsm[b * sq:(b + 1) * sq] = self.rsm[src * sq:(src + 1) * sq]
        return m

    def finish(self):
        if self.hmm is None:
            raise ValueError("Did not find HMM for alignment of read to a reference specified in json polish params")
        if self.use_repeat_counts:
            if not self.use_rle:
                raise ValueError("Trying to use repeat counts in read to reference alignment but not using "
                                 "run length encoding")
            if self.rsm is None:
                raise ValueError("Did not find repeat counts specified in json but trying to use repeat counts")
        if self.p.diagonal_expansion % 2 != 0:
            raise ValueError("pairwiseAlignmentParameters.diagonalExpansion must be even")
        self.models = (self._flatten(True), self._flatten(False))
        return self

    @property
    def max_repeat_count_exclusive(self):
        """maxRL handed to rleString_constructSymbolString at impl/poa.c:642-645."""
        return abi.MAX_RL if self.use_rle else 2


def load_params(path):
    """params_readParams (impl/parser.c:643-650), hot-path fields only."""
    pm = PolishModel()
    pm._read(path)
    return pm.finish()


def default_nucleotide_model():
    """stateMachine3_constructNucleotide(threeState): impl/stateMachine.c:409-432, :612-644."""
    m = abi.Model()
    m.t[abi.T_MATCH_CONTINUE] = -0.030064059121770816
    m.t[abi.T_MATCH_FROM_GAP_X] = -1.272871422049609
    m.t[abi.T_MATCH_FROM_GAP_Y] = -1.272871422049609
    m.t[abi.T_GAP_OPEN_X] = -4.21256642
    m.t[abi.T_GAP_OPEN_Y] = -4.21256642
    m.t[abi.T_GAP_EXTEND_X] = -0.3388262689231553
    m.t[abi.T_GAP_EXTEND_Y] = -0.3388262689231553
    m.t[abi.T_GAP_SWITCH_TO_X] = -4.910694825551255
    m.t[abi.T_GAP_SWITCH_TO_Y] = -4.910694825551255
    em, tv, ts = -1.8917761142, -4.3459578861, -3.760242452
    tbl = [em, tv, ts, tv, tv, em, tv, ts, ts, tv, em, tv, tv, ts, tv, em]
    for i in range(16):
        m.match[i] = tbl[i]
    for i in range(4):
        m.gap_x[i] = -1.3862943611
        m.gap_y[i] = -1.3862943611
    m.rle = 0
    m.max_rl = abi.MAX_RL
    return m


def load_flat_model(path):
    """Reads a flattened model file (2 x margin_b200_model, margin_b200_params, 3 int32 flags).

    This is the form in which a host program hands its already-parsed PolishParams to the library when the
    params JSON is not around (e.g. tests/golden/model_*.bin, produced by the reference's own parser).
    """
    import ctypes as C

    with open(path, "rb") as fh:
        raw = fh.read()
    ms, ps = C.sizeof(abi.Model), C.sizeof(abi.Params)
    if len(raw) != 2 * ms + ps + 12:
        raise ValueError("%s: unexpected size %d" % (path, len(raw)))
    pm = PolishModel()
    pm.models = (abi.Model.from_buffer_copy(raw[0:ms]), abi.Model.from_buffer_copy(raw[ms:2 * ms]))
    pm.p = abi.Params.from_buffer_copy(raw[2 * ms:2 * ms + ps])
    flags = [int.from_bytes(raw[2 * ms + ps + 4 * i:2 * ms + ps + 4 * i + 4], "little", signed=True) for i in range(3)]
    pm.use_rle, pm.use_repeat_counts = bool(flags[0]), bool(flags[1])
    pm.rsm = [] if pm.use_repeat_counts else None
    return pm
