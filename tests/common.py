"""Helpers shared by the parity tests."""
import ctypes as C
import os

import numpy as np

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load_golden_cases(name):
    z = np.load(os.path.join(GOLDEN, "posteriors_%s.npz" % name))
    cases = []
    for i in range(int(z["ncase"][0])):
        meta = z["meta%d" % i]
        cases.append(dict(sx=z["sx%d" % i], sy=z["sy%d" % i], a=z["a%d" % i], slot=int(meta[0]), rl=int(meta[1]),
                          rr=int(meta[2]), min_diags=int(meta[3]), dynamic=int(meta[4]),
                          lists=[z["l%d_%d" % (l, i)] for l in range(3)], fwd=float(z["fwd%d" % i][0]),
                          kl=z["kl%d" % i], kr=z["kr%d" % i]))
    return cases


def compare_lists(got, want, tol_w=1):
    """Tuple lists must have identical (x, y) sequences in identical order; weights may differ by at most
    tol_w units of 1e-7 (device exp() vs glibc exp() can differ in the last ulp; 0 = bit-exact).
    Returns the number of weights that differ."""
    ndiff = 0
    for l in range(3):
        g, w = got[l], want[l]
        assert len(g) == len(w), "list %d: %d tuples vs %d" % (l, len(g), len(w))
        assert np.array_equal(g["x"], w["x"]) and np.array_equal(g["y"], w["y"]), "list %d: coordinates differ" % l
        dw = np.abs(g["w"].astype(np.int64) - w["w"].astype(np.int64))
        assert dw.max(initial=0) <= tol_w, "list %d: weight differs by %d" % (l, dw.max())
        ndiff += int((dw > 0).sum())
    return ndiff


def load_dump(path):
    """A posterior batch written by the margin-side binding under MARGIN_B200_DUMP (compat/pairwiseAligner_b200.c):
    (items, models, params, want_indels)."""
    from margin_b200 import abi
    raw = open(path, "rb").read()
    head = np.frombuffer(raw, np.int64, 8)
    assert head[0] == 0x42323030, "not a margin_b200 dump"
    n, want, tx, ty, ta, sz_model, sz_params = (int(v) for v in head[1:8])
    assert sz_model == C.sizeof(abi.Model) and sz_params == C.sizeof(abi.Params)
    o = 64
    params = abi.Params.from_buffer_copy(raw[o:o + sz_params]); o += sz_params
    models = []
    for _ in range(2):
        models.append(abi.Model.from_buffer_copy(raw[o:o + sz_model])); o += sz_model

    def take(dt, count):
        nonlocal o
        a = np.frombuffer(raw, dt, count, o).copy()
        o += a.nbytes
        return a
    off_x, off_y, off_a = take(np.int64, n + 1), take(np.int64, n + 1), take(np.int64, n + 1)
    slot, rl, rr = take(np.uint8, n), take(np.uint8, n), take(np.uint8, n)
    sx, sy = take(np.uint16, tx), take(np.uint16, ty)
    anc = take(np.int32, 3 * ta).reshape(-1, 3)
    items = [(sx[off_x[i]:off_x[i + 1]], sy[off_y[i]:off_y[i + 1]], anc[off_a[i]:off_a[i + 1]], int(slot[i]),
              int(rl[i]), int(rr[i])) for i in range(n)]
    return items, models, params, want
