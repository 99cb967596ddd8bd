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


from margin_b200.replay import load_dump  # noqa: E402,F401
