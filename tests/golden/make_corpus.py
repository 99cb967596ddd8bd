"""Digests of the parity corpus (tests/gen.py: corpus_case) from the UNMODIFIED reference (oracle/_ref, built from
/root/reference).  Run in the build container only:  python tests/golden/make_corpus.py

Writes tests/golden/corpus_digests.npz: per case and list (match, gapX, gapY) the tuple count, an order-sensitive
hash of the (x, y) sequence and the sum of the weights, plus the band-cell count of the case.  The inputs are not
stored: every case is a pure function of its index."""
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", ".."))
sys.path.insert(0, os.path.join(HERE, ".."))

import gen  # noqa: E402
from margin_b200 import abi, model  # noqa: E402
from oracle.oracle_py import Reference  # noqa: E402


def main():
    R = Reference()
    pm = model.load_flat_model(os.path.join(HERE, "model_r94g360.bin"))
    sms = [R.sm_from_model(pm.models[s]) for s in (0, 1)]
    n = gen.CORPUS_SIZE
    cnt = np.zeros((n, 3), np.int64)
    hsh = np.zeros((n, 3), np.uint64)
    wsum = np.zeros((n, 3), np.int64)
    t0 = time.time()
    by_ps = {}
    for i in range(n):
        item, ps = gen.corpus_case(i)
        by_ps.setdefault(ps, []).append((i, item))
    from concurrent.futures import ThreadPoolExecutor
    for ps, cases in sorted(by_ps.items()):
        md, dyn = gen.CORPUS_PARAM_SETS[ps]
        p = pm.p.copy(min_diags_between_traceback=md, dynamic_anchor_expansion=dyn)

        def one(case):
            i, it = case
            ls = R.posteriors(sms[it[3]], p, it[0], it[1], it[2], it[4], it[5], 1)  # ctypes call: releases the GIL
            return i, [gen.list_digest(ls[l]) for l in range(3)]
        with ThreadPoolExecutor(os.cpu_count() or 1) as pool:
            for i, dg in pool.map(one, cases):
                for l in range(3):
                    cnt[i, l], h, wsum[i, l] = dg[l]
                    hsh[i, l] = np.uint64(h)
        print("parameter set %d: %d cases, %.0f s" % (ps, len(cases), time.time() - t0), flush=True)
    np.savez_compressed(os.path.join(HERE, "corpus_digests.npz"), count=cnt, xy_hash=hsh, weight_sum=wsum,
                        size=np.array([n]))
    print("wrote %d cases, %d tuples" % (n, int(cnt.sum())))


if __name__ == "__main__":
    main()
