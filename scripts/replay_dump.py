"""Replays batches dumped by the margin-side binding (MARGIN_B200_DUMP=<prefix>, compat/pairwiseAligner_b200.c)
through the library under several settings and compares every tuple list with the CPU oracle (test infrastructure;
this script is a debugging aid, not part of the product path).

    MARGIN_B200_DUMP=/tmp/d oracle/_ref/margin_b200cli polish ... ; python scripts/replay_dump.py /tmp/d.0.0.bin
"""
import ctypes as C
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from margin_b200 import abi  # noqa: E402
from margin_b200.aligner import PairHmmAligner  # noqa: E402
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
from common import load_dump  # noqa: E402


def score(lists):
    return sum(int(l[0]["w"].sum()) for l in lists)


def compare(name, got, want, tol):
    bad = 0
    worst = 0
    for i, (g, w) in enumerate(zip(got, want)):
        for l in range(3):
            if len(g[l]) != len(w[l]) or not (np.array_equal(g[l]["x"], w[l]["x"]) and np.array_equal(g[l]["y"], w[l]["y"])):
                sg = set(zip(g[l]["x"].tolist(), g[l]["y"].tolist()))
                sw = set(zip(w[l]["x"].tolist(), w[l]["y"].tolist()))
                miss, extra = sw - sg, sg - sw
                if bad < 12:
                    ys = sorted(p[1] for p in miss | extra)
                    print("  %s: alignment %d list %d: %d vs %d tuples, %d missing %d extra, y range %s" % (
                        name, i, l, len(g[l]), len(w[l]), len(miss), len(extra), (ys[0], ys[-1]) if ys else None))
                bad += 1
                continue
            d = np.abs(g[l]["w"].astype(np.int64) - w[l]["w"].astype(np.int64)).max(initial=0)
            worst = max(worst, int(d))
            if d > tol:
                if bad < 12:
                    k = int(np.argmax(np.abs(g[l]["w"].astype(np.int64) - w[l]["w"].astype(np.int64))))
                    print("  %s: alignment %d list %d: weight differs by %d at (%d, %d)" % (name, i, l, d, g[l]["x"][k], g[l]["y"][k]))
                bad += 1
    print("%s: %d lists differ, largest weight difference %d (1e-7 units)" % (name, bad, worst))
    return bad


def main():
    path = sys.argv[1]
    items, models, params, want = load_dump(path)
    lens = [len(it[1]) for it in items]
    print("%s: %d alignments, read length median %d max %d, %d anchors, want_indels=%d" % (
        path, len(items), int(np.median(lens)), max(lens), sum(len(it[2]) for it in items), want))
    from oracle.oracle_py import Oracle
    t0 = time.time()
    batch = abi.Batch(items)
    ref, tot = Oracle().posteriors_batch(models, params, batch, want, threads=os.cpu_count() or 1)
    print("oracle: %.1f s, %d cells, score %d" % (time.time() - t0, tot["cells"], score(ref)))
    settings = [("plain triples", dict(MARGIN_B200_PIPELINE="0"), False, abi.PAIRS_TRIPLES),
                ("plain compact+packed", dict(MARGIN_B200_PIPELINE="0"), True, abi.PAIRS_PACKED64),
                ("pipelined compact+packed", dict(MARGIN_B200_PIPELINE_FORCE="1"), True, abi.PAIRS_PACKED64),
                ("pipelined again (warm)", None, True, abi.PAIRS_PACKED64)]
    al = None
    for name, env, compact, fmt in settings:
        if env is not None:
            for k in ("MARGIN_B200_PIPELINE", "MARGIN_B200_PIPELINE_FORCE"):
                os.environ.pop(k, None)
            os.environ.update(env)
            if al is not None:
                al.close()
            al = PairHmmAligner(models, params, device=0)  # the environment is read when a context is created
        al.set_pair_format(fmt)
        b = abi.Batch(items, compact=compact)
        got = al.posteriors_batch(b, bool(want))
        print("%s: compact=%s score %d tiles %d" % (name, b.compact, score(got.lists), got.stats["tiles"]))
        compare(name, got.lists, ref, tol=1000)
    al.close()


if __name__ == "__main__":
    main()
