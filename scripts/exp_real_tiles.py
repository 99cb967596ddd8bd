"""Debugging aid: one alignment of a dumped batch (scripts/replay_dump.py) under several warm-up lengths."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from margin_b200 import abi  # noqa: E402
from margin_b200.aligner import PairHmmAligner  # noqa: E402
from replay_dump import load_dump, compare  # noqa: E402
from oracle.oracle_py import Oracle  # noqa: E402

items, models, params, want = load_dump(sys.argv[1])
idx = [int(v) for v in sys.argv[2].split(",")]
yc = int(sys.argv[3])
sub = [items[i] for i in idx]
ref, _ = Oracle().posteriors_batch(models, params, abi.Batch(sub), want, threads=4)
it = sub[0]
a = it[2]
k = int(np.searchsorted(a[:, 1], yc))
print("lx %d ly %d anchors %d; anchors around y=%d:" % (len(it[0]), len(it[1]), len(a), yc))
print(a[max(0, k - 6):k + 6].tolist())
al = PairHmmAligner(models, params, device=0)
kl, kr = al.bands(abi.Batch(sub))[0]
d0 = int(a[max(0, k - 1), 0] + a[max(0, k - 1), 1])
w = (kr - kl) // 2 + 1
print("band widths (cells) on diagonals %d..%d:" % (d0 - 40, d0 + 300), w[d0 - 40:d0 + 300:10].tolist(), "max over job", int(w.max()))
al.close()
for W in (128, 256, 512, 1024, 2048):
    for L in (0, 400):
        os.environ["MARGIN_B200_WARMUP"] = str(W)
        os.environ["MARGIN_B200_PIPELINE"] = "0"
        if L:
            os.environ["MARGIN_B200_TILE_LEN"] = str(L)
        else:
            os.environ.pop("MARGIN_B200_TILE_LEN", None)
        al = PairHmmAligner(models, params, device=0)
        got = al.posteriors_batch(abi.Batch(sub), bool(want))
        compare("W=%d L=%s tiles=%d" % (W, L or "auto", got.stats["tiles"]), got.lists, ref, tol=1000)
        al.close()
