"""Diagnostic: where do tiled-mode weights differ from sequential mode?"""
import os, sys, subprocess
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))

def run(mode, warm, tile_len, tag):
    os.environ["MARGIN_B200_MODE"] = mode
    os.environ["MARGIN_B200_WARMUP"] = str(warm)
    os.environ["MARGIN_B200_TILE_LEN"] = str(tile_len)
    import gen
    from margin_b200 import model
    from margin_b200.aligner import PairHmmAligner
    pm = model.load_flat_model(os.path.join(ROOT, "tests", "golden", "model_r94g360.bin"))
    rng = np.random.default_rng(5)
    items = []
    while len(items) < 32:
        it = gen.random_case(rng, int(rng.integers(1500, 4000)), anchor_keep=0.9, expansions=(4,), ragged=False)
        if len(it[2]) > 100:
            items.append(it)
    al = PairHmmAligner(pm.models, pm.p, 0)
    res = al.posteriors_batch(items, True)
    np.savez("/tmp/diag_%s.npz" % tag, **{"l%d_%d" % (l, i): res[i][l] for i in range(len(items)) for l in range(3)})
    print(tag, res.stats, flush=True)
    al.close()

if __name__ == "__main__":
    if len(sys.argv) > 1:
        run(sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), sys.argv[4])
    else:
        for args in (("sequential", 64, 500, "seq"), ("tiles", 64, 1000000, "nocut"), ("tiles", 64, 500, "t500")):
            subprocess.check_call([sys.executable, __file__] + [str(a) for a in args])
        base = np.load("/tmp/diag_seq.npz")
        for tag in ("nocut", "t500"):
            z = np.load("/tmp/diag_%s.npz" % tag)
            tot = nd = 0
            hist = np.zeros(50, np.int64)
            for i in range(32):
                a, b = base["l0_%d" % i], z["l0_%d" % i]
                assert np.array_equal(a["x"], b["x"])
                dw = b["w"].astype(np.int64) - a["w"].astype(np.int64)
                tot += len(dw); nd += int((dw != 0).sum())
                dg = (a["x"] + a["y"] + 2)[dw != 0]
                hist += np.bincount((dg % 500) // 10, minlength=50)
                if i == 0:
                    idx = np.flatnonzero(dw != 0)[:12]
                    print(tag, "first diffs (x,y,w_seq,dw):", [(int(a["x"][k]), int(a["y"][k]), int(a["w"][k]), int(dw[k])) for k in idx])
            print(tag, "match tuples", tot, "differ", nd, "hist of (x+y+2)%500 //10:", hist.tolist(), flush=True)
