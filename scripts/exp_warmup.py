"""Experiment: tiled mode vs sequential mode on the GPU for several warm-up lengths (weights that differ, max |dw|)."""
import os, sys, subprocess, json
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))

def run(mode, warm, tile_len):
    os.environ["MARGIN_B200_MODE"] = mode
    os.environ["MARGIN_B200_WARMUP"] = str(warm)
    os.environ["MARGIN_B200_TILE_LEN"] = str(tile_len)
    import gen
    from margin_b200 import model
    from margin_b200.aligner import PairHmmAligner
    pm = model.load_flat_model(os.path.join(ROOT, "tests", "golden", "model_r94g360.bin"))
    rng = np.random.default_rng(5)
    items = [gen.random_case(rng, int(rng.integers(1500, 6000)), anchor_keep=0.9, expansions=(4,), ragged=False) for _ in range(48)]
    al = PairHmmAligner(pm.models, pm.p, 0)
    res = al.posteriors_batch(items, True)
    out = [np.concatenate([r[l]["w"] for r in res.lists]) for l in range(3)]
    xy = [np.concatenate([r[l]["x"] for r in res.lists]) for l in range(3)]
    al.close()
    return out, xy, res.stats

if __name__ == "__main__":
    if len(sys.argv) > 1:
        mode, warm, tl = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
        out, xy, st = run(mode, warm, tl)
        np.savez("/tmp/exp_%s_%d_%d.npz" % (mode, warm, tl), w0=out[0], w1=out[1], w2=out[2], x0=xy[0], x1=xy[1], x2=xy[2])
        print(mode, warm, tl, st)
    else:
        subprocess.check_call([sys.executable, __file__, "sequential", "64", "500"])
        base = np.load("/tmp/exp_sequential_64_500.npz")
        for tl in (500, 200):
            for warm in (16, 32, 64, 128, 256):
                subprocess.check_call([sys.executable, __file__, "tiles", str(warm), str(tl)])
                z = np.load("/tmp/exp_tiles_%d_%d.npz" % (warm, tl))
                rep = []
                for l in range(3):
                    if len(z["w%d" % l]) != len(base["w%d" % l]) or not np.array_equal(z["x%d" % l], base["x%d" % l]):
                        rep.append("list%d: SET DIFFERS (%d vs %d)" % (l, len(z["w%d" % l]), len(base["w%d" % l])))
                        continue
                    dw = np.abs(z["w%d" % l].astype(np.int64) - base["w%d" % l].astype(np.int64))
                    rep.append("list%d: n=%d ndiff=%d max=%d" % (l, len(dw), int((dw > 0).sum()), int(dw.max(initial=0))))
                print("tile_len", tl, "warmup", warm, "|", " ; ".join(rep), flush=True)
