"""Generates the committed fixtures from the UNMODIFIED reference (oracle/_ref, built from /root/reference).

Run in the build container only:  python tests/golden/make_golden.py
* model_<name>.bin      the reference parser's state machines for a shipped params file, flattened
                        (2 x margin_b200_model + margin_b200_params + 3 int32 flags)
* posteriors_<name>.npz seeded alignment problems with the reference's own outputs (three lists in stList order,
                        forward log-probabilities, band geometry)
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", ".."))
sys.path.insert(0, os.path.join(HERE, ".."))

import gen  # noqa: E402
from oracle.oracle_py import Reference  # noqa: E402

PARAMS = {
    "r94g360": "/root/reference/params/polish/ont/r9.4/allParams.np.human.r94-g360.json",
    "r94g344": "/root/reference/params/polish/ont/r9.4/allParams.np.human.r94-g344.json",
    "hifi": "/root/reference/params/polish/pacbio/hifi/allParams.hifi.json",
    "no_rle": "/root/reference/params/polish/misc/allParams.no_rle.json",
}


def main():
    R = Reference()
    for name, path in PARAMS.items():
        if not os.path.exists(path):
            print("skip", name)
            continue
        ref = R.load_params(path)
        with open(os.path.join(HERE, "model_%s.bin" % name), "wb") as fh:
            fh.write(bytes(ref["models"][0]))
            fh.write(bytes(ref["models"][1]))
            fh.write(bytes(ref["p"]))
            fh.write(np.array([ref["use_rle"], ref["use_repeat_counts"], ref["max_rl"]], np.int32).tobytes())
        rng = np.random.default_rng(20260104)
        out = {}
        ncase = 24
        for i in range(ncase):
            lx = int(rng.integers(0, 700)) if i else 0
            rle = bool(ref["use_repeat_counts"])
            sx, sy, a, slot, rl, rr = gen.random_case(rng, lx, rle=rle, n_frac=0.02)
            p = ref["p"].copy(min_diags_between_traceback=int(rng.choice([5000, 100, 150])))
            if i % 5 == 4:
                p.dynamic_anchor_expansion = 0
            lists = R.posteriors(ref["sms"][slot], p, sx, sy, a, rl, rr, 1)
            fwd = R.forward(ref["sms"][slot], p, sx[:60], sy[:60], (), rl, rr)
            kl, kr = R.band(a, len(sx), len(sy), p.dynamic_anchor_expansion, p.diagonal_expansion)
            out["sx%d" % i], out["sy%d" % i], out["a%d" % i] = sx, sy, a
            out["meta%d" % i] = np.array([slot, rl, rr, p.min_diags_between_traceback, p.dynamic_anchor_expansion], np.int64)
            for l in range(3):
                out["l%d_%d" % (l, i)] = lists[l]
            out["fwd%d" % i] = np.array([fwd])
            out["kl%d" % i], out["kr%d" % i] = kl.astype(np.int32), kr.astype(np.int32)
        out["ncase"] = np.array([ncase])
        np.savez_compressed(os.path.join(HERE, "posteriors_%s.npz" % name), **out)
        print("wrote", name)


if __name__ == "__main__":
    main()
