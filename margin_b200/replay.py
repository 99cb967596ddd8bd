"""Batches dumped by the margin-side binding (MARGIN_B200_DUMP=<prefix>, compat/pairwiseAligner_b200.c: every posterior
batch `margin polish` sends to the library, with its models and parameters) and how to read them back: the input of the
real-data parity test (tests/test_polish_e2e.py) and of `bench.py --workload c1_replay`."""
import ctypes as C
import glob
import os
import subprocess

import numpy as np


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


def dump_bundled_region(out_prefix, ref_dir, threads=2, timeout=900):
    """Runs the reference CLI linked against this library (oracle/_ref/margin_b200cli) on the bundled ONT region
    (BASELINE.json configs[0]) with MARGIN_B200_DUMP set; returns the dump files in call order per thread."""
    data = os.path.join(ref_dir, "data")
    cmd = [os.path.join(ref_dir, "margin_b200cli"), "polish", os.path.join(data, "HG002.r94g360.chr20_59M_100k.bam"),
           os.path.join(data, "hg38.chr20_59M_100k.fa"),
           os.path.join(data, "params", "polish", "ont", "r9.4", "allParams.np.human.r94-g360.json"),
           "-t", str(threads), "-o", out_prefix + ".polish"]
    env = dict(os.environ, MARGIN_B200_DUMP=out_prefix)
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, env=env, timeout=timeout)
    if r.returncode != 0:
        raise RuntimeError("margin_b200cli failed: " + r.stdout[-2000:])

    def order(path):
        t, n = os.path.basename(path).split(".")[-3:-1]
        return int(t), int(n)
    return sorted(glob.glob(out_prefix + ".*.*.bin"), key=order)
