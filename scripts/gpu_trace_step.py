"""Where a resident bench step spends its time: the 100 kb chunk with / without re-upload, L2 flush, per-launch events,
and the four bench chunks in rotation.  Prints the library's own device time (first kernel to last) per step."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from margin_b200 import abi, model
from margin_b200.aligner import PairHmmAligner
pm = model.load_flat_model(bench.MODEL_FILE)
chunks = [abi.Batch(bench.build_chunk(pm, 42, k), compact=True).to_pinned() for k in range(4)]
al = PairHmmAligner(pm.models, pm.p, device=0)
al.set_pair_format(abi.PAIRS_DELTA32)
flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda:0")


def run(name, which, upload, do_flush, events, n=12):
    al.set_kernel_timing(events)
    ms, per = [], {}
    for i in range(n + 4):
        k = which[i % len(which)]
        if upload or i == 0:
            al.upload(chunks[k])
        if do_flush:
            flush.zero_()
            torch.cuda.synchronize()
        st = al.posteriors_resident(True)
        if i >= 4:
            ms.append(st["gpu_ms"])
            per.setdefault(k, []).append(st["gpu_ms"])
    kt = {a: round(b[0] / n, 3) for a, b in al.kernel_times().items()} if events else {}
    print("%-34s %.3f ms/step  per chunk %s  %s" % (name, sum(ms) / len(ms), {k: round(sum(v) / len(v), 3) for k, v in per.items()}, kt))


run("chunk 1, resident, repeated", [1], False, False, False)
run("chunk 1, re-upload", [1], True, False, False)
run("chunk 1, re-upload + L2 flush", [1], True, True, False)
run("chunk 1, + per-launch events", [1], True, True, True)
run("4 chunks in rotation", [0, 1, 2, 3], True, False, False)
run("4 chunks + flush", [0, 1, 2, 3], True, True, False)
run("4 chunks + flush + events", [0, 1, 2, 3], True, True, True)
