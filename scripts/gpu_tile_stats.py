"""Tile statistics of the four bench chunks (MARGIN_B200_TILE_STATS=1) and the device time of a resident step."""
import os, sys
os.environ["MARGIN_B200_TILE_STATS"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from margin_b200 import abi, model
from margin_b200.aligner import PairHmmAligner
pm = model.load_flat_model(bench.MODEL_FILE)
al = PairHmmAligner(pm.models, pm.p, device=0)
al.set_pair_format(abi.PAIRS_DELTA32)
for k in range(4):
    b = abi.Batch(bench.build_chunk(pm, 42, k), compact=True).to_pinned()
    al.upload(b)
    for i in range(3):
        st = al.posteriors_resident(True)
    sys.stderr.write("chunk %d: %d reads, %d cells, %.3f ms\n" % (k, b.n, st["band_cells"], st["gpu_ms"]))
