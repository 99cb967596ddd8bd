"""The chunk scheduler with the real aligner on cuda:0: several contexts take turns on the device (compute token
inside margin_b200_posteriors_batch) and every chunk's lists still match the oracle."""
import numpy as np
import pytest

import gen
from common import compare_lists
from margin_b200 import abi, scheduler
from margin_b200.aligner import PairHmmAligner

pytestmark = pytest.mark.gpu


def _chunk(i):
    rng = np.random.default_rng(900 + i)
    return [gen.random_case(rng, int(rng.integers(50, 900))) for _ in range(5 + i % 3)]


def test_chunks_through_three_contexts_match_the_oracle(oracle, golden_model):
    pm = golden_model("r94g360")
    p = pm.p.copy(min_diags_between_traceback=300)
    n_chunks = 7
    sizes = [scheduler.estimate_chunk_size(_chunk(i)) for i in range(n_chunks)]
    mine = scheduler.shard_chunks(sizes, 1)[0]
    assert sorted(mine) == list(range(n_chunks)) and sizes[mine[0]] == max(sizes)

    def make(device):
        al = PairHmmAligner(pm.models, p, device=device)
        al.set_pair_format(abi.PAIRS_PACKED64)
        return al

    def run(al, index, items):
        return al.posteriors_batch(abi.Batch(items, compact=True), True)

    res = scheduler.ChunkScheduler(make, run, device=0, contexts_per_gpu=3).process(mine, _chunk)
    merged = scheduler.gather_in_chunk_order(None, res, n_chunks)
    for i in range(n_chunks):
        items = _chunk(i)
        assert len(merged[i]) == len(items)
        for k, it in enumerate(items):
            want = oracle.posteriors(pm.models[it[3]], p, it[0], it[1], it[2], it[4], it[5], 1)
            compare_lists(merged[i][k], want, tol_w=1)
