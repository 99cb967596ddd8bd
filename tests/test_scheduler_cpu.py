"""Host logic of the multi-GPU path on CPU: world_size-2 gloo processes shard chunks with no data-path collective
(SURVEY.md §8 E-1).  The per-chunk "aligner" here is a stand-in checksum so that only the scheduling, the
per-context threading and the rank-0 gather are under test; the kernels are covered by the -m gpu tests."""
import os
import socket
import sys

import numpy as np
import pytest

from margin_b200 import scheduler

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_order_is_size_descending_like_the_reference():
    # polish.c:456-462 sorts chunks by estimated depth, largest first
    assert scheduler.order_chunks([5, 9, 9, 1]) == [1, 2, 0, 3]


def test_shard_covers_every_chunk_once_and_balances():
    rng = np.random.default_rng(3)
    sizes = rng.integers(1, 1000, 50).tolist()
    for world in (1, 2, 4, 8):
        parts = scheduler.shard_chunks(sizes, world)
        flat = sorted(i for p in parts for i in p)
        assert flat == list(range(50))
        loads = [sum(sizes[i] for i in p) for p in parts]
        assert max(loads) - min(loads) <= max(sizes)  # LPT bound
        for p in parts:  # each rank processes its own share largest first
            assert [sizes[i] for i in p] == sorted((sizes[i] for i in p), reverse=True)
    assert scheduler.shard_chunks([], 2) == [[], []]
    with pytest.raises(ValueError):
        scheduler.shard_chunks([1], 0)


def _chunk_items(i):
    rng = np.random.default_rng(100 + i)
    n = 3 + i % 4
    return [(rng.integers(0, 4, 20 + 7 * i).astype(np.uint16), rng.integers(0, 4, 25 + 5 * i).astype(np.uint16),
             np.zeros((0, 3), np.int32), 0, 0, 0) for _ in range(n)]


class _FakeAligner:
    live = 0
    peak = 0

    def __init__(self, device):
        self.device = device
        _FakeAligner.live += 1
        _FakeAligner.peak = max(_FakeAligner.peak, _FakeAligner.live)

    def close(self):
        _FakeAligner.live -= 1


def _run(al, index, items):
    return (index, len(items), int(sum(int(it[0].sum()) + int(it[1].sum()) for it in items)))


def test_scheduler_runs_chunks_through_several_contexts():
    _FakeAligner.live = _FakeAligner.peak = 0
    s = scheduler.ChunkScheduler(_FakeAligner, _run, device=0, contexts_per_gpu=3)
    res = s.process(list(range(9)), _chunk_items)
    assert sorted(res) == list(range(9))
    for i in range(9):
        assert res[i] == _run(None, i, _chunk_items(i))
    assert _FakeAligner.live == 0 and 1 <= _FakeAligner.peak <= 3


def test_scheduler_errors_are_loud():
    def boom(al, i, items):
        if i == 2:
            raise RuntimeError("kernel failed on chunk 2")
        return i
    s = scheduler.ChunkScheduler(_FakeAligner, boom, contexts_per_gpu=2)
    with pytest.raises(RuntimeError, match="chunk 2"):
        s.process([0, 1, 2, 3], _chunk_items)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _gloo_worker(rank, world, port, n_chunks, out_dir):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        sizes = [scheduler.estimate_chunk_size(_chunk_items(i)) for i in range(n_chunks)]
        mine = scheduler.shard_chunks(sizes, world)[rank]
        s = scheduler.ChunkScheduler(_FakeAligner, _run, device=rank, contexts_per_gpu=2)
        local = s.process(mine, _chunk_items)
        merged = scheduler.gather_in_chunk_order(dist, local, n_chunks)
        if rank == 0:
            np.save(os.path.join(out_dir, "merged.npy"), np.array(merged, dtype=np.int64))
        np.save(os.path.join(out_dir, "mine%d.npy" % rank), np.array(mine, dtype=np.int64))
        dist.barrier()
    finally:
        dist.destroy_process_group()


def test_two_ranks_shard_chunks_over_gloo(tmp_path):
    import torch.multiprocessing as mp
    n_chunks, world = 11, 2
    mp.spawn(_gloo_worker, args=(world, _free_port(), n_chunks, str(tmp_path)), nprocs=world, join=True)
    merged = np.load(tmp_path / "merged.npy")
    assert merged.shape == (n_chunks, 3)
    for i in range(n_chunks):
        assert tuple(merged[i]) == _run(None, i, _chunk_items(i))
    a, b = np.load(tmp_path / "mine0.npy"), np.load(tmp_path / "mine1.npy")
    assert sorted(a.tolist() + b.tolist()) == list(range(n_chunks))
    assert len(a) and len(b)
