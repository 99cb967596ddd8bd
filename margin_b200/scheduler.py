"""Chunk scheduler: shards independent polishing chunks over the GPUs of one box.

The reference polishes chunks in an OpenMP ``schedule(dynamic,1)`` loop over chunks ordered by estimated depth,
largest first (impl/polish.c:456-462 ``shuffleChunksMethod: size_desc``, loop at :476); chunks share only
read-only parameters and meet again in the host-side stitcher (:926).  Here the unit is the same and there is
**no data-path collective**:

* across processes (one per GPU, ``torch.distributed``): :func:`shard_chunks` is the static equivalent of the
  reference's dynamic schedule — longest-processing-time-first onto the least loaded rank, deterministic, so
  every rank derives the same assignment from the chunk sizes alone and nothing has to be exchanged;
* inside a process: :class:`ChunkScheduler` keeps ``contexts_per_gpu`` aligner contexts (each owns a stream and its
  pinned/device arenas, include/margin_b200.h) busy from a size-descending queue, so a second chunk's host work
  and copies overlap the first one's kernels;
* :func:`gather_in_chunk_order` brings small per-chunk summaries (not tuple lists) to rank 0 for the stitcher's
  bookkeeping — control plane only (``all_gather_object``).

Nothing here computes alignments: a ``run(aligner, chunk_index, items)`` callable does, on the GPU.
"""
import queue
import threading


def order_chunks(sizes):
    """Chunk indices, largest estimated size first, ties by index (polish.c:456-462)."""
    return sorted(range(len(sizes)), key=lambda i: (-int(sizes[i]), i))


def shard_chunks(sizes, world_size):
    """[rank] -> chunk indices (each list in processing order).  Greedy LPT: the next-largest chunk goes to the
    rank with the least work so far (ties: lowest rank) — what a dynamic schedule over equally fast workers does."""
    if world_size < 1:
        raise ValueError("world_size must be >= 1")
    load = [0] * world_size
    out = [[] for _ in range(world_size)]
    for i in order_chunks(sizes):
        r = min(range(world_size), key=lambda k: (load[k], k))
        out[r].append(i)
        load[r] += max(int(sizes[i]), 1)
    return out


def estimate_chunk_size(items):
    """Work estimate of a chunk before any band exists: summed diagonals (lX + lY) of its alignments."""
    return int(sum(len(it[0]) + len(it[1]) for it in items))


class ChunkScheduler:
    """Runs this rank's chunks through ``contexts_per_gpu`` aligner contexts on one device.

    ``make_aligner(device)`` builds one context (PairHmmAligner in production); ``run(aligner, index, items)``
    processes one chunk and returns its result.  Results come back as {chunk index: result}.  An exception in any
    worker stops the queue and is re-raised (errors are loud; there is no CPU fallback to hide behind).
    """

    def __init__(self, make_aligner, run, device=0, contexts_per_gpu=2):
        if contexts_per_gpu < 1:
            raise ValueError("contexts_per_gpu must be >= 1")
        self.make_aligner = make_aligner
        self.run = run
        self.device = device
        self.contexts_per_gpu = contexts_per_gpu

    def process(self, chunk_indices, load_chunk):
        """``load_chunk(index)`` -> items of that chunk (called on the worker thread, so input preparation of one
        chunk overlaps the kernels of another)."""
        q = queue.Queue()
        for i in chunk_indices:
            q.put(i)
        results, errors = {}, []
        lock = threading.Lock()

        def worker():
            al = None
            try:
                al = self.make_aligner(self.device)
                while not errors:
                    try:
                        i = q.get_nowait()
                    except queue.Empty:
                        break
                    r = self.run(al, i, load_chunk(i))
                    with lock:
                        results[i] = r
            except BaseException as e:  # noqa: BLE001 - re-raised on the caller's thread
                with lock:
                    errors.append(e)
            finally:
                if al is not None and hasattr(al, "close"):
                    al.close()

        n = max(1, min(self.contexts_per_gpu, len(chunk_indices)))
        threads = [threading.Thread(target=worker, daemon=True) for _ in range(n)]
        for t in threads:
            t.start()
        for t in threads:
            t.join()
        if errors:
            raise errors[0]
        return results


def gather_in_chunk_order(dist, local_results, n_chunks):
    """All ranks' {chunk: summary} merged into a list in chunk order (every rank gets it; rank 0 stitches).
    ``dist`` is ``torch.distributed`` (gloo or nccl) or None for a single process."""
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        parts = [local_results]
    else:
        parts = [None] * dist.get_world_size()
        dist.all_gather_object(parts, local_results)
    merged = {}
    for p in parts:
        for k, v in p.items():
            if k in merged:
                raise RuntimeError("chunk %d was processed by two ranks" % k)
            merged[k] = v
    missing = [i for i in range(n_chunks) if i not in merged]
    if missing:
        raise RuntimeError("chunks never processed: %s" % missing[:8])
    return [merged[i] for i in range(n_chunks)]
