"""margin_b200_forward_tiny_noalloc (include/margin_b200.h): the forward-only likelihoods of small matrices with the
scratch in a fixed device array inside the library and no allocation — an entry point without a counterpart in the
reference.  Collected last on purpose.

Checked through the C ABI against margin_b200_forward_batch (the existing entry point, itself bit-exact against the oracle
in tests/test_gpu_parity.py) on two calls in a row that share the scratch array: the second batch is smaller and
different, so stale contents of the first call would show.  Tolerance: none, as for every FP64 forward likelihood here."""
import numpy as np
import pytest

import gen
from margin_b200 import abi
from margin_b200.aligner import AlignerError, PairHmmAligner

pytestmark = pytest.mark.gpu


def _bubble_items(rng, n, max_len=31):
    items = []
    for _ in range(n):
        lx = int(rng.integers(0, 44))
        ly = int(rng.integers(0, max_len + 1)) if lx > max_len else int(rng.integers(0, 44))  # shorter side <= 31
        sx = gen.random_symbols(rng, lx, True, 0.02)
        sy = gen.random_symbols(rng, ly, True, 0.02)
        items.append((sx, sy, np.zeros((0, 3), np.int32), int(rng.integers(0, 2)), 0, 0))
    return items


def _run_noalloc(al, items):
    import torch
    b = abi.Batch(items)
    dev = torch.device("cuda", 0)
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a).view(np.int16 if a.dtype == np.uint16 else a.dtype)).to(dev)
    sx, sy, ox, oy, slot = t(b.sym_x), t(b.sym_y), t(b.off_x), t(b.off_y), t(b.model_slot)
    out = torch.full((b.n,), float("nan"), dtype=torch.float64, device=dev)
    torch.cuda.synchronize(dev)
    al.forward_tiny_noalloc(b.n, sx.data_ptr(), ox.data_ptr(), int(b.off_x[b.n]), sy.data_ptr(), oy.data_ptr(), int(b.off_y[b.n]),
                            slot.data_ptr(), out.data_ptr())
    return out.cpu().numpy()


def test_no_alloc_forward_matches_forward_batch_on_two_calls_in_a_row(golden_model):
    pm = golden_model("r94g360")
    rng = np.random.default_rng(20261105)
    first, second = _bubble_items(rng, 3000), _bubble_items(rng, 700)
    al = PairHmmAligner(pm.models, pm.p, device=0)
    want = [al.forward_batch(first), al.forward_batch(second)]
    got = [_run_noalloc(al, first), _run_noalloc(al, second)]
    for g, w in zip(got, want):
        assert g.shape == w.shape and not np.isnan(g).any()
        assert np.array_equal(g, w)
    # and once more in the other order, on a fresh context (the scratch array outlives contexts)
    al.close()
    al = PairHmmAligner(pm.models, pm.p, device=0)
    assert np.array_equal(_run_noalloc(al, second), want[1])
    assert np.array_equal(_run_noalloc(al, first), want[0])
    al.close()


def test_no_alloc_forward_rejects_what_does_not_fit(golden_model):
    pm = golden_model("r94g360")
    al = PairHmmAligner(pm.models, pm.p, device=0)
    with pytest.raises(AlignerError, match="exceed the fixed scratch"):
        al.forward_tiny_noalloc(65537, 0, 1, 10, 0, 1, 10, None, 1)  # rejected before any pointer is touched
    with pytest.raises(AlignerError, match="exceed the fixed scratch"):
        al.forward_tiny_noalloc(10, 1, 1, (4 << 20) + 1, 1, 1, 10, None, 1)
    rng = np.random.default_rng(3)
    big = [(gen.random_symbols(rng, 40), gen.random_symbols(rng, 40), np.zeros((0, 3), np.int32), 0, 0, 0)]
    with pytest.raises(AlignerError, match="swept in registers"):
        _run_noalloc(al, big + _bubble_items(rng, 5))
    ok = _bubble_items(rng, 50)  # the context and the scratch are usable afterwards
    assert np.array_equal(_run_noalloc(al, ok), al.forward_batch(ok))
    al.close()
