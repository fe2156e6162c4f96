"""GPU: the allocation-free NMS entry point (b2_nms_single_noalloc) against the regular one.

Some users of this library cannot allocate GPU memory when they call it — a perception loop that sized its memory
once at start-up and forbids cudaMalloc afterwards, or a host application that owns the GPU heap — so this entry
point keeps its scratch inside the library.  Called through the C ABI, twice in a row on different inputs (the second
call sees the first call's leftovers in the shared scratch), and compared bit for bit with b2_nms_batched.
Collected last on purpose: if it fails under `-x`, every other test has already run."""
import numpy as np
import pytest
import torch

from fcaf3d_b200 import _lib, ops
from fcaf3d_b200.scenes import random_boxes

pytestmark = pytest.mark.gpu


def _sorted(n, seed, rotated, dev):
    b, s = random_boxes(n, seed, rotated)
    return torch.from_numpy(b[np.argsort(-s, kind="stable")]).to(dev)


def _noalloc(boxes, thr, rotated, keep, count):
    _lib.call("b2_nms_single_noalloc", boxes.data_ptr(), boxes.shape[0], float(thr), int(rotated), keep.data_ptr(),
              count.data_ptr())


def test_two_calls_in_a_row_match_the_regular_entry_point(dev):
    assert _lib.lib().b2_nms_single_noalloc_max_boxes() == 2048
    cases = [(2048, 1, True), (700, 2, False), (65, 3, True), (1, 4, False), (0, 5, True)]
    inputs = [(_sorted(max(n, 1), seed, rot, dev)[:n].contiguous(), rot) for n, seed, rot in cases]
    # outputs come from the caller and are allocated before the calls under test
    outs = [(torch.full((max(b.shape[0], 1),), 7, dtype=torch.uint8, device=dev),
             torch.full((1,), -1, dtype=torch.int32, device=dev)) for b, _ in inputs]
    for (boxes, rot), (keep, count) in zip(inputs, outs):
        _noalloc(boxes, 0.5, rot, keep, count)          # back to back, same stream, shared scratch
    torch.cuda.synchronize()
    for (boxes, rot), (keep, count) in zip(inputs, outs):
        n = boxes.shape[0]
        want, _, _, wcount = ops.nms_batched(boxes, None, [0, n], 0.5, rot, sort=False, greedy=False)
        assert torch.equal(keep[:n], want), (n, rot)
        assert count.item() == wcount[0].item() == int(want.sum())


def test_size_limit_is_rejected_with_a_clear_error(dev):
    boxes = _sorted(2049, 9, False, dev)
    keep = torch.zeros(2049, dtype=torch.uint8, device=dev)
    count = torch.zeros(1, dtype=torch.int32, device=dev)
    with pytest.raises(_lib.B2Error, match="at most 2048 boxes"):
        _noalloc(boxes, 0.5, False, keep, count)
