"""k7_nms_static: the allocation-free variant of k7_nms (include/k7det.h), through the C-ABI.

It has no counterpart in the reference; its oracle is the
existing entry point k7_nms, whose own parity with torchvision.ops.nms is pinned in test_gpu_parity.py.
NMS output is index work, so the comparison is bit-exact.  The file sorts last on purpose:
a failure here must not hide the parity tests.
"""
import numpy as np
import pytest
import torch

import b200det
from b200det import _lib as L

pytestmark = pytest.mark.gpu


def _boxes(n, seed):
    rng = np.random.default_rng(seed)
    xy = rng.uniform(0, 400, (n, 2)).astype(np.float32)
    wh = rng.uniform(1, 90, (n, 2)).astype(np.float32)
    boxes = np.concatenate([xy, xy + wh], 1)
    scores = np.round(rng.uniform(0, 1, n), 3).astype(np.float32)      # exact ties included
    return boxes, scores


def _with_workspace(lib, b, s, n, thr):
    dev = b.device
    keep = torch.full((max(n, 1),), -1, dtype=torch.int64, device=dev)
    count = torch.full((1,), -1, dtype=torch.int32, device=dev)
    ws = torch.empty(lib.k7_nms_workspace_bytes(n), dtype=torch.uint8, device=dev)
    L.check(lib.k7_nms(L.ptr(b), L.ptr(s), n, thr, L.ptr(keep), L.ptr(count), L.ptr(ws), ws.numel(),
                       L.stream_ptr(dev)), "k7_nms")
    return keep, count


def _static(lib, b, s, n, thr, keep, count):
    """Outputs are the caller's; nothing is allocated between here and the launch."""
    L.check(lib.k7_nms_static(L.ptr(b), L.ptr(s), n, thr, L.ptr(keep), L.ptr(count), L.stream_ptr(b.device)),
            "k7_nms_static")


def test_static_scratch_nms_matches_k7_nms_on_consecutive_calls():
    lib = L.lib()
    dev = torch.device("cuda:0")
    nmax = lib.k7_nms_static_max_n()
    assert nmax == 4096
    with torch.cuda.device(dev):
        # large first, small second (stale scratch from call 1 is still there), then the limit itself
        cases = [(nmax - 7, 0.5, 11), (300, 0.3, 12), (nmax, 0.5, 13), (1, 0.5, 14), (2049, 0.7, 15)]
        inputs = []
        for n, thr, seed in cases:
            bx, sc = _boxes(n, seed)
            inputs.append((torch.from_numpy(bx).to(dev), torch.from_numpy(sc).to(dev)))
        outs = [(torch.full((n,), -1, dtype=torch.int64, device=dev), torch.full((1,), -1, dtype=torch.int32, device=dev))
                for n, _, _ in cases]
        torch.cuda.synchronize()
        before = torch.cuda.memory_allocated(dev)
        for (n, thr, _), (b, s), (keep, count) in zip(cases, inputs, outs):       # back to back on one stream
            _static(lib, b, s, n, thr, keep, count)
        torch.cuda.synchronize()
        assert torch.cuda.memory_allocated(dev) == before
        for (n, thr, _), (b, s), (keep, count) in zip(cases, inputs, outs):
            want_keep, want_count = _with_workspace(lib, b, s, n, thr)
            k = int(want_count.item())
            assert int(count.item()) == k, (n, thr)
            assert 0 < k <= n
            assert torch.equal(keep[:k], want_keep[:k]), (n, thr)
        # the same input twice in a row gives the same answer
        (n, thr, _), (b, s) = cases[0], inputs[0]
        keep2 = torch.full((n,), -1, dtype=torch.int64, device=dev)
        count2 = torch.full((1,), -1, dtype=torch.int32, device=dev)
        _static(lib, b, s, n, thr, keep2, count2)
        _static(lib, b, s, n, thr, outs[0][0], outs[0][1])
        torch.cuda.synchronize()
        k = int(count2.item())
        assert k == int(outs[0][1].item()) and torch.equal(keep2[:k], outs[0][0][:k])


def test_static_scratch_nms_agrees_with_torchvision():
    from torchvision.ops import nms as tv_nms
    lib = L.lib()
    dev = torch.device("cuda:0")
    with torch.cuda.device(dev):
        for n, thr, seed in [(1500, 0.5, 21), (64, 0.0, 22)]:
            bx, sc = _boxes(n, seed)
            b, s = torch.from_numpy(bx).to(dev), torch.from_numpy(sc).to(dev)
            keep = torch.empty(n, dtype=torch.int64, device=dev)
            count = torch.empty(1, dtype=torch.int32, device=dev)
            _static(lib, b, s, n, thr, keep, count)
            want = tv_nms(torch.from_numpy(bx), torch.from_numpy(sc), thr).numpy()
            assert np.array_equal(keep[:int(count.item())].cpu().numpy(), want), (n, thr)


def test_static_scratch_nms_rejects_sizes_beyond_its_limit_and_handles_empty():
    lib = L.lib()
    dev = torch.device("cuda:0")
    nmax = lib.k7_nms_static_max_n()
    with torch.cuda.device(dev):
        n = nmax + 1
        bx, sc = _boxes(n, 31)
        b, s = torch.from_numpy(bx).to(dev), torch.from_numpy(sc).to(dev)
        keep = torch.full((n,), -1, dtype=torch.int64, device=dev)
        count = torch.full((1,), -1, dtype=torch.int32, device=dev)
        rc = lib.k7_nms_static(L.ptr(b), L.ptr(s), n, 0.5, L.ptr(keep), L.ptr(count), L.stream_ptr(dev))
        assert rc != 0
        msg = lib.k7_last_error().decode()
        assert "k7_nms_static" in msg and str(nmax) in msg
        with pytest.raises(Exception):
            L.check(rc, "k7_nms_static")
        torch.cuda.synchronize()
        assert int(count.item()) == -1 and int(keep.min().item()) == -1          # nothing was launched
        _static(lib, b, s, 0, 0.5, keep, count)                                 # n = 0: count only
        torch.cuda.synchronize()
        assert int(count.item()) == 0
