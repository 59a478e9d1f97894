"""world_size-2 test of the multi-GPU host logic on CPU (gloo).

The sharded path is: every rank runs match + loss on ITS images and produces the four sums
{sum cls_i, sum reg_i, sum num_pos_i, B_local}; ONE all-reduce combines them; the batch mean of
losses.py:168 is sums[0:2] / sums[3].  The kernels need a GPU, so here the per-shard sums come
from the oracle (the checker) and the test covers what runs on the host: shard_range,
combine_loss_sums over a real process group, mean_losses, and the claim the design rests on —
sharded sums + all-reduce == the reference's mean over the concatenated batch, with gradients
pre-scaled by the GLOBAL batch so backward needs no collective.
"""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import b200det
from b200det import distributed as D
from oracle import restate as R

H, W, B, C, M = 96, 128, 4, 5, 6


def _inputs():
    gen = torch.Generator().manual_seed(11)
    anc = torch.from_numpy(R.retina_anchors(H, W))[None]
    A = anc.shape[1]
    ann = b200det.synth.make_annotations(B, M, C, H, W, gen, gt_min=8.0, gt_max=60.0)
    ann[1] = -1.0                                  # an image without any valid row (losses.py:52-60)
    cls, reg = b200det.synth.make_head_outputs(B, A, C, gen, logit_mean=-2.0, logit_std=1.5)
    return anc, cls, reg, ann


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        torch.set_num_threads(1)
        anc, cls, reg, ann = _inputs()
        lo, hi = D.shard_range(B, rank, world)
        c = cls[lo:hi].clone().requires_grad_(True)
        r = reg[lo:hi].clone().requires_grad_(True)
        out = R.focal_loss(c, r, anc, ann[lo:hi])
        sums = torch.stack([out.cls_per_image.sum(), out.reg_per_image.sum(),
                            torch.tensor(float(sum(out.num_pos))), torch.tensor(float(hi - lo))]).detach()
        D.combine_loss_sums(sums)                                  # the path's only collective
        lc, lr = D.mean_losses(sums)
        # local gradient of the GLOBAL mean: d/dx (sum_local / B_global); no collective needed
        ((out.cls_per_image.sum() + out.reg_per_image.sum()) / sums[3]).backward()
        q.put((rank, lc.numpy(), lr.numpy(), sums.numpy(), c.grad.numpy(), r.grad.numpy()))
    finally:
        dist.destroy_process_group()


def test_sharded_sums_allreduce_equals_full_batch():
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = sorted((q.get(timeout=120) for _ in range(world)), key=lambda t: t[0])
    for p in procs:
        p.join(60)
        assert p.exitcode == 0

    anc, cls, reg, ann = _inputs()
    c = cls.clone().requires_grad_(True)
    r = reg.clone().requires_grad_(True)
    want = R.focal_loss(c, r, anc, ann)
    (want.cls_mean + want.reg_mean).sum().backward()
    for rank, lc, lr, sums, dc, dr in got:
        np.testing.assert_allclose(lc, want.cls_mean.detach().numpy(), rtol=1e-6)
        np.testing.assert_allclose(lr, want.reg_mean.detach().numpy(), rtol=1e-6)
        assert sums[3] == B and sums[2] == sum(want.num_pos)
        lo, hi = D.shard_range(B, rank, world)
        np.testing.assert_allclose(dc, c.grad[lo:hi].numpy(), rtol=1e-5, atol=1e-12)
        np.testing.assert_allclose(dr, r.grad[lo:hi].numpy(), rtol=1e-5, atol=1e-12)
    # both ranks hold the same all-reduced payload
    assert np.array_equal(got[0][3], got[1][3])


def test_shard_range():
    assert [D.shard_range(64, r, 8) for r in (0, 7)] == [(0, 8), (56, 64)]
    with pytest.raises(ValueError):
        D.shard_range(10, 0, 4)


def test_combine_is_identity_without_process_group():
    s = torch.tensor([1.0, 2.0, 3.0, 4.0])
    assert torch.equal(D.combine_loss_sums(s.clone()), s)
    lc, lr = D.mean_losses(s)
    assert lc.shape == (1,) and lr.shape == (1,) and float(lc) == 0.25 and float(lr) == 0.5
