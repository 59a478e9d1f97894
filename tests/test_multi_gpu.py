"""Two ranks on two GPUs: the loss-sum exchange fused into k7_loss over peer memory (K7PeerReduce)
against NCCL's all_reduce and against the unsharded batch.  Skipped on boxes with one GPU."""
import os
import socket

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out_dir):
    import torch.distributed as dist
    import b200det
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    dev = torch.device("cuda", rank)
    H, W, Bg, C, M = 128, 160, 6, 5, 7
    gen = torch.Generator().manual_seed(11)
    img = torch.empty(Bg, 3, H, W, device=dev)
    anchors = b200det.Anchors()(img)
    A = anchors.shape[1]
    ann = b200det.synth.make_annotations(Bg, M, C, H, W, gen, gt_min=8.0, gt_max=70.0)
    ann[4] = -1.0                                             # an image without annotations
    cls, reg = b200det.synth.make_head_outputs(Bg, A, C, gen, logit_mean=-3.0, logit_std=1.5)
    lo, hi = b200det.distributed.shard_range(Bg, rank, world)
    fused = b200det.FocalLoss(sharded=True, peer_reduce=True)
    nccl = b200det.FocalLoss(sharded=True, peer_reduce=False)
    rec = {}
    for step in range(5):                                     # both slot parities, several times
        scale = 1.0 + 0.25 * step                             # different numbers every step
        c = (cls[lo:hi] * (1.0 / scale)).to(dev).requires_grad_(True)
        r = (reg[lo:hi] * scale).to(dev).requires_grad_(True)
        lc, lr = fused(c, r, anchors, ann[lo:hi].to(dev))
        (lc + lr).sum().backward()
        c2 = c.detach().clone().requires_grad_(True)
        r2 = r.detach().clone().requires_grad_(True)
        lc2, lr2 = nccl(c2, r2, anchors, ann[lo:hi].to(dev))
        (lc2 + lr2).sum().backward()
        assert torch.equal(lc, lc2) and torch.equal(lr, lr2), (step, lc, lc2)      # two ranks: a+b == b+a
        assert torch.equal(c.grad, c2.grad) and torch.equal(r.grad, r2.grad)
        rec[f"lc{step}"], rec[f"lr{step}"] = lc.item(), lr.item()
        if step == 4:                                         # the unsharded batch on one GPU
            cf = (cls * (1.0 / scale)).to(dev).requires_grad_(True)
            rf = (reg * scale).to(dev).requires_grad_(True)
            lcf, lrf = b200det.FocalLoss()(cf, rf, anchors, ann.to(dev))
            (lcf + lrf).sum().backward()
            np.testing.assert_allclose(lc.item(), lcf.item(), rtol=2e-6)
            np.testing.assert_allclose(lr.item(), lrf.item(), rtol=2e-6)
            np.testing.assert_allclose(c.grad.cpu().numpy(), cf.grad[lo:hi].cpu().numpy(), rtol=1e-6, atol=1e-12)
    assert fused.exchange == "p2p-fused" and nccl.exchange == "nccl-allreduce"
    # the two-phase form (post now, collect on another stream), mixed with ordinary calls on the same mailboxes
    for step in range(6):
        scale = 1.0 + 0.125 * step
        c = (cls[lo:hi] * (1.0 / scale)).to(dev).requires_grad_(True)
        r = (reg[lo:hi] * scale).to(dev).requires_grad_(True)
        a = ann[lo:hi].to(dev)
        if step == 3:                                         # an ordinary call in between
            lc, lr = fused(c, r, anchors, a)
            (lc + lr).sum().backward()
        else:
            pend = fused.start(c, r, anchors, a)
            pend.backward()
            lc, lr = pend.losses()
        c2 = c.detach().clone().requires_grad_(True)
        r2 = r.detach().clone().requires_grad_(True)
        lc2, lr2 = nccl(c2, r2, anchors, a)
        (lc2 + lr2).sum().backward()
        assert torch.equal(lc, lc2) and torch.equal(lr, lr2), ("deferred", step, lc, lc2)
        assert torch.equal(c.grad, c2.grad) and torch.equal(r.grad, r2.grad)
        rec[f"dlc{step}"] = lc.item()
    # per-level head outputs through the same exchange
    sp = b200det.spec_of(anchors)
    K = sp.cell_anchors
    z = torch.logit(cls[lo:hi].clamp(1e-6, 1 - 1e-6)).to(dev)
    lv = [z[:, sp.offset[l]:sp.offset[l + 1], :].reshape(hi - lo, sp.fh[l], sp.fw[l], K * C).permute(0, 3, 1, 2).contiguous()
          for l in range(sp.num_levels)]
    fl = b200det.FocalLoss(sharded=True, peer_reduce=True, from_logits=True)
    a1, b1 = fl(lv, reg[lo:hi].to(dev), anchors, ann[lo:hi].to(dev))
    a2, b2 = b200det.FocalLoss(sharded=True, peer_reduce=False, from_logits=True)(lv, reg[lo:hi].to(dev), anchors,
                                                                                 ann[lo:hi].to(dev))
    assert torch.equal(a1, a2) and torch.equal(b1, b2)
    rec["lv"] = a1.item()
    torch.cuda.synchronize()
    np.save(os.path.join(out_dir, f"rank{rank}.npy"), np.array([rec[k] for k in sorted(rec)]))
    dist.barrier()
    for m in (fused, fl):
        m._peer.close()
    dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_fused_peer_exchange_two_ranks(tmp_path):
    import torch.multiprocessing as mp
    mp.spawn(_worker, args=(2, _free_port(), str(tmp_path)), nprocs=2, join=True)
    r0, r1 = np.load(tmp_path / "rank0.npy"), np.load(tmp_path / "rank1.npy")
    assert np.array_equal(r0, r1) and np.all(np.isfinite(r0))      # identical bits on every rank
