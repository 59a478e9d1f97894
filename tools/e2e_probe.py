#!/usr/bin/env python
"""Where the end-to-end step of bench.py spends its time (C2): copy only, + loss, + detect, + host reads."""
import os, sys, time
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import b200det

wl = b200det.synth.WORKLOADS["c2_effdet_d0_512"]
dev = torch.device("cuda", 0)
B = wl.batch
cls, reg, ann = b200det.synth.make_inputs(wl, seed=0, device=dev, batch=B)
img = torch.empty(B, 3, wl.height, wl.width, device=dev)
anc = b200det.Anchors(**wl.anchor_kwargs())
loss_mod = b200det.FocalLoss()
h_cls, h_reg, h_ann = (t.cpu().pin_memory() for t in (cls, reg, ann))
d_cls, d_reg, d_ann = torch.empty_like(cls), torch.empty_like(reg), torch.empty_like(ann)
hw = (wl.height, wl.width)


def copies():
    d_cls.copy_(h_cls, non_blocking=True)
    d_reg.copy_(h_reg, non_blocking=True)
    d_ann.copy_(h_ann, non_blocking=True)


def loss():
    c = d_cls.detach().requires_grad_(True)
    r = d_reg.detach().requires_grad_(True)
    lc, lr = loss_mod(c, r, anc(img), d_ann)
    (lc + lr).sum().backward()
    return lc, lr


def v_copy():
    copies()


def v_copy_loss():
    copies()
    lc, lr = loss()
    torch.cat([lc.detach(), lr.detach()]).cpu()


def v_copy_loss_detect():
    copies()
    lc, lr = loss()
    torch.cat([lc.detach(), lr.detach()]).cpu()
    b200det.detect(d_cls, d_reg, anc(img), hw, sync=True)


def v_full():
    copies()
    lc, lr = loss()
    torch.cat([lc.detach(), lr.detach()]).cpu()
    det = b200det.detect(d_cls, d_reg, anc(img), hw, sync=True)
    kmax = max(int(det.count.max()), 1)
    return det.scores[:, :kmax].cpu(), det.labels[:, :kmax].cpu(), det.boxes[:, :kmax].cpu()


def v_compute_only():
    lc, lr = loss()
    torch.cat([lc.detach(), lr.detach()]).cpu()
    b200det.detect(d_cls, d_reg, anc(img), hw, sync=True)


for name, fn in [("copy", v_copy), ("copy+loss", v_copy_loss), ("copy+loss+detect", v_copy_loss_detect), ("full", v_full),
                 ("compute only (no copy)", v_compute_only)]:
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(8):
        fn()
    torch.cuda.synchronize()
    print(f"{name}: {(time.perf_counter() - t0) / 8 * 1e3:.3f} ms")
