#!/usr/bin/env python
"""Run every stage of the hot path a few times on one workload — the target of the ncu captures.

    python tools/prof_kernels.py [workload] [--batch B] [--reps R] [--stages match,loss_fwd,loss_fwd_bwd,detect]

Under ncu (B200_PROFILING.md):
    ncu --set full --clock-control none --import-source on -k regex:k7_ -c 40 -o gpurun_out/X \
        python tools/prof_kernels.py c2_effdet_d0_512 --reps 2
Without ncu it prints CUDA-event times per stage (a sanity check, not a bench number)."""
import argparse
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import b200det                                    # noqa: E402
from b200det import losses as BL                  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("workload", nargs="?", default="c2_effdet_d0_512")
    ap.add_argument("--batch", type=int, default=0)
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--stages", default="match,match_exhaustive,loss_fwd,loss_fwd_bwd,detect")
    args = ap.parse_args()
    wl = b200det.synth.WORKLOADS[args.workload]
    B = args.batch or wl.batch
    dev = torch.device("cuda", 0)
    cls, reg, ann = b200det.synth.make_inputs(wl, seed=0, device=dev, batch=B)
    img = torch.empty(B, 3, wl.height, wl.width, device=dev)
    anc = b200det.Anchors(**wl.anchor_kwargs())(img)
    A = wl.anchors
    kw = dict(alpha=0.25, beta=1.0 / 9.0, target_div=(0.1, 0.1, 0.2, 0.2), yx_targets=False, mean_divisor=B)
    sp = b200det.spec_of(anc)
    thr = (0.35, 0.45)
    dreg0 = torch.empty_like(reg)

    def best():
        """A fresh assignment of the production path (every scatter call starts a new generation of the array)."""
        arr, gen, consume = BL.match_scatter(sp, ann, A, *thr, zero_fill=dreg0)
        return dict(assign=None, num_pos=None, best=(arr, gen, consume, *thr))

    def loss(c, want_grad, **extra):
        a = cur[0]
        return BL.loss_stage(anc, c, reg, ann, a["assign"], a["num_pos"], want_grad=want_grad, best=a["best"],
                             dreg_zeroed=dreg0 if want_grad else None, **extra, **kw)

    loss_mod = b200det.FocalLoss()
    cls_g, reg_g = cls.detach().requires_grad_(True), reg.detach().requires_grad_(True)

    def train():
        cls_g.grad = None
        reg_g.grad = None
        lc, lr = loss_mod(cls_g, reg_g, anc, ann)
        (lc + lr).sum().backward()

    def fwd_only():
        with torch.no_grad():
            return loss_mod(cls, reg, anc, ann)

    cur = [best()]
    stages = {
        "match": lambda: BL.match_scatter(sp, ann, A, *thr),
        "match_zero": lambda: BL.match_scatter(sp, ann, A, *thr, zero_fill=dreg0),
        "module_fwd": fwd_only,
        "module_fwd_bwd": train,
        "match_exhaustive": lambda: BL.match(anc, ann, A, *thr, zero_fill=dreg0),
        "loss_fwd": lambda: loss(cls, False),
        "loss_fwd_bwd": lambda: loss(cls, True),
        "detect": lambda: b200det.detect(cls, reg, anc, (wl.height, wl.width), sync=False),
    }
    if "logits" in args.stages:
        z = torch.logit(cls)                      # the head's pre-sigmoid tensor
        stages.update({
            "loss_fwd_logits": lambda: loss(z, False, from_logits=True),
            "loss_fwd_bwd_logits": lambda: loss(z, True, from_logits=True),
            "detect_logits": lambda: b200det.detect(z, reg, anc, (wl.height, wl.width), sync=False, from_logits=True),
            "torch_sigmoid": lambda: torch.sigmoid(z),
        })
    if "levels" in args.stages:
        # the head's raw per-level NCHW tensors holding the same logits
        K, Cn = sp.cell_anchors, wl.classes
        zz = torch.logit(cls)
        lv = [zz[:, sp.offset[l]:sp.offset[l + 1], :].reshape(B, sp.fh[l], sp.fw[l], K * Cn).permute(0, 3, 1, 2).contiguous()
              for l in range(sp.num_levels)]
        del zz

        def torch_layout():
            return torch.cat([t.permute(0, 2, 3, 1).contiguous().view(B, -1, Cn) for t in lv], dim=1)

        def loss_lv(want_grad):
            a = cur[0]
            return BL.loss_stage_levels(sp, lv, reg, ann, None, a["num_pos"], want_grad=want_grad, best=a["best"],
                                        from_logits=True, **kw)

        stages.update({
            "loss_fwd_levels": lambda: loss_lv(False),
            "loss_fwd_bwd_levels": lambda: loss_lv(True),
            "detect_levels": lambda: b200det.detect(lv, reg, anc, (wl.height, wl.width), sync=False, from_logits=True),
            "torch_layout": torch_layout,
        })
    for name in args.stages.split(","):
        fn = stages[name]
        if name.startswith("loss"):
            cur[0] = best()                       # the match stages above have moved the array on
        fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(args.reps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        print(f"{args.workload} B={B} {name}: {e0.elapsed_time(e1) / args.reps * 1e3:.1f} us/call")


if __name__ == "__main__":
    main()
