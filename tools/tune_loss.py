#!/usr/bin/env python
"""Sweep the streaming geometry of k7_loss (K7_SLICE x K7_WARPS x K7_STAGES) on one workload.
Usage (GPU box): python tools/tune_loss.py [workload]"""
import itertools
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import b200det                                    # noqa: E402
from b200det import losses as BL                  # noqa: E402


def timed(fn, steps=20, warmup=3):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "c2_effdet_d0_512"
    wl = b200det.synth.WORKLOADS[name]
    B = int(os.environ.get("BATCH", wl.batch))
    dev = torch.device("cuda", 0)
    cls, reg, ann = b200det.synth.make_inputs(wl, seed=0, device=dev, batch=B)
    img = torch.empty(B, 3, wl.height, wl.width, device=dev)
    anc = b200det.Anchors(**wl.anchor_kwargs())(img)
    A, C, M = wl.anchors, wl.classes, wl.max_gt
    arr, gen, consume = BL.match_scatter(b200det.spec_of(anc), ann, A, 0.35, 0.45)
    num_pos = None
    best = (arr, gen, consume, 0.35, 0.45)
    kw = dict(alpha=0.25, beta=1.0 / 9.0, target_div=(0.1, 0.1, 0.2, 0.2), yx_targets=False, mean_divisor=B)
    bf = B * (A * (4 * C + 16) + 20 * M)
    bb = B * (A * (8 * C + 32) + 20 * M)
    print(f"{name} B={B} A={A} C={C}: fwd {bf/1e9:.3f} GB, fwd+bwd {bb/1e9:.3f} GB")
    combos = [(512, w, s) for w, s in itertools.product((12, 14, 16), (4, 6, 8))]
    combos += [(1024, w, s) for w, s in itertools.product((6, 8, 10, 12), (3, 4, 6))]
    for slice_, warps, stages in combos:
        os.environ["K7_SLICE"], os.environ["K7_WARPS"], os.environ["K7_STAGES"] = str(slice_), str(warps), str(stages)
        try:
            f = timed(lambda: BL.loss_stage(anc, cls, reg, ann, None, num_pos, want_grad=False, best=best, **kw))
            g = timed(lambda: BL.loss_stage(anc, cls, reg, ann, None, num_pos, want_grad=True, best=best, **kw))
        except Exception as e:                     # geometry that does not fit
            print(f"slice={slice_} warps={warps:2d} stages={stages}: {e}")
            continue
        print(f"slice={slice_:4d} warps={warps:2d} stages={stages}: fwd {f*1e3:7.1f} us {bf/f/1e6:7.0f} GB/s | "
              f"fwd+bwd {g*1e3:7.1f} us {bb/g/1e6:7.0f} GB/s")


if __name__ == "__main__":
    main()
