"""Seeded synthetic head outputs and ground truth of the shapes BASELINE.json names.

Layout contracts restated from the reference:
  classifications [B,A,C] f32, POST-SIGMOID probabilities, class innermost
      (ksevendet/architecture/head/simple_head.py:289-311, architectures/retinanet.py:163-178)
  regressions     [B,A,4] f32 raw deltas
  annotations     [B,M,5] f32 rows (x1,y1,x2,y2,label); pad rows are all -1
      (datasettool/dataloader.py:655-691)

Generation recipe: SURVEY.md §8(d).  Everything is produced with a
``torch.Generator`` so a (config, seed) pair names one input set on any device.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field
from typing import List, Optional, Sequence

import torch


@dataclass(frozen=True)
class Workload:
    name: str
    height: int
    width: int
    batch: int
    classes: int
    max_gt: int
    gt_mode: str = "uniform"          # "uniform" U{1..max_gt} rows, "fixed" = max_gt rows
    logit_mean: float = -6.5          # sparse-realistic; dense = -4.6
    logit_std: float = 1.0
    gt_min: float = 16.0
    gt_max_frac: float = 0.5
    gt_max_abs: Optional[float] = None
    # anchor recipe (Anchors ctor kwargs of the reference)
    pyramid_levels: Sequence[int] = (3, 4, 5, 6, 7)
    ratios: Sequence[float] = (0.5, 1.0, 2.0)
    sizes: Optional[Sequence[float]] = None
    n_scales: int = 3

    def anchor_kwargs(self) -> dict:
        kw = dict(pyramid_levels=list(self.pyramid_levels), ratios=list(self.ratios))
        if self.sizes is not None:
            kw["sizes"] = list(self.sizes)
        return kw

    @property
    def cells(self) -> int:
        return sum(-(-self.height // 2 ** l) * -(-self.width // 2 ** l) for l in self.pyramid_levels)

    @property
    def anchors(self) -> int:
        return self.cells * len(self.ratios) * self.n_scales


# BASELINE.json configs[0..4].  "9/cell" follows BASELINE;
# the *_coded variants follow the reference's classes as written (6/cell).
WORKLOADS = {
    "c1_retinanet_r50_512": Workload("c1_retinanet_r50_512", 512, 512, 2, 80, 20, gt_mode="fixed"),
    "c2_effdet_d0_512": Workload("c2_effdet_d0_512", 512, 512, 64, 90, 100),
    "c3_ksevendet_thermal_512x640": Workload(
        "c3_ksevendet_thermal_512x640", 512, 640, 256, 1, 100, gt_mode="fixed", logit_mean=-4.6,
        gt_min=12.0, gt_max_abs=96.0, ratios=(1.0, 2.0),
        sizes=tuple(2 ** (l + 1) * 1.5 for l in (3, 4, 5, 6, 7))),
    "c4_effdet_d3_896": Workload("c4_effdet_d3_896", 896, 896, 128, 90, 100),
    "c5_effdet_d7_1536": Workload("c5_effdet_d7_1536", 1536, 1536, 64, 90, 100),
    "c2_effdet_d0_512_coded": Workload(
        "c2_effdet_d0_512_coded", 512, 512, 64, 90, 100, ratios=(1.0, 2.0),
        sizes=tuple(2 ** (l + 1) * 1.25 for l in (3, 4, 5, 6, 7))),
}


def make_annotations(batch: int, max_gt: int, classes: int, height: int, width: int,
                     gen: torch.Generator, mode: str = "uniform", gt_min: float = 16.0,
                     gt_max: Optional[float] = None, device="cpu") -> torch.Tensor:
    """[B,max_gt,5] f32; pad rows -1 (dataloader.py:678-686)."""
    gt_max = 0.5 * min(height, width) if gt_max is None else gt_max
    dev = torch.device(device)
    if mode == "fixed":
        counts = torch.full((batch,), max_gt, dtype=torch.int64)
    else:
        counts = torch.randint(1, max_gt + 1, (batch,), generator=gen, device=gen.device).cpu()
    u = torch.rand((batch, max_gt, 5), generator=gen, device=gen.device).to(dev)
    lo, hi = math.log(gt_min), math.log(gt_max)
    w = torch.exp(lo + (hi - lo) * u[..., 0])
    aspect = torch.exp(math.log(0.5) + (math.log(2.0) - math.log(0.5)) * u[..., 1])
    h = torch.clamp(w * aspect, gt_min, gt_max)
    w = torch.minimum(w, torch.tensor(float(width - 1), device=dev))
    h = torch.minimum(h, torch.tensor(float(height - 1), device=dev))
    x1 = u[..., 2] * (width - 1 - w)
    y1 = u[..., 3] * (height - 1 - h)
    lab = torch.floor(u[..., 4] * classes).clamp(max=classes - 1)
    ann = torch.stack([x1, y1, x1 + w, y1 + h, lab], dim=-1).float()
    idx = torch.arange(max_gt)[None, :].expand(batch, -1)
    pad = (idx >= counts[:, None]).to(dev)
    ann[pad] = -1.0
    return ann.contiguous()


def make_head_outputs(batch: int, anchors: int, classes: int, gen: torch.Generator,
                      logit_mean: float = -6.5, logit_std: float = 1.0, device="cpu"):
    """(classifications [B,A,C] probabilities, regressions [B,A,4])."""
    dev = torch.device(device)
    cls = torch.empty((batch, anchors, classes), dtype=torch.float32, device=dev)
    # generated per image to bound the peak temporary (D7: 159 MB / image)
    for b in range(batch):
        z = torch.randn((anchors, classes), generator=gen, device=gen.device, dtype=torch.float32)
        cls[b] = torch.sigmoid(z.mul_(logit_std).add_(logit_mean)).to(dev)
    reg = (torch.randn((batch, anchors, 4), generator=gen, device=gen.device,
                       dtype=torch.float32) * 0.5).to(dev)
    return cls, reg


def make_inputs(wl: Workload, seed: int = 0, device="cpu", batch: Optional[int] = None):
    """(classifications, regressions, annotations) for a workload."""
    dev = torch.device(device)
    gen = torch.Generator(device=dev if dev.type == "cuda" else "cpu")
    gen.manual_seed(20260000 + seed)
    B = wl.batch if batch is None else batch
    gt_max = wl.gt_max_abs if wl.gt_max_abs is not None else wl.gt_max_frac * min(wl.height, wl.width)
    ann = make_annotations(B, wl.max_gt, wl.classes, wl.height, wl.width, gen, wl.gt_mode,
                           wl.gt_min, gt_max, device=dev)
    cls, reg = make_head_outputs(B, wl.anchors, wl.classes, gen, wl.logit_mean, wl.logit_std, device=dev)
    return cls, reg, ann
