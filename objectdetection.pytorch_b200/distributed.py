"""Multi-GPU use of the hot path: one process per GPU, images sharded along dim 0.

The reference's only multi-GPU mechanism is single-process nn.DataParallel, which
scatters the batch, runs FocalLoss per replica and gathers the two 1-element
losses to GPU 0 for a .mean() (train_ksevendet.py:299-302, 358-361).  Here every
rank runs the match + loss kernels on its own shard and the four local sums
{sum cls_i, sum reg_i, sum num_pos_i, B_local} are combined with ONE all-reduce
(NCCL over NVLink on the GPU box, gloo in the CPU tests).  Gradients are
pre-scaled by the global batch inside the loss kernel, so the backward of the
path needs no collective.  The eval side needs no collective at all.
"""
from __future__ import annotations

import torch
import torch.distributed as dist


def shard_range(batch: int, rank: int, world: int):
    """Images [lo, hi) owned by `rank` (equal contiguous shards, like DataParallel's scatter)."""
    if batch % world != 0:
        raise ValueError(f"batch {batch} must divide evenly over {world} ranks")
    per = batch // world
    return rank * per, (rank + 1) * per


def combine_loss_sums(sums: torch.Tensor, group=None) -> torch.Tensor:
    """sums = [sum cls_i, sum reg_i, sum num_pos_i, B_local] -> all-reduced copy (in place)."""
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(sums, op=dist.ReduceOp.SUM, group=group)
    return sums


def mean_losses(sums: torch.Tensor):
    """(cls_mean[1], reg_mean[1]) of losses.py:168 from (all-reduced) sums."""
    m = sums[:2] / sums[3]
    return m[0:1], m[1:2]


def sharded_focal_loss(classifications, regressions, anchors, annotations, group=None, **loss_kwargs):
    """FocalLoss over a batch whose dim 0 is already this rank's shard."""
    from .losses import FocalLoss
    return FocalLoss(process_group=group, sharded=True, **loss_kwargs)(classifications, regressions, anchors, annotations)
