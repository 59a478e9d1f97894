"""CPU oracle for the detection-head hot path of Kai7/ObjectDetection.Pytorch.

TEST INFRASTRUCTURE ONLY.  Nothing under ``oracle/`` is part of the product:
only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline /
``--impl reference`` legs may import it, and only as the checker or the timed
CPU baseline.  The product (``b200det``) never imports this file and has no
CPU path.

What this is: a restatement, in float32 torch-CPU / float64 numpy arithmetic,
of the reference algorithm, evaluated in the SAME ORDER OF IEEE OPERATIONS as
the reference so that integer results (assignment, argmax, survivor order, NMS
keep) are bit-identical and float results agree to rounding.  Unlike the
reference it returns the intermediates the contract is stated on (IoU max /
argmax / state / num_pos / survivor order / keep), which the reference computes
but never returns.

Reference lines restated (paths relative to /root/reference):
  anchors           architectures/anchors.py:38-119 (Anchors.forward,
                    generate_anchors, shift); ksevendet/architecture/anchors.py:14-21
  yx anchors        architectures/utils_efficientdet.py:74-137
  IoU               architectures/losses.py:6-23; architectures/loss_efficientdet.py:10-24
  assignment+loss   architectures/losses.py:30-168; architectures/loss_efficientdet.py:31-160
  decode / clip     architectures/utils.py:26-71; architectures/utils_efficientdet.py:8-50
  eval tail         architectures/retinanet.py:297-318; utils/utils.py:76-114
  evaluator tails   datasettool/coco_eval.py:25-57; ksevendata_eval.py:24-61; csv_eval.py:93-121
  NMS               torchvision.ops.nms (third party, NOT vendored by the reference;
                    pinned here to torchvision 0.26.0 as installed).  `greedy_nms`
                    restates its published CPU algorithm (stable descending sort,
                    greedy, strict '>' on inter/(area_i+area_j-inter), no '+1').

Parity pinning: the reference ships no tests/golden vectors for this path
(SURVEY.md §4), so the oracle is pinned against outputs of the reference
itself: ``tests/golden/make_golden.py`` imports the unmodified reference
modules from /root/reference (+ torchvision.ops.nms), runs them on seeded
inputs and commits the results under ``tests/golden/``;
``tests/test_oracle_golden.py`` checks this file against those fixtures.
Per-category NMS (``batched_nms``) is pinned the same way against torchvision
0.26.0's ``batched_nms`` (tests/golden/batched_nms.npz).
PARITY UNPINNED for ONE function: ``soft_nms`` - neither the reference nor
torchvision implements Soft-NMS, so it restates the published algorithm
(Bodla et al., ICCV 2017) and is checked only on cases with known answers.
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import List, Optional, Sequence, Tuple

import numpy as np
import torch

# --------------------------------------------------------------------------
# Variants.  "retina" = architectures/{anchors,losses,utils}.py (what every
# shipped model uses); "effdet" = architectures/{loss,utils}_efficientdet.py.
# --------------------------------------------------------------------------


@dataclass(frozen=True)
class LossVariant:
    neg_thr: float
    pos_thr: float
    yx_anchors: bool          # anchors stored (y1,x1,y2,x2)
    scale_targets: bool       # divide targets by [.1,.1,.2,.2]
    yx_targets: bool          # target column order (dy,dx,dh,dw)


RETINA = LossVariant(0.35, 0.45, False, True, False)       # losses.py:79-80,148
EFFDET = LossVariant(0.4, 0.5, True, False, True)          # loss_efficientdet.py:75,77,127


# --------------------------------------------------------------------------
# Anchors
# --------------------------------------------------------------------------

def cell_anchor_table(base_size, ratios, scales) -> np.ndarray:
    """[len(ratios)*len(scales), 4] float64 boxes centred on the origin,
    ratio-major / scale-minor (anchors.py:62-97)."""
    ratios = np.asarray(ratios, dtype=np.float64)
    scales = np.asarray(scales, dtype=np.float64)
    n = len(ratios) * len(scales)
    side = base_size * np.tile(scales, len(ratios))          # w = h = size*scale
    area = side * side
    rr = np.repeat(ratios, len(scales))
    w = np.sqrt(area / rr)
    h = w * rr
    out = np.zeros((n, 4))
    out[:, 0] = 0.0 - w * 0.5
    out[:, 1] = 0.0 - h * 0.5
    out[:, 2] = w - w * 0.5
    out[:, 3] = h - h * 0.5
    return out


def retina_anchors(height: int, width: int, pyramid_levels=None, strides=None,
                   sizes=None, ratios=None, scales=None) -> np.ndarray:
    """[A,4] float32, (x1,y1,x2,y2); level-major, y, x, anchor-in-cell
    (anchors.py:38-60, 99-119).  All arithmetic in float64, one final cast."""
    levels = [3, 4, 5, 6, 7] if pyramid_levels is None else list(pyramid_levels)
    strides = [2 ** l for l in levels] if strides is None else list(strides)
    sizes = [2 ** (l + 2) for l in levels] if sizes is None else list(sizes)
    ratios = np.array([0.5, 1, 2]) if ratios is None else np.asarray(ratios)
    scales = (np.array([2 ** 0, 2 ** (1.0 / 3.0), 2 ** (2.0 / 3.0)])
              if scales is None else np.asarray(scales))
    parts = []
    for stride, size in zip(strides, sizes):
        fh = (height + stride - 1) // stride                  # anchors.py:43
        fw = (width + stride - 1) // stride
        cell = cell_anchor_table(size, ratios, scales)        # [K,4]
        sx = (np.arange(0, fw) + 0.5) * stride
        sy = (np.arange(0, fh) + 0.5) * stride
        gx, gy = np.meshgrid(sx, sy)
        sh = np.stack([gx.ravel(), gy.ravel(), gx.ravel(), gy.ravel()], axis=1)
        parts.append((sh[:, None, :] + cell[None, :, :]).reshape(-1, 4))
    return np.concatenate(parts, axis=0).astype(np.float32)


def effdet_anchors(height: int, width: int, anchor_scale=4.0, pyramid_levels=None,
                   strides=None, scales=None, ratios=None) -> np.ndarray:
    """[A,4] float32, (y1,x1,y2,x2); level-major, cell, (scale-major, ratio-minor)
    (utils_efficientdet.py:74-137)."""
    levels = [3, 4, 5, 6, 7] if pyramid_levels is None else list(pyramid_levels)
    strides = [2 ** l for l in levels] if strides is None else list(strides)
    scales = ([2 ** 0, 2 ** (1.0 / 3.0), 2 ** (2.0 / 3.0)] if scales is None else list(scales))
    ratios = ([(1.0, 1.0), (1.4, 0.7), (0.7, 1.4)] if ratios is None else list(ratios))
    parts = []
    for stride in strides:
        if width % stride != 0:
            raise ValueError('input size must be divided by the stride.')
        xs = np.arange(stride / 2, width, stride)
        ys = np.arange(stride / 2, height, stride)
        gx, gy = np.meshgrid(xs, ys)
        gx = gx.reshape(-1)
        gy = gy.reshape(-1)
        per_k = []
        for sc in scales:
            for ra in ratios:
                base = anchor_scale * stride * sc
                hx = base * ra[0] / 2.0
                hy = base * ra[1] / 2.0
                per_k.append(np.stack([gy - hy, gx - hx, gy + hy, gx + hx], axis=1))
        parts.append(np.stack(per_k, axis=1).reshape(-1, 4))
    return np.concatenate(parts, axis=0).astype(np.float32)


# --------------------------------------------------------------------------
# IoU + assignment
# --------------------------------------------------------------------------

def pair_iou(anchors: torch.Tensor, gt: torch.Tensor, yx_anchors: bool = False) -> torch.Tensor:
    """[A,M] IoU, op order of losses.py:6-23 (loss_efficientdet.py:10-24 for yx)."""
    if yx_anchors:
        ay1, ax1, ay2, ax2 = (anchors[:, i] for i in range(4))
    else:
        ax1, ay1, ax2, ay2 = (anchors[:, i] for i in range(4))
    area_g = (gt[:, 2] - gt[:, 0]) * (gt[:, 3] - gt[:, 1])
    iw = torch.min(ax2[:, None], gt[:, 2]) - torch.max(ax1[:, None], gt[:, 0])
    ih = torch.min(ay2[:, None], gt[:, 3]) - torch.max(ay1[:, None], gt[:, 1])
    iw = iw.clamp(min=0)
    ih = ih.clamp(min=0)
    if yx_anchors:
        area_a = (anchors[:, 2] - anchors[:, 0]) * (anchors[:, 3] - anchors[:, 1])
    else:
        area_a = (ax2 - ax1) * (ay2 - ay1)
    union = area_a[:, None] + area_g - iw * ih
    union = union.clamp(min=1e-8)
    return (iw * ih) / union


NEG, IGN, POS = 0, 1, 2


@dataclass
class Match:
    valid_rows: torch.Tensor      # [Mv] int64 original row index of each kept GT
    iou_max: torch.Tensor         # [A] f32
    argmax: torch.Tensor          # [A] int64 index into the FILTERED GT list
    gt_row: torch.Tensor          # [A] int64 original row index (valid_rows[argmax])
    state: torch.Tensor           # [A] int8 NEG/IGN/POS
    num_pos: int


def match_image(anchors: torch.Tensor, annot: torch.Tensor, v: LossVariant) -> Optional[Match]:
    """Assignment of one image (losses.py:49-82).  None when the image has no
    valid annotation row (losses.py:52-60: both losses are 0 and nothing else
    is computed)."""
    rows = torch.nonzero(annot[:, 4] != -1).flatten()
    if rows.numel() == 0:
        return None
    gt = annot[rows]
    iou = pair_iou(anchors, gt[:, :4], v.yx_anchors)
    iou_max, argmax = torch.max(iou, dim=1)
    state = torch.full((anchors.shape[0],), IGN, dtype=torch.int8)
    state[torch.lt(iou_max, v.neg_thr)] = NEG
    pos = torch.ge(iou_max, v.pos_thr)
    state[pos] = POS
    return Match(rows, iou_max, argmax, rows[argmax], state, int(pos.sum()))


# --------------------------------------------------------------------------
# Loss
# --------------------------------------------------------------------------

@dataclass
class LossOut:
    cls_mean: torch.Tensor                 # [1]
    reg_mean: torch.Tensor                 # [1]
    cls_per_image: torch.Tensor            # [B]
    reg_per_image: torch.Tensor            # [B]
    num_pos: List[int]
    matches: List[Optional[Match]]


def focal_loss(classifications: torch.Tensor, regressions: torch.Tensor,
               anchors: torch.Tensor, annotations: torch.Tensor,
               v: LossVariant = RETINA, alpha: float = 0.25) -> LossOut:
    """losses.py:30-168 with intermediates.  Differentiable through torch autograd
    (the graph is the same sequence of ATen ops as the reference)."""
    a = anchors[0] if anchors.dim() == 3 else anchors
    if v.yx_anchors:
        aw = a[:, 3] - a[:, 1]
        ah = a[:, 2] - a[:, 0]
        acx = a[:, 1] + 0.5 * aw
        acy = a[:, 0] + 0.5 * ah
    else:
        aw = a[:, 2] - a[:, 0]
        ah = a[:, 3] - a[:, 1]
        acx = a[:, 0] + 0.5 * aw
        acy = a[:, 1] + 0.5 * ah
    B = classifications.shape[0]
    cls_l, reg_l, npos, matches = [], [], [], []
    for j in range(B):
        m = match_image(a, annotations[j], v)
        matches.append(m)
        if m is None:
            cls_l.append(torch.zeros((), dtype=torch.float32))
            reg_l.append(torch.zeros((), dtype=torch.float32))
            npos.append(0)
            continue
        p = torch.clamp(classifications[j], 1e-4, 1.0 - 1e-4)
        gt = annotations[j][m.valid_rows]
        pos = m.state == POS
        t = torch.full(p.shape, -1.0)
        t[m.state == NEG, :] = 0
        t[pos, :] = 0
        picked = gt[m.argmax]
        t[pos, picked[pos, 4].long()] = 1
        af = torch.full(p.shape, alpha)
        af = torch.where(t == 1.0, af, 1.0 - af)
        fw = torch.where(t == 1.0, 1.0 - p, p)
        fw = af * torch.pow(fw, 2.0)
        bce = -(t * torch.log(p) + (1.0 - t) * torch.log(1.0 - p))
        el = fw * bce
        el = torch.where(t != -1.0, el, torch.zeros_like(el))
        n = pos.sum()
        cls_l.append(el.sum() / torch.clamp(n.float(), min=1.0))
        npos.append(int(n))
        if int(n) > 0:
            g = picked[pos]
            gw = g[:, 2] - g[:, 0]
            gh = g[:, 3] - g[:, 1]
            gcx = g[:, 0] + 0.5 * gw
            gcy = g[:, 1] + 0.5 * gh
            gw = gw.clamp(min=1)
            gh = gh.clamp(min=1)
            dx = (gcx - acx[pos]) / aw[pos]
            dy = (gcy - acy[pos]) / ah[pos]
            dw = torch.log(gw / aw[pos])
            dh = torch.log(gh / ah[pos])
            if v.yx_targets:
                tg = torch.stack((dy, dx, dh, dw)).t()
            else:
                tg = torch.stack((dx, dy, dw, dh)).t()
            if v.scale_targets:
                tg = tg / torch.tensor([[0.1, 0.1, 0.2, 0.2]])
            d = torch.abs(tg - regressions[j][pos])
            sl = torch.where(d <= 1.0 / 9.0, 0.5 * 9.0 * torch.pow(d, 2), d - 0.5 / 9.0)
            reg_l.append(sl.mean())
        else:
            reg_l.append(torch.zeros((), dtype=torch.float32))
    cls_pi = torch.stack(cls_l)
    reg_pi = torch.stack(reg_l)
    return LossOut(cls_pi.mean(dim=0, keepdim=True), reg_pi.mean(dim=0, keepdim=True),
                   cls_pi, reg_pi, npos, matches)


# --------------------------------------------------------------------------
# Decode / clip
# --------------------------------------------------------------------------

def decode_boxes(anchors: torch.Tensor, deltas: torch.Tensor,
                 mean=(0.0, 0.0, 0.0, 0.0), std=(0.1, 0.1, 0.2, 0.2)) -> torch.Tensor:
    """utils.py:26-53: anchors [1|B,A,4] (x1,y1,x2,y2), deltas [B,A,4] -> [B,A,4]."""
    mean = torch.tensor(np.array(mean).astype(np.float32))
    std = torch.tensor(np.array(std).astype(np.float32))
    w = anchors[:, :, 2] - anchors[:, :, 0]
    h = anchors[:, :, 3] - anchors[:, :, 1]
    cx = anchors[:, :, 0] + 0.5 * w
    cy = anchors[:, :, 1] + 0.5 * h
    dx = deltas[:, :, 0] * std[0] + mean[0]
    dy = deltas[:, :, 1] * std[1] + mean[1]
    dw = deltas[:, :, 2] * std[2] + mean[2]
    dh = deltas[:, :, 3] * std[3] + mean[3]
    pcx = cx + dx * w
    pcy = cy + dy * h
    pw = torch.exp(dw) * w
    ph = torch.exp(dh) * h
    return torch.stack([pcx - 0.5 * pw, pcy - 0.5 * ph, pcx + 0.5 * pw, pcy + 0.5 * ph], dim=2)


def decode_boxes_yx(anchors: torch.Tensor, deltas: torch.Tensor) -> torch.Tensor:
    """utils_efficientdet.py:8-35: anchors (y1,x1,y2,x2), deltas (dy,dx,dh,dw)
    -> boxes (x1,y1,x2,y2)."""
    yc = (anchors[..., 0] + anchors[..., 2]) / 2
    xc = (anchors[..., 1] + anchors[..., 3]) / 2
    ha = anchors[..., 2] - anchors[..., 0]
    wa = anchors[..., 3] - anchors[..., 1]
    w = deltas[..., 3].exp() * wa
    h = deltas[..., 2].exp() * ha
    y = deltas[..., 0] * ha + yc
    x = deltas[..., 1] * wa + xc
    return torch.stack([x - w / 2., y - h / 2., x + w / 2., y + h / 2.], dim=2)


def clip_boxes_(boxes: torch.Tensor, height: int, width: int) -> torch.Tensor:
    """utils.py:61-71: one-sided, in place."""
    boxes[:, :, 0] = torch.clamp(boxes[:, :, 0], min=0)
    boxes[:, :, 1] = torch.clamp(boxes[:, :, 1], min=0)
    boxes[:, :, 2] = torch.clamp(boxes[:, :, 2], max=width - 1)
    boxes[:, :, 3] = torch.clamp(boxes[:, :, 3], max=height - 1)
    return boxes


# --------------------------------------------------------------------------
# NMS (restatement of torchvision's CPU kernel) + eval tail
# --------------------------------------------------------------------------

def greedy_nms(boxes: np.ndarray, scores: np.ndarray, thr: float) -> np.ndarray:
    """Indices kept, in descending-score order.  float32 arithmetic throughout."""
    boxes = np.asarray(boxes, dtype=np.float32)
    scores = np.asarray(scores, dtype=np.float32)
    n = boxes.shape[0]
    if n == 0:
        return np.zeros((0,), dtype=np.int64)
    order = np.argsort(-scores, kind='stable')               # ties: lower index first
    x1, y1, x2, y2 = (boxes[:, i] for i in range(4))
    area = (x2 - x1) * (y2 - y1)
    dead = np.zeros(n, dtype=bool)
    thr32 = np.float32(thr)
    keep = []
    for pos in range(n):
        i = order[pos]
        if dead[i]:
            continue
        keep.append(i)
        rest = order[pos + 1:]
        if rest.size == 0:
            break
        w = np.maximum(np.float32(0), np.minimum(x2[i], x2[rest]) - np.maximum(x1[i], x1[rest]))
        h = np.maximum(np.float32(0), np.minimum(y2[i], y2[rest]) - np.maximum(y1[i], y1[rest]))
        inter = w * h
        with np.errstate(divide='ignore', invalid='ignore'):
            ovr = inter / (area[i] + area[rest] - inter)
        dead[rest[ovr > thr32]] = True
    return np.asarray(keep, dtype=np.int64)


def batched_nms(boxes: np.ndarray, scores: np.ndarray, idxs: np.ndarray, thr: float) -> np.ndarray:
    """Per-category NMS - the reference is class-agnostic, so the behaviour restated is
    torchvision 0.26.0 ``torchvision/ops/boxes.py::_batched_nms_vanilla``: nms() on the boxes of every
    category by itself, the union of the kept indices ordered by descending score.  Ties: torchvision's
    final ``sort(descending=True)`` is not a stable sort; here equal scores keep their index order
    (what the device path does)."""
    boxes = np.asarray(boxes, dtype=np.float32).reshape(-1, 4)
    scores = np.asarray(scores, dtype=np.float32)
    idxs = np.asarray(idxs)
    keep_mask = np.zeros(scores.shape[0], dtype=bool)
    for c in np.unique(idxs):
        cur = np.nonzero(idxs == c)[0]
        keep_mask[cur[greedy_nms(boxes[cur], scores[cur], thr)]] = True
    kept = np.nonzero(keep_mask)[0]
    return kept[np.argsort(-scores[kept], kind='stable')].astype(np.int64)


def batched_nms_coordinate_trick(boxes: np.ndarray, scores: np.ndarray, idxs: np.ndarray, thr: float) -> np.ndarray:
    """torchvision 0.26.0 ``_batched_nms_coordinate_trick``: boxes shifted by idxs * (max coordinate + 1)
    in float32, then plain NMS."""
    boxes = np.asarray(boxes, dtype=np.float32).reshape(-1, 4)
    if boxes.size == 0:
        return np.zeros((0,), dtype=np.int64)
    off = np.asarray(idxs).astype(np.float32) * (boxes.max() + np.float32(1))
    return greedy_nms(boxes + off[:, None], scores, thr)


def soft_nms(boxes: np.ndarray, scores: np.ndarray, method: str = "gaussian", iou_thr: float = 0.3,
             sigma: float = 0.5, score_thr: float = 0.001):
    """Soft-NMS as published (Bodla, Singh, Chellappa, Davis: "Soft-NMS - Improving Object Detection With One
    Line of Code", ICCV 2017, algorithm 1 / eq. 2-3).  PARITY UNPINNED: neither the reference nor torchvision
    implements it, so there is no third-party output to pin this restatement against; it follows the paper with
    torchvision.ops.nms's overlap (inter / (area_i + area_j - inter), float32).  Returns (indices in pick order,
    the score each box had when picked).  Ties: lowest index; disjoint / degenerate pairs do not decay."""
    boxes = np.asarray(boxes, dtype=np.float32).reshape(-1, 4)
    s = np.asarray(scores, dtype=np.float32).copy()
    n = boxes.shape[0]
    s[np.isnan(s)] = -np.inf
    x1, y1, x2, y2 = (boxes[:, i] for i in range(4))
    area = (x2 - x1) * (y2 - y1)
    keep, out = [], []
    thr32, sig32, sthr32 = np.float32(iou_thr), np.float32(sigma), np.float32(score_thr)
    for _ in range(n):
        i = int(np.argmax(s))                      # first maximum = lowest index on ties
        if not (s[i] >= sthr32):
            break
        keep.append(i)
        out.append(s[i])
        s[i] = -np.inf
        live = np.nonzero(s > -np.inf)[0]
        if live.size == 0:
            break
        w = np.maximum(np.float32(0), np.minimum(x2[i], x2[live]) - np.maximum(x1[i], x1[live]))
        h = np.maximum(np.float32(0), np.minimum(y2[i], y2[live]) - np.maximum(y1[i], y1[live]))
        inter = w * h
        with np.errstate(divide='ignore', invalid='ignore'):
            iou = inter / (area[i] + area[live] - inter)
        wgt = np.ones_like(iou)
        pos = iou > 0
        if method == "linear":
            m = pos & (iou > thr32)
            wgt[m] = np.float32(1) - iou[m]
        else:
            wgt[pos] = np.exp(-(iou[pos] * iou[pos]) / sig32).astype(np.float32)
        s[live] = s[live] * wgt
    return np.asarray(keep, dtype=np.int64), np.asarray(out, dtype=np.float32)


@dataclass
class Detections:
    survivors: torch.Tensor      # [S] int64 anchor ids passing the score threshold (ascending)
    keep_rank: torch.Tensor      # [K] int64 index into survivors, as torchvision.nms returns
    keep_anchor: torch.Tensor    # [K] int64 anchor ids (survivors[keep_rank])
    scores: torch.Tensor         # [K] f32
    labels: torch.Tensor         # [K] int64
    boxes: torch.Tensor          # [K,4] f32


def detect_image(cls: torch.Tensor, boxes: torch.Tensor, score_thr: float = 0.05,
                 iou_thr: float = 0.5, use_torchvision: bool = True) -> Detections:
    """retinanet.py:300-318 for one image: cls [A,C] probabilities, boxes [A,4]
    decoded+clipped.  (The empty case is reported as K=0 here; the zeros triple
    of retinanet.py:308 is a module-level quirk.)"""
    score = torch.max(cls, dim=1)[0]
    mask = score > score_thr
    surv = torch.nonzero(mask).flatten()
    b = boxes[surv]
    s = score[surv]
    if surv.numel() == 0:
        keep = torch.zeros((0,), dtype=torch.int64)
    elif use_torchvision:
        from torchvision.ops import nms
        keep = nms(b, s, iou_thr)
    else:
        keep = torch.from_numpy(greedy_nms(b.numpy(), s.numpy(), iou_thr))
    if keep.numel():
        ks, kl = cls[surv][keep].max(dim=1)
    else:
        ks = torch.zeros((0,), dtype=torch.float32)
        kl = torch.zeros((0,), dtype=torch.int64)
    return Detections(surv, keep, surv[keep], ks, kl, b[keep])


def detect_batch(classifications: torch.Tensor, regressions: torch.Tensor,
                 anchors: torch.Tensor, height: int, width: int,
                 score_thr: float = 0.05, iou_thr: float = 0.5, yx: bool = False,
                 use_torchvision: bool = True) -> Tuple[torch.Tensor, List[Detections]]:
    """utils/utils.py:76-114 semantics for every image of the batch.
    Returns (decoded+clipped boxes [B,A,4], per-image detections)."""
    anc = anchors if anchors.dim() == 3 else anchors[None]
    boxes = decode_boxes_yx(anc, regressions) if yx else decode_boxes(anc, regressions)
    boxes = clip_boxes_(boxes, height, width)
    dets = [detect_image(classifications[i], boxes[i], score_thr, iou_thr, use_torchvision)
            for i in range(classifications.shape[0])]
    return boxes, dets


# --------------------------------------------------------------------------- head outputs
def head_logits(logits_per_level, num_anchors: int, num_classes: int) -> torch.Tensor:
    """The layout step of the classification head without its activation
    (architectures/retinanet.py:162-178, ksevendet/architecture/head/simple_head.py:289-307):
    per level NCHW [B, num_anchors*num_classes, H, W] -> permute(0,2,3,1) -> [B, H*W*num_anchors, C],
    levels concatenated along dim 1 (retinanet.py:286).  Anchor order: cell row, cell column, in-cell
    anchor - the order Anchors.forward generates (anchors.py:99-119)."""
    out = []
    for z in logits_per_level:
        b = z.shape[0]
        assert z.shape[1] == num_anchors * num_classes
        out.append(z.permute(0, 2, 3, 1).contiguous().view(b, -1, num_classes))
    return torch.cat(out, dim=1)


def head_probabilities(logits_per_level, num_anchors: int, num_classes: int, act_first: bool = True) -> torch.Tensor:
    """What the reference hands to FocalLoss / the eval tail: `output_act = nn.Sigmoid()` on the NCHW
    logits, then the layout (retinanet.py:159-178; act_first=True), or the layout and then `header_act`
    on the concatenated [B,A,C] tensor (simple_head.py:289-310; act_first=False).  Mathematically the
    same; on CPU the two orders can differ in the last bit of single elements (ATen's vectorised
    sigmoid handles the tail of a buffer with another code path), so each is restated in its own order."""
    if act_first:
        return head_logits([torch.sigmoid(z) for z in logits_per_level], num_anchors, num_classes)
    return torch.sigmoid(head_logits(logits_per_level, num_anchors, num_classes))


# ------------------------------------------------------------------ evaluator tails
def eval_rows_coco(scores, labels, boxes, counts, scales, image_ids, category_of_label,
                   threshold: float = 0.05, valid_class_id=None) -> np.ndarray:
    """datasettool/coco_eval.py:25-57 and ksevendata_eval.py:24-61 for a batch: per image the model's
    (scores[n], labels[n], boxes[n,4]) in descending score order become result rows
    (image_id, x, y, w, h, score, category_id) — float64 [N,7], images in order.
      boxes /= scale                      float32 tensor divided by a Python float (coco_eval.py:30)
      boxes[:, 2] -= boxes[:, 0] ...      xyxy -> xywh in float32 (coco_eval.py:34-35)
      labels outside valid_class_id are skipped BEFORE the threshold test (ksevendata_eval.py:46-47)
      the loop BREAKS at the first row whose float(score) < threshold (coco_eval.py:45-46)."""
    rows = []
    for b in range(len(counts)):
        n = int(counts[b])
        if n == 0:
            continue
        sc = torch.as_tensor(np.asarray(scores[b][:n])).float().clone()
        lb = np.asarray(labels[b][:n])
        bx = torch.as_tensor(np.asarray(boxes[b][:n])).float().clone()
        bx /= float(scales[b])
        bx[:, 2] -= bx[:, 0]
        bx[:, 3] -= bx[:, 1]
        for i in range(n):
            score, label = float(sc[i]), int(lb[i])
            if valid_class_id is not None and len(valid_class_id) and label not in valid_class_id:
                continue
            if score < threshold:
                break
            rows.append([float(image_ids[b]), *[float(v) for v in bx[i].tolist()], score, float(category_of_label[label])])
    return np.array(rows, np.float64).reshape(-1, 7)


def eval_rows_csv(scores, labels, boxes, counts, scales, num_classes: int, score_threshold: float = 0.05,
                  max_detections: int = 100):
    """datasettool/csv_eval.py:93-121 (_get_detections) for a batch: rows (image index, x1, y1, x2, y2,
    score, label) float64 [N,7] ordered by (image, label, rank) and the exclusive offsets of every
    (image, label) bucket [B*num_classes + 1] — all_detections[b][c] == rows[off[b*nc+c]:off[b*nc+c+1], 1:6].
      boxes /= scale                       float32 numpy array by a Python float (csv_eval.py:96)
      indices = where(scores > thr)        float32 comparison (the Python float is cast down)
      argsort(-scores)[:max_detections]    numpy's default sort is not stable; the inputs are already in
                                           descending order, equal scores keep their input order here."""
    rows, offs = [], [0]
    for b in range(len(counts)):
        n = int(counts[b])
        sc = np.asarray(scores[b][:n], np.float32)
        lb = np.asarray(labels[b][:n])
        bx = np.asarray(boxes[b][:n], np.float32).copy()
        bx /= np.float32(scales[b])
        idx = np.where(sc > np.float32(score_threshold))[0]
        order = np.argsort(-sc[idx], kind="stable")[:max_detections] if max_detections > 0 else np.argsort(-sc[idx], kind="stable")
        sel = idx[order]
        for c in range(num_classes):
            for i in sel[lb[sel] == c]:
                rows.append([float(b), *[float(v) for v in bx[i]], float(sc[i]), float(c)])
            offs.append(len(rows))
    return np.array(rows, np.float64).reshape(-1, 7), np.array(offs, np.int32)


# ------------------------------------------------------------------ collater, annotation half
def collate_annotations(annots) -> torch.Tensor:
    """datasettool/dataloader.py:672-688: pad every image's [n_i,5] rows to the batch maximum with -1;
    a batch without any annotation gives [B,1,5] of -1."""
    annots = [torch.as_tensor(np.asarray(a, dtype=np.float32)).reshape(-1, 5) for a in annots]
    max_num = max(a.shape[0] for a in annots)
    if max_num > 0:
        out = torch.ones((len(annots), max_num, 5)) * -1
        for i, a in enumerate(annots):
            if a.shape[0] > 0:
                out[i, :a.shape[0], :] = a
        return out
    return torch.ones((len(annots), 1, 5)) * -1
