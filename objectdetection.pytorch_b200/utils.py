"""Inference-side drop-ins.

  BBoxTransform, ClipBoxes      architectures/utils.py:6-71
  BBoxTransformYX               architectures/utils_efficientdet.py:7-35
  detect()                      the eval tail of every detector for the WHOLE batch
                                (architectures/retinanet.py:297-318, utils/utils.py:76-114)
  postprocess()                 utils/utils.py:76-114 (same signature, same list-of-dicts return)
  nms()                         torchvision.ops.nms call sites (retinanet.py:314 ...)
  eval_tail() / patch_model()   the (scores, labels, boxes) return of model(img, return_loss=False)

All of them call libk7det.so; none has a PyTorch/CPU fallback.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass
from typing import List, Optional

import numpy as np
import torch
import torch.nn as nn

from . import _lib as L
from .anchors import spec_of
from .losses import _WS


def _decode_params(variant, mean, std, clip, height, width) -> L.K7DecodeParams:
    p = L.K7DecodeParams()
    p.variant = variant
    for i in range(4):
        p.mean[i] = float(mean[i])
        p.std[i] = float(std[i])
    p.clip, p.height, p.width = int(bool(clip)), int(height), int(width)
    return p


def _to4(v, default):
    if v is None:
        return default
    if isinstance(v, torch.Tensor):
        return [float(x) for x in v.detach().cpu().reshape(-1).tolist()]
    return [float(x) for x in np.asarray(v).reshape(-1).tolist()]


def decode(boxes, deltas: torch.Tensor, *, variant=L.DECODE_RETINA, mean=(0, 0, 0, 0), std=(.1, .1, .2, .2),
           clip_hw=None) -> torch.Tensor:
    """k7_decode wrapper: boxes = anchors [1|B,A,4] tensor (or anything carrying a k7_spec)."""
    L.require_cuda(deltas)
    d = L.as_f32_contig(deltas)
    B, A = d.shape[0], d.shape[1]
    dev = d.device
    spec = spec_of(boxes)
    anc_t, stride = None, 0
    order = L.BOX_YXYX if variant == L.DECODE_EFFDET else L.BOX_XYXY
    if spec is None:
        L.require_cuda(boxes)
        anc_t = L.as_f32_contig(boxes)
        if anc_t.dim() != 3 or anc_t.shape[1] != A or anc_t.shape[0] not in (1, B):
            raise ValueError("boxes must be [1,A,4] or [B,A,4] matching deltas [B,A,4]")
        stride = 0 if anc_t.shape[0] == 1 else A * 4
    out = torch.empty((B, A, 4), dtype=torch.float32, device=dev)
    h, w = clip_hw if clip_hw is not None else (0, 0)
    p = _decode_params(variant, mean, std, clip_hw is not None, h, w)
    with torch.cuda.device(dev):
        L.check(L.lib().k7_decode(C.byref(spec) if spec is not None else None, L.ptr(anc_t), order, stride,
                                  L.ptr(d), B, A, C.byref(p), L.ptr(out), L.stream_ptr(dev)), "k7_decode")
    return out


class BBoxTransform(nn.Module):
    """architectures/utils.py:6-53.  mean/std are plain attributes (no buffers), as in the reference."""

    def __init__(self, mean=None, std=None):
        super().__init__()
        self.mean = torch.from_numpy(np.array([0, 0, 0, 0]).astype(np.float32)) if mean is None else mean
        self.std = torch.from_numpy(np.array([0.1, 0.1, 0.2, 0.2]).astype(np.float32)) if std is None else std
        self._mean4 = _to4(self.mean, [0.0] * 4)
        self._std4 = _to4(self.std, [0.1, 0.1, 0.2, 0.2])

    def forward(self, boxes, deltas):
        return decode(boxes, deltas, variant=L.DECODE_RETINA, mean=self._mean4, std=self._std4)


class BBoxTransformYX(nn.Module):
    """architectures/utils_efficientdet.py:7-35: anchors (y1,x1,y2,x2), regression (dy,dx,dh,dw)."""

    def forward(self, anchors, regression):
        return decode(anchors, regression, variant=L.DECODE_EFFDET)


class ClipBoxes(nn.Module):
    """architectures/utils.py:56-71: in place, returns the same tensor."""

    def __init__(self, width=None, height=None):
        super().__init__()

    def forward(self, boxes, img):
        L.require_cuda(boxes)
        if boxes.dtype != torch.float32 or not boxes.is_contiguous() or boxes.data_ptr() % 16:
            raise ValueError("ClipBoxes works in place on a contiguous float32 [B,A,4] tensor")
        _, _, height, width = img.shape
        B, A = boxes.shape[0], boxes.shape[1]
        with torch.cuda.device(boxes.device):
            L.check(L.lib().k7_clip(L.ptr(boxes), B, A, int(height), int(width), L.stream_ptr(boxes.device)), "k7_clip")
        return boxes


# ------------------------------------------------------------------ fused eval tail
@dataclass
class BatchDetections:
    """Kept detections of a batch, row-padded to `cap` per image.  count[b] rows are valid."""
    count: torch.Tensor          # [B] int64 (host)
    survivors: torch.Tensor      # [B] int64 (host): boxes above the score threshold
    scores: torch.Tensor         # [B,cap] f32, descending per image
    labels: torch.Tensor         # [B,cap] i64
    boxes: torch.Tensor          # [B,cap,4] f32 (x1,y1,x2,y2)
    anchor_ids: torch.Tensor     # [B,cap] i32
    survivor_rank: torch.Tensor  # [B,cap] i32: what torchvision.ops.nms would have returned

    def image(self, b: int):
        k = int(self.count[b])
        return self.scores[b, :k], self.labels[b, :k], self.boxes[b, :k]


def detect(classifications: torch.Tensor, regressions: torch.Tensor, anchors, image_hw, *,
           score_thr: float = 0.05, iou_thr: float = 0.5, variant=L.DECODE_RETINA, mean=(0, 0, 0, 0),
           std=(.1, .1, .2, .2), clip: bool = True, pre_nms_topk: int = 0, max_detections: int = 0,
           cap: Optional[int] = None, sync: bool = True, from_logits: bool = False,
           class_aware: bool = False, want_rank: Optional[bool] = None, nms_chain: str = "auto") -> BatchDetections:
    """Fused decode + clip + threshold + sort + NMS for every image of the batch.
    class_aware: suppress only inside a class (torchvision.ops.batched_nms with idxs = labels); the
    reference's own eval tail is class-agnostic (retinanet.py:314), which is the default.
    from_logits: `classifications` are pre-sigmoid logits; scores are torch.sigmoid of the row maxima
    and threshold / label ties are decided on those float32 probabilities, as the reference would.
    `classifications` may also be the LIST of the head's raw per-level NCHW outputs
    [B, anchors*classes, H_l, W_l] (then `anchors` must come from b200det.Anchors).
    want_rank: also return survivor_rank (the index torchvision.ops.nms would have returned: one more small
    kernel for the popcount prefix of the survivor bitmap); default: only when sync=True.
    nms_chain: "auto" (per-image kernels for many or small images, multi-SM kernels for a few large ones),
    "per_image" or "multi_sm" - same result, K7DetectParams.nms_chain.
    sync=False returns at once, count / survivors stay on the device and at most `cap` (default 1024) kept
    boxes per image are returned - a larger count[b] says how many there were."""
    if want_rank is None:
        want_rank = bool(sync)
    levels = list(classifications) if isinstance(classifications, (list, tuple)) else None
    L.require_cuda(regressions)
    reg = L.as_f32_contig(regressions)
    spec = spec_of(anchors)
    anc_t = None
    order = L.BOX_YXYX if variant == L.DECODE_EFFDET else L.BOX_XYXY
    if levels is not None:
        # the head's raw per-level outputs [B, anchors*classes, H_l, W_l]
        B, Cn = L.check_levels(levels, spec)
        A = spec.total
        dev = reg.device
        cls = None
    else:
        L.require_cuda(classifications)
        cls = L.as_f32_contig(classifications)
        B, A, Cn = cls.shape
        dev = cls.device
        if spec is None:
            L.require_cuda(anchors)
            anc_t = L.as_f32_contig(anchors.reshape(-1, 4))
            if anc_t.shape[0] != A:
                raise ValueError("anchor count mismatch")
    lib = L.lib()
    ws_bytes = lib.k7_detect_workspace_bytes(B, A, Cn)
    ws = _WS.get("detect", ws_bytes, dev)
    counts = torch.empty((B, 2), dtype=torch.int32, device=dev)
    p = L.K7DetectParams()
    p.decode = _decode_params(variant, mean, std, clip, image_hw[0], image_hw[1])
    p.score_thr, p.iou_thr = float(score_thr), float(iou_thr)
    p.pre_nms_topk, p.max_detections = int(pre_nms_topk), int(max_detections)
    p.from_logits = int(from_logits)
    p.class_aware = int(class_aware)
    try:
        p.nms_chain = {"auto": 0, "per_image": 1, "multi_sm": 2}[nms_chain]
    except KeyError:
        raise ValueError("nms_chain must be 'auto', 'per_image' or 'multi_sm'")
    if cap is None:
        cap = max_detections if max_detections > 0 else min(A, 1024)
    cap = max(1, min(int(cap), A))

    def gather(capn):
        scores = torch.empty((B, capn), dtype=torch.float32, device=dev)
        labels = torch.empty((B, capn), dtype=torch.int64, device=dev)
        boxes = torch.empty((B, capn, 4), dtype=torch.float32, device=dev)
        aids = torch.empty((B, capn), dtype=torch.int32, device=dev)
        ranks = torch.empty((B, capn), dtype=torch.int32, device=dev) if want_rank else None
        L.check(lib.k7_detect_gather(L.ptr(ws), ws.numel(), B, A, Cn, capn, L.ptr(scores), L.ptr(labels),
                                     L.ptr(boxes), L.ptr(aids), L.ptr(ranks), L.stream_ptr(dev)), "k7_detect_gather")
        return scores, labels, boxes, aids, ranks

    with torch.cuda.device(dev):
        if levels is not None:
            L.check(lib.k7_detect_levels(C.byref(spec), L.ptr_array(levels), L.ptr(reg), B, Cn, C.byref(p), L.ptr(counts),
                                         L.ptr(ws), ws.numel(), L.stream_ptr(dev)), "k7_detect_levels")
        else:
            L.check(lib.k7_detect(C.byref(spec) if spec is not None else None, L.ptr(anc_t), order, L.ptr(cls),
                                  L.ptr(reg), B, A, Cn, C.byref(p), L.ptr(counts), L.ptr(ws), ws.numel(),
                                  L.stream_ptr(dev)), "k7_detect")
        out = gather(cap)
        if not sync:
            return BatchDetections(counts[:, 1], counts[:, 0], *out)
        host = counts.cpu()                       # the one synchronisation of the eval path
        kmax = int(host[:, 1].max())
        if kmax > cap:                            # rare: more kept boxes than the first guess
            out = gather(kmax)
    return BatchDetections(host[:, 1].long(), host[:, 0].long(), *out)


def nms(boxes: torch.Tensor, scores: torch.Tensor, iou_threshold: float) -> torch.Tensor:
    """torchvision.ops.nms(boxes, scores, iou_threshold) on device: int64 indices, descending score."""
    L.require_cuda(boxes, scores)
    b = L.as_f32_contig(boxes)
    s = L.as_f32_contig(scores)
    n = b.shape[0]
    dev = b.device
    lib = L.lib()
    keep = torch.empty(max(n, 1), dtype=torch.int64, device=dev)
    count = torch.empty(1, dtype=torch.int32, device=dev)
    ws = _WS.get("nms", lib.k7_nms_workspace_bytes(n), dev)
    with torch.cuda.device(dev):
        L.check(lib.k7_nms(L.ptr(b), L.ptr(s), n, float(iou_threshold), L.ptr(keep), L.ptr(count), L.ptr(ws),
                           ws.numel(), L.stream_ptr(dev)), "k7_nms")
    return keep[:int(count.item())]


def batched_nms(boxes: torch.Tensor, scores: torch.Tensor, idxs: torch.Tensor, iou_threshold: float,
                strategy: str = "vanilla") -> torch.Tensor:
    """torchvision.ops.batched_nms(boxes, scores, idxs, iou_threshold) on device: NMS
    inside each category, int64 indices in descending-score order.
    strategy "vanilla": exact per-category NMS (torchvision's _batched_nms_vanilla) in ONE pass of the
    bitmask kernels, the category test fused behind the overlap filter;
    "coordinate_trick": torchvision's _batched_nms_coordinate_trick (boxes shifted by
    idxs * (boxes.max() + 1), then plain NMS) - the float32 shift rounds the coordinates, so its result can
    differ from the exact one exactly as torchvision's does;
    "auto": torchvision 0.26's own rule on CUDA (boxes.numel() > 100_000 -> vanilla)."""
    L.require_cuda(boxes, scores, idxs)
    if strategy == "auto":
        strategy = "vanilla" if boxes.numel() > 100_000 else "coordinate_trick"
    if strategy == "coordinate_trick":
        if boxes.numel() == 0:
            return torch.empty((0,), dtype=torch.int64, device=boxes.device)
        max_coordinate = boxes.max()
        offsets = idxs.to(boxes) * (max_coordinate + torch.tensor(1).to(boxes))
        return nms(boxes + offsets[:, None], scores, iou_threshold)
    if strategy != "vanilla":
        raise ValueError("strategy must be 'vanilla', 'coordinate_trick' or 'auto'")
    b = L.as_f32_contig(boxes)
    s = L.as_f32_contig(scores)
    i = idxs.to(torch.int64).contiguous()
    n = b.shape[0]
    if s.shape[0] != n or i.shape[0] != n:
        raise ValueError("boxes, scores and idxs must have the same length")
    dev = b.device
    lib = L.lib()
    keep = torch.empty(max(n, 1), dtype=torch.int64, device=dev)
    count = torch.empty(1, dtype=torch.int32, device=dev)
    ws = _WS.get("nms", lib.k7_nms_workspace_bytes(n), dev)
    with torch.cuda.device(dev):
        L.check(lib.k7_batched_nms(L.ptr(b), L.ptr(s), L.ptr(i), n, float(iou_threshold), L.ptr(keep), L.ptr(count),
                                   L.ptr(ws), ws.numel(), L.stream_ptr(dev)), "k7_batched_nms")
    return keep[:int(count.item())]


def soft_nms(boxes: torch.Tensor, scores: torch.Tensor, method: str = "gaussian", iou_threshold: float = 0.3,
             sigma: float = 0.5, score_threshold: float = 0.001):
    """Soft-NMS (Bodla et al., ICCV 2017) on device: returns (keep int64 [K] in pick order, scores float32 [K] - the
    decayed score each box had when it was picked).  method "linear": s_j *= 1 - iou where iou > iou_threshold;
    "gaussian": s_j *= exp(-iou^2 / sigma); picking stops below score_threshold.  An extension:
    the reference's detectors use hard NMS only."""
    L.require_cuda(boxes, scores)
    if method not in ("linear", "gaussian"):
        raise ValueError("method must be 'linear' or 'gaussian'")
    b = L.as_f32_contig(boxes)
    s = L.as_f32_contig(scores)
    n = b.shape[0]
    if s.shape[0] != n:
        raise ValueError("boxes and scores must have the same length")
    dev = b.device
    keep = torch.empty(max(n, 1), dtype=torch.int64, device=dev)
    out = torch.empty(max(n, 1), dtype=torch.float32, device=dev)
    count = torch.empty(1, dtype=torch.int32, device=dev)
    p = L.K7SoftNmsParams(0 if method == "linear" else 1, float(iou_threshold), float(sigma), float(score_threshold))
    with torch.cuda.device(dev):
        L.check(L.lib().k7_soft_nms(L.ptr(b), L.ptr(s), n, C.byref(p), L.ptr(keep), L.ptr(out), L.ptr(count),
                                    L.stream_ptr(dev)), "k7_soft_nms")
    k = int(count.item())
    return keep[:k], out[:k]


def _variant_of(regressBoxes) -> int:
    return L.DECODE_EFFDET if isinstance(regressBoxes, BBoxTransformYX) or \
        type(regressBoxes).__module__.endswith("utils_efficientdet") else L.DECODE_RETINA


def postprocess(x, anchors, regression, classification, regressBoxes, clipBoxes, threshold, iou_threshold,
                from_logits: bool = False, class_aware: bool = False):
    """utils/utils.py:76-114 — same arguments, same return (list of dicts of numpy arrays).  The
    reference's duplicate empty dict for an all-empty batch (utils.py:83-88) is not reproduced."""
    variant = _variant_of(regressBoxes)
    mean = _to4(getattr(regressBoxes, "mean", None), [0.0] * 4)
    std = _to4(getattr(regressBoxes, "std", None), [0.1, 0.1, 0.2, 0.2])
    det = detect(classification, regression, anchors, (int(x.shape[2]), int(x.shape[3])), score_thr=threshold,
                 iou_thr=iou_threshold, variant=variant, mean=mean, std=std, clip=clipBoxes is not None,
                 from_logits=from_logits, class_aware=class_aware)
    out = []
    sc, lb, bx = det.scores.cpu().numpy(), det.labels.cpu().numpy(), det.boxes.cpu().numpy()
    for b in range(classification.shape[0]):
        k = int(det.count[b])
        if k == 0:
            out.append({'rois': np.array(()), 'class_ids': np.array(()), 'scores': np.array(())})
        else:
            out.append({'rois': bx[b, :k], 'class_ids': lb[b, :k], 'scores': sc[b, :k]})
    return out


def eval_tail(classification, regression, anchors, img_batch, regressBoxes=None, score_thr=0.05, iou_thr=0.5,
              from_logits: bool = False):
    """What model(img, return_loss=False) returns after the heads (retinanet.py:297-318):
    [scores[K], labels[K] (int64), boxes[K,4]] of IMAGE 0, or the zeros triple when nothing passes."""
    variant = _variant_of(regressBoxes) if regressBoxes is not None else L.DECODE_RETINA
    mean = _to4(getattr(regressBoxes, "mean", None), [0.0] * 4)
    std = _to4(getattr(regressBoxes, "std", None), [0.1, 0.1, 0.2, 0.2])
    det = detect(classification[:1], regression[:1], anchors, (int(img_batch.shape[2]), int(img_batch.shape[3])),
                 score_thr=score_thr, iou_thr=iou_thr, variant=variant, mean=mean, std=std, from_logits=from_logits)
    if int(det.survivors[0]) == 0:
        dev = classification.device                # retinanet.py:308
        return [torch.zeros([1], device=dev), torch.zeros([1], device=dev), torch.zeros([1, 4], device=dev)]
    s, l, b = det.image(0)
    return [s, l, b]


def strip_head_sigmoid(model) -> bool:
    """Replace the classification head's final sigmoid by the identity so the head returns logits:
    ``classificationModel.output_act`` (architectures/retinanet.py:140,159) or the shared head's
    ``header_act`` (ksevendet/architecture/head/simple_head.py:310).  Returns True if one was found."""
    found = False
    for mod in model.modules():
        for name in ("output_act", "header_act"):
            if isinstance(getattr(mod, name, None), nn.Sigmoid):
                setattr(mod, name, nn.Identity())
                found = True
    return found


def patch_model(model, score_thr: float = 0.05, iou_thr: float = 0.5, fuse_sigmoid: bool = False):
    """Swap a reference detector's hot-path attributes for the B200 modules and route its eval
    branch through eval_tail().  Works on every detector forward of the reference:

        retinanet.py:272, architectures/ksevendet.py:282   forward(img, annotations=None, return_head=False, return_loss=True)
        ksevendet/architecture/ksevendet.py:222            the same with return_loss=False as the default
        efficientdet.py:528                                forward(images, annotations=None, return_loss=True)   (no return_head)

    The only seam ALL of them share after the conv heads is
    ``self.focalLoss(classification, regression, anchors, annotations)`` (retinanet.py:294,
    efficientdet.py:545, ksevendet.py:312 / :274): an eval call runs the model's own forward with
    return_loss=True while the swapped-in FocalLoss is in capture mode (thread-local), which hands
    back the head outputs instead of a loss; the fused tail then produces the reference's
    ``[scores[K], labels[K], boxes[K,4]]`` (or its zeros triple).  The model's own defaults are kept
    (inspect.signature), ``return_head`` is only forwarded to forwards that have it, and the
    override lives on a subclass of the model's class, so nn.DataParallel replicas get it too.
    fuse_sigmoid: also remove the classification head's final sigmoid (strip_head_sigmoid) and run
    FocalLoss / the eval tail on the logits; ``return_head=True`` then returns logits."""
    import inspect
    from .anchors import Anchors
    from .losses import FocalLoss
    for name in ("anchors", "focalLoss", "regressBoxes", "clipBoxes"):
        if not hasattr(model, name):
            raise ValueError(f"patch_model: the model has no attribute '{name}' (retinanet.py:220-230)")
    from_logits = bool(fuse_sigmoid) and strip_head_sigmoid(model)
    old = model.anchors
    model.anchors = Anchors(pyramid_levels=list(old.pyramid_levels), strides=list(old.strides),
                            sizes=list(old.sizes), ratios=np.asarray(old.ratios), scales=np.asarray(old.scales))
    model.focalLoss = FocalLoss(from_logits=from_logits)
    model.regressBoxes = BBoxTransform()
    model.clipBoxes = ClipBoxes()

    base = type(model)
    if getattr(base, "_k7_patched", False):          # patched twice: rebuild on the original class
        base = base.__mro__[1]
    inner = base.forward
    sig = inspect.signature(inner)
    names = list(sig.parameters)                      # self, img_batch|images, annotations, ...
    if "return_loss" not in sig.parameters or len(names) < 3:
        raise ValueError("patch_model: forward(img, annotations=None, ..., return_loss=...) expected "
                         f"(retinanet.py:272, efficientdet.py:528); got {sig}")
    has_head = "return_head" in sig.parameters
    img_name = names[1]

    def forward(self, *args, **kwargs):
        bound = sig.bind(self, *args, **kwargs)
        bound.apply_defaults()                        # the model's OWN defaults (ksevendet: return_loss=False)
        a = bound.arguments
        if a["return_loss"] or (has_head and a["return_head"]) or getattr(self, "convert_onnx", False):
            return inner(self, *args, **kwargs)
        img_batch = a[img_name]
        extra = {"return_head": False} if has_head else {}
        with self.focalLoss.capture_head():           # retinanet.py:294 then returns the head outputs
            cap = inner(self, img_batch, None, return_loss=True, **extra)
        return eval_tail(cap.classification, cap.regression, cap.anchors, img_batch, self.regressBoxes,
                         score_thr, iou_thr, from_logits=from_logits)

    forward.__signature__ = sig
    model.__class__ = type(base.__name__, (base,), {"forward": forward, "_k7_patched": True,
                                                    "__module__": base.__module__})
    return model
