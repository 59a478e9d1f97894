"""FocalLoss — drop-in for architectures/losses.py:25-168 (and, through ctor kwargs, for
architectures/loss_efficientdet.py:27-160).

    FocalLoss()(classifications[B,A,C], regressions[B,A,4], anchors[1,A,4], annotations[B,M,5])
        -> (classification_loss[1], regression_loss[1])        differentiable

Two CUDA stages (csrc/k7_anchor_match.cu, csrc/k7_loss.cu) replace the reference's
per-image Python loop of ~45 ATen launches: `k7_match` (IoU + assignment, 4 B/anchor out)
and `k7_loss` (one streaming pass that produces the loss sums and, when an input
requires grad, d(loss)/d(classifications|regressions) in the same pass).  The
backward therefore only hands the saved gradients back (scaled on device when the
incoming gradient is not 1).  No host synchronisation happens in forward or backward.

Sharded training (one process per GPU, images split along dim 0): pass
``process_group=`` (or call ``b200det.distributed.sharded_focal_loss``): the local
sums are combined with ONE 4-float all-reduce and gradients are pre-scaled by the
global batch, so backward needs no collective.
"""
from __future__ import annotations

import contextlib
import ctypes as C
import threading
from collections import namedtuple
from typing import Optional

import torch
import torch.nn as nn

from . import _lib as L
from .anchors import spec_of


class _Workspace:
    """Scratch tensors, grown on demand and reused across calls.  One set per (device, CUDA stream):
    two callers that overlap on different streams of one device (an eval thread next to a training
    thread, detect() of two half batches) never share a buffer - the library itself keeps no state,
    so this is the only place where calls could meet (re-entrancy)."""

    def __init__(self):
        self.bufs = {}
        self.best = {}

    @staticmethod
    def _key(name, device):
        device = torch.device(device)
        if device.index is None:
            device = torch.device("cuda", torch.cuda.current_device())
        return (name, device, torch.cuda.current_stream(device).cuda_stream)

    def get(self, name: str, nbytes: int, device) -> torch.Tensor:
        key = self._key(name, device)
        t = self.bufs.get(key)
        if t is None or t.numel() < nbytes:
            t = torch.empty(max(nbytes, 256), dtype=torch.uint8, device=device)
            self.bufs[key] = t
        return t

    def best_array(self, B: int, A: int, device):
        """The buffer of k7_match_scatter (uint64 [B*A] maxima + the list of touched anchors behind them, as
        int64 words) + the generation of this call + whether k7_loss has to zero what it reads.  Zeroed when
        created and whenever B or A change (the list of one shape must not be read as maxima of another);
        every call gets the next generation (stale entries are ignored by their tag); after K7_BEST_GEN_MAX
        calls the array is zeroed again.  While a CUDA graph is being captured the generation would be frozen
        into the graph, so the replays reuse ONE generation and k7_loss clears the entries it consumes
        (K7LossParams.consume_best)."""
        key = self._key("best", device)
        ent = self.best.get(key)
        words = (L.lib().k7_match_scatter_workspace_bytes(B, A) + 7) // 8
        if ent is None or ent[0].numel() < words:
            ent = [torch.zeros(words, dtype=torch.int64, device=device), 0, (B, A)]
            self.best[key] = ent
        if ent[1] >= L.BEST_GEN_MAX or ent[2] != (B, A):
            ent[0].zero_()
            ent[1] = 0
            ent[2] = (B, A)
        ent[1] += 1
        return ent[0], ent[1], bool(torch.cuda.is_current_stream_capturing())


_WS = _Workspace()


def scatter_ok(spec, M: int, neg_thr: float, pos_thr: float) -> bool:
    """Can the ground-truth-first assignment (k7_match_scatter) serve this call?"""
    return spec is not None and M <= 65535 and 1e-3 < neg_thr < 1.0 and pos_thr >= neg_thr


def match_scatter(spec, annotations: torch.Tensor, A: int, neg_thr: float, pos_thr: float,
                  zero_fill: Optional[torch.Tensor] = None):
    """k7_match_scatter: returns (best buffer as int64 words - [B*A] maxima, then the touched list -, generation,
    consume).  The positives are counted by the loss stage, which walks the touched list.
    zero_fill: optional [B,A,4] float32 tensor the kernel fills with zeros on the way (the d(reg) buffer of the
    loss stage that follows)."""
    L.require_cuda(annotations)
    ann = L.as_f32_contig(annotations)
    B, M = ann.shape[0], ann.shape[1]
    dev = ann.device
    best, gen, consume = _WS.best_array(B, A, dev)
    p = L.K7MatchParams(neg_thr, pos_thr)
    with torch.cuda.device(dev):
        L.check(L.lib().k7_match_scatter(C.byref(spec), L.ptr(ann), B, A, M, C.byref(p), L.ptr(best), gen,
                                         L.ptr(zero_fill), L.stream_ptr(dev)), "k7_match_scatter")
    return best, gen, consume


def match(anchors, annotations: torch.Tensor, A: int, neg_thr: float, pos_thr: float,
          box_order: int = L.BOX_XYXY, debug: bool = False, zero_fill: Optional[torch.Tensor] = None,
          with_bits: bool = False):
    """Assignment stage.  Returns (assign[B*A] int32, num_pos[B] int32[, iou_max, argmax][, ignore_bits]).
    zero_fill: optional [B,A,4] float32 tensor the kernel fills with zeros on the way (the d(reg)
    buffer of the loss stage that follows).  with_bits: also return the one-bit-per-anchor ignore
    bitmap [B, K7_IGNORE_WORDS(A)] int32 that the loss stage streams."""
    L.require_cuda(annotations)
    ann = L.as_f32_contig(annotations)
    B, M = ann.shape[0], ann.shape[1]
    dev = ann.device
    spec = spec_of(anchors)
    anc_t = None
    if spec is None:
        anc_t = L.as_f32_contig(anchors.reshape(-1, 4))
        L.require_cuda(anc_t)
    else:
        box_order = spec.box_order
    assign = torch.empty(B * A, dtype=torch.int32, device=dev)
    num_pos = torch.empty(B, dtype=torch.int32, device=dev)
    iou_max = torch.empty((B, A), dtype=torch.float32, device=dev) if debug else None
    argmax = torch.empty((B, A), dtype=torch.int32, device=dev) if debug else None
    bits = torch.empty((B, L.ignore_words(A)), dtype=torch.int32, device=dev) if with_bits else None
    p = L.K7MatchParams(neg_thr, pos_thr)
    with torch.cuda.device(dev):
        L.check(L.lib().k7_match(C.byref(spec) if spec is not None else None, L.ptr(anc_t), box_order,
                                 L.ptr(ann), B, A, M, C.byref(p), L.ptr(assign), L.ptr(num_pos),
                                 L.ptr(iou_max), L.ptr(argmax), L.ptr(zero_fill), L.ptr(bits),
                                 L.stream_ptr(dev)), "k7_match")
    out = (assign, num_pos)
    if debug:
        out += (iou_max, argmax)
    if with_bits:
        out += (bits,)
    return out


def _loss_params(*, alpha, beta, target_div, yx_targets, mean_divisor, want_grad, dreg_zeroed, from_logits, peer_reduce,
                 best, assign_out, spare_sms=0):
    p = L.K7LossParams()
    p.spare_sms = int(spare_sms)
    p.dreg_prezeroed = int(want_grad and dreg_zeroed is not None)
    p.alpha, p.beta = alpha, beta
    for i in range(4):
        p.target_div[i] = target_div[i]
    p.yx_targets, p.mean_divisor = int(yx_targets), int(mean_divisor)
    p.from_logits = int(from_logits)
    if peer_reduce is not None:                # K7PeerReduce: `sums` come back as the GLOBAL sums
        p.peer_reduce = C.pointer(peer_reduce)
    if best is not None:                       # (array, generation, consume, neg_thr, pos_thr) of match_scatter
        p.match_generation, p.consume_best = int(best[1]), int(best[2])
        p.neg_thr, p.pos_thr = float(best[3]), float(best[4])
    if assign_out is not None:
        p.assign_out = assign_out.data_ptr()
    return p


def loss_stage(anchors, cls, reg, ann, assign, num_pos=None, *, alpha, beta, target_div, yx_targets,
               mean_divisor, want_grad, box_order=L.BOX_XYXY, dreg_zeroed: Optional[torch.Tensor] = None,
               ignore_bits: Optional[torch.Tensor] = None, from_logits: bool = False, peer_reduce=None,
               best=None, assign_out: Optional[torch.Tensor] = None, spare_sms: int = 0):
    """Streaming loss stage.  Returns (per_image[B,2], sums[4], dcls|None, dreg|None).
    `num_pos` [B] int32 is an OUTPUT (the positives are counted while the assignment is walked); pass a
    tensor to receive it.  The assignment comes as match()'s `assign` words or as match_scatter()'s array:
    best = (array, generation, consume, neg_thr, pos_thr) with assign=None.
    dreg_zeroed: a [B,A,4] buffer match() already filled with zeros (saves a 16 B/anchor store);
    ignore_bits: accepted for compatibility, unused (ignored rows are handled by the row kernel);
    from_logits: `cls` holds pre-sigmoid logits, the sigmoid is fused and dcls is d(loss)/d(logit);
    assign_out: optional [B*A] int32 that receives k7_match's words when the input is `best`."""
    B, A, Cn = cls.shape
    M = ann.shape[1]
    dev = cls.device
    spec = spec_of(anchors)
    anc_t = None
    if spec is None:
        anc_t = L.as_f32_contig(anchors.reshape(-1, 4))
    else:
        box_order = spec.box_order
    lib = L.lib()
    ws_bytes = lib.k7_loss_workspace_bytes(B, A, Cn)
    ws = _WS.get("loss", ws_bytes, dev)
    per_image = torch.empty((B, 2), dtype=torch.float32, device=dev)
    sums = torch.empty(4, dtype=torch.float32, device=dev)
    if num_pos is None:
        num_pos = torch.empty(B, dtype=torch.int32, device=dev)
    dcls = torch.empty_like(cls) if want_grad else None
    dreg = None
    if want_grad:
        dreg = dreg_zeroed if dreg_zeroed is not None else torch.empty_like(reg)
    p = _loss_params(alpha=alpha, beta=beta, target_div=target_div, yx_targets=yx_targets, mean_divisor=mean_divisor,
                     want_grad=want_grad, dreg_zeroed=dreg_zeroed, from_logits=from_logits, peer_reduce=peer_reduce,
                     best=best, assign_out=assign_out, spare_sms=spare_sms)
    with torch.cuda.device(dev):
        L.check(lib.k7_loss(C.byref(spec) if spec is not None else None, L.ptr(anc_t), box_order,
                            L.ptr(cls), L.ptr(reg), L.ptr(ann), L.ptr(assign), L.ptr(best[0]) if best is not None else None,
                            L.ptr(num_pos), B, A, Cn, M, C.byref(p), L.ptr(per_image), L.ptr(sums), L.ptr(dcls), L.ptr(dreg),
                            L.ptr(ws), ws.numel(), L.stream_ptr(dev)), "k7_loss")
    return per_image, sums, dcls, dreg


def _scale_grads(pairs, dev):
    """grad *= incoming gradient (a device scalar; a no-op inside the kernel when it is 1) for every
    (grad, scale) pair, one launch per eight tensors."""
    lib = L.lib()
    with torch.cuda.device(dev):
        st = L.stream_ptr(dev)
        for i in range(0, len(pairs), 8):
            chunk = pairs[i:i + 8]
            g = L.ptr_array([t for t, _ in chunk])
            sc = L.ptr_array([x for _, x in chunk])
            n = (C.c_size_t * len(chunk))(*[t.numel() for t, _ in chunk])
            L.check(lib.k7_scale_inplace_multi(g, n, sc, len(chunk), st), "k7_scale_inplace_multi")


def _assignment(anchor_src, ann, A, cfg, dreg0):
    """The assignment stage of a FocalLoss call: (assign|None, best|None, num_pos, dreg_zeroed|None).
    Anchors from a recipe: ground-truth-first k7_match_scatter; a foreign anchor tensor (or thresholds
    it cannot serve): the exhaustive k7_match, which also zero-fills d(reg) on its way."""
    sp = spec_of(anchor_src)
    num_pos = torch.empty(ann.shape[0], dtype=torch.int32, device=ann.device)     # written by the loss stage
    if scatter_ok(sp, ann.shape[1], cfg["neg_thr"], cfg["pos_thr"]):
        arr, gen, consume = match_scatter(sp, ann, A, cfg["neg_thr"], cfg["pos_thr"], zero_fill=dreg0)
        return None, (arr, gen, consume, cfg["neg_thr"], cfg["pos_thr"]), num_pos, dreg0
    assign, _ = match(anchor_src, ann, A, cfg["neg_thr"], cfg["pos_thr"], cfg["box_order"], zero_fill=dreg0)
    return assign, None, num_pos, dreg0


def _exchange(cfg):
    """(world, peer mailbox|None, group) of a sharded call."""
    group = cfg.get("process_group")
    world = 1
    if group is not None or cfg.get("sharded"):
        import torch.distributed as dist
        world = dist.get_world_size(group)
    peer = cfg["peer"]() if world > 1 and cfg.get("peer") else None   # fused exchange over peer memory, or None
    return world, peer, group


class _FocalLossFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, cls, reg, anchors, ann, cfg):
        want_grad = cls.requires_grad or reg.requires_grad
        A = cls.shape[1]
        dreg0 = torch.empty_like(reg) if want_grad else None
        assign, best, num_pos, dreg_zeroed = _assignment(anchors, ann, A, cfg, dreg0)
        world, peer, group = _exchange(cfg)
        B = cls.shape[0]
        per_image, sums, dcls, dreg = loss_stage(
            anchors, cls, reg, ann, assign, num_pos, alpha=cfg["alpha"], beta=cfg["beta"],
            target_div=cfg["target_div"], yx_targets=cfg["yx_targets"], mean_divisor=cfg["divisor"](B, world),
            want_grad=want_grad, box_order=cfg["box_order"], dreg_zeroed=dreg_zeroed, best=best,
            from_logits=cfg.get("from_logits", False), peer_reduce=peer.next() if peer is not None else None,
            spare_sms=cfg.get("spare_sms", 0))
        if want_grad and dreg_zeroed is None:
            dreg0 = None                         # loss_stage allocated d(reg) itself
        if world > 1 and peer is None:
            import torch.distributed as dist
            dist.all_reduce(sums, op=dist.ReduceOp.SUM, group=group)   # the path's only collective
        out = sums[:2] / sums[3]                   # mean over (global) images, losses.py:168
        ctx.save_for_backward(dcls, dreg) if want_grad else None
        ctx.has_grad = want_grad
        ctx.device = cls.device
        ctx.consumed = False
        ctx.set_materialize_grads(False)    # no zero tensors for the outputs nobody differentiates
        ctx.mark_non_differentiable(num_pos, per_image, sums)
        return out[0:1], out[1:2], num_pos, per_image, sums

    @staticmethod
    def backward(ctx, g_cls, g_reg, *_):
        if not ctx.has_grad:
            return None, None, None, None, None
        _once(ctx)
        dcls, dreg = ctx.saved_tensors
        pairs = []
        if g_cls is not None:
            pairs.append((dcls, g_cls.reshape(-1).float().contiguous()))
        else:
            dcls = None
        if g_reg is not None:
            pairs.append((dreg, g_reg.reshape(-1).float().contiguous()))
        else:
            dreg = None
        _scale_grads(pairs, ctx.device)
        return dcls, dreg, None, None, None


def _once(ctx):
    """The saved gradients are scaled IN PLACE by the incoming gradient (no second copy of the largest
    tensor of the step), so a second backward through the same graph (retain_graph=True) would scale
    them twice: refuse it instead of returning wrong numbers."""
    if ctx.consumed:
        raise RuntimeError("b200det.FocalLoss: backward was already run on this graph; its saved gradients were scaled "
                           "in place.  Call the loss again instead of backward(retain_graph=True) twice.")
    ctx.consumed = True


def loss_stage_levels(spec, levels, reg, ann, assign, num_pos=None, *, alpha, beta, target_div, yx_targets, mean_divisor,
                      want_grad, dreg_zeroed: Optional[torch.Tensor] = None, ignore_bits: Optional[torch.Tensor] = None,
                      from_logits: bool = False, peer_reduce=None, best=None, spare_sms: int = 0):
    """loss_stage on the head's raw per-level NCHW outputs (k7_loss_levels).  Returns
    (per_image[B,2], sums[4], dlevels|None (list, same layouts as `levels`), dreg|None)."""
    B, Cn = L.check_levels(levels, spec)
    M = ann.shape[1]
    dev = reg.device
    lib = L.lib()
    ws = _WS.get("loss_levels", lib.k7_loss_levels_workspace_bytes(C.byref(spec), B, Cn), dev)
    per_image = torch.empty((B, 2), dtype=torch.float32, device=dev)
    sums = torch.empty(4, dtype=torch.float32, device=dev)
    if num_pos is None:
        num_pos = torch.empty(B, dtype=torch.int32, device=dev)
    dlevels = [torch.empty_like(t) for t in levels] if want_grad else None
    dreg = None
    if want_grad:
        dreg = dreg_zeroed if dreg_zeroed is not None else torch.empty_like(reg)
    p = _loss_params(alpha=alpha, beta=beta, target_div=target_div, yx_targets=yx_targets, mean_divisor=mean_divisor,
                     want_grad=want_grad, dreg_zeroed=dreg_zeroed, from_logits=from_logits, peer_reduce=peer_reduce,
                     best=best, assign_out=None, spare_sms=spare_sms)
    lv, dlv = L.ptr_array(levels), L.ptr_array(dlevels)
    with torch.cuda.device(dev):
        L.check(lib.k7_loss_levels(C.byref(spec), lv, L.ptr(reg), L.ptr(ann), L.ptr(assign),
                                   L.ptr(best[0]) if best is not None else None,
                                   L.ptr(num_pos), B, Cn, M, C.byref(p), L.ptr(per_image), L.ptr(sums), dlv, L.ptr(dreg),
                                   L.ptr(ws), ws.numel(), L.stream_ptr(dev)), "k7_loss_levels")
    return per_image, sums, dlevels, dreg


class _FocalLossLevelsFn(torch.autograd.Function):
    """FocalLoss on per-level NCHW head outputs; inputs (reg, spec, ann, cfg, *levels)."""

    @staticmethod
    def forward(ctx, reg, spec, ann, cfg, *levels):
        want_grad = reg.requires_grad or any(t.requires_grad for t in levels)
        A = spec.total
        dreg0 = torch.empty_like(reg) if want_grad else None
        assign, best, num_pos, dreg_zeroed = _assignment(spec, ann, A, cfg, dreg0)
        world, peer, group = _exchange(cfg)
        B = reg.shape[0]
        per_image, sums, dlevels, dreg = loss_stage_levels(
            spec, list(levels), reg, ann, assign, num_pos, alpha=cfg["alpha"], beta=cfg["beta"],
            target_div=cfg["target_div"], yx_targets=cfg["yx_targets"], mean_divisor=cfg["divisor"](B, world),
            want_grad=want_grad, dreg_zeroed=dreg_zeroed, best=best, from_logits=cfg.get("from_logits", False),
            peer_reduce=peer.next() if peer is not None else None, spare_sms=cfg.get("spare_sms", 0))
        if world > 1 and peer is None:
            import torch.distributed as dist
            dist.all_reduce(sums, op=dist.ReduceOp.SUM, group=group)
        out = sums[:2] / sums[3]
        if want_grad:
            ctx.save_for_backward(dreg, *dlevels)
        ctx.has_grad = want_grad
        ctx.device = reg.device
        ctx.consumed = False
        ctx.set_materialize_grads(False)
        ctx.n_levels = len(levels)
        ctx.mark_non_differentiable(num_pos, per_image, sums)
        return out[0:1], out[1:2], num_pos, per_image, sums

    @staticmethod
    def backward(ctx, g_cls, g_reg, *_):
        if not ctx.has_grad:
            return (None,) * (4 + ctx.n_levels)
        _once(ctx)
        dreg, *dlevels = ctx.saved_tensors
        pairs = []
        if g_cls is not None:
            gc = g_cls.reshape(-1).float().contiguous()
            pairs += [(t, gc) for t in dlevels]
        else:
            dlevels = [None] * len(dlevels)
        if g_reg is not None:
            pairs.append((dreg, g_reg.reshape(-1).float().contiguous()))
        else:
            dreg = None
        _scale_grads(pairs, ctx.device)
        return (dreg, None, None, None, *dlevels)


class PendingLoss:
    """What FocalLoss.start() returns: the loss stage has run (gradients are ready, this rank's sums are posted to the
    peers); the GLOBAL sums are being collected on another stream.

      backward(scale=1.0)   feed d(loss)/d(classifications|regressions) to autograd (those inputs that require grad)
      losses()              (classification_loss[1], regression_loss[1]) of the global batch: makes the CURRENT stream
                            wait for the collection - call it where the value is needed (logging, the end of the step)"""

    def __init__(self, inputs, grads, sums, event, B_global_fn, module):
        self._inputs, self._grads, self._sums, self._event, self._module = inputs, grads, sums, event, module
        self.num_positives = module.last_num_positives
        self.per_image = module.last_per_image

    def backward(self, scale: float = 1.0):
        ts, gs = [], []
        grads, self._grads = self._grads, (None, None)
        for t, g in zip(self._inputs, grads):
            if t is None or g is None or not t.requires_grad:
                continue
            if scale != 1.0:
                g = g * scale
            if t.is_leaf and t.grad is None and g.shape == t.shape:
                t.grad = g                     # a leaf: the buffer itself (autograd would copy it: 8 C bytes per anchor)
            else:
                ts.append(t)
                gs.append(g)
        if ts:
            torch.autograd.backward(ts, gs)

    def losses(self):
        if self._event is not None:
            torch.cuda.current_stream(self._sums.device).wait_event(self._event)
            self._event = None
        m = self._sums[:2] / self._sums[3]
        self._module.last_sums = self._sums
        return m[0:1], m[1:2]


HeadOutputs = namedtuple("HeadOutputs", "classification regression anchors")
_CAPTURE = threading.local()      # ids of the FocalLoss modules in capture mode, per thread (see FocalLoss.capture_head)


def _capturing() -> set:
    s = getattr(_CAPTURE, "ids", None)
    if s is None:
        s = _CAPTURE.ids = set()
    return s


class FocalLoss(nn.Module):
    """Same call signature as the reference.  Constructor keywords expose the constants the
    reference hard-codes; defaults are architectures/losses.py's (0.35/0.45, alpha .25,
    smooth-L1 beta 1/9, targets / [.1,.1,.2,.2], xyxy anchors).  ``FocalLoss.efficientdet()``
    gives architectures/loss_efficientdet.py's (0.4/0.5, yx anchors, unscaled (dy,dx,dh,dw)).
    ``sharded=True`` / ``process_group=``: one process per GPU, dim 0 is this rank's shard; the four batch
    sums are exchanged inside k7_loss over NVLink peer memory (``peer_reduce``), gradients need no collective.
    ``spare_sms=n``: the streaming kernel leaves n SMs free (K7LossParams.spare_sms) for SM-bound work the caller
    overlaps on another stream - the NMS kernels of a ``detect`` on few images; up to ~16 cost nothing.
    ``from_logits=True``: `classifications` are the head's PRE-sigmoid logits (drop
    ``output_act`` of retinanet.py:140 / ``header_act`` of simple_head.py:310, see
    ``b200det.strip_head_sigmoid``); the sigmoid is fused into the streaming kernel and the gradient
    comes back with respect to the logits, so neither the sigmoid pass nor its backward touch HBM."""

    def __init__(self, neg_thr: float = 0.35, pos_thr: float = 0.45, alpha: float = 0.25,
                 gamma: float = 2.0, beta: float = 1.0 / 9.0, target_std=(0.1, 0.1, 0.2, 0.2),
                 box_order: str = "xyxy", yx_targets: bool = False, process_group=None,
                 sharded: bool = False, from_logits: bool = False, peer_reduce="auto", ddp_average: bool = False,
                 spare_sms: int = 0):
        super().__init__()
        if gamma != 2.0:
            raise ValueError("only gamma == 2 (the reference's constant, losses.py:32) is implemented")
        if box_order not in ("xyxy", "yxyx"):
            raise ValueError("box_order must be 'xyxy' or 'yxyx'")
        self.cfg = dict(neg_thr=float(neg_thr), pos_thr=float(pos_thr), alpha=float(alpha), beta=float(beta),
                        target_div=tuple(float(x) for x in (target_std if target_std is not None else (1, 1, 1, 1))),
                        box_order=L.BOX_XYXY if box_order == "xyxy" else L.BOX_YXYX,
                        yx_targets=bool(yx_targets), process_group=process_group, sharded=bool(sharded),
                        from_logits=bool(from_logits), spare_sms=int(spare_sms))
        # What the gradients are divided by.  The reference's nn.DataParallel takes the mean of the GLOBAL
        # batch (train_ksevendet.py:358-361), so a rank's gradient is d(global mean)/d(its shard) =
        # .../(B_local * world) and the ranks' parameter gradients must be SUMMED.  ddp_average=True is for
        # torch DistributedDataParallel, which AVERAGES them: each rank then scales by its own B_local only
        # and the average over ranks is again the gradient of the global mean.  Both need equal shards
        # (check_equal_shards()).
        self.ddp_average = bool(ddp_average)
        self.cfg["divisor"] = (lambda B, world: B) if self.ddp_average else (lambda B, world: B * world)
        # sharded: the exchange of the four batch sums.  "auto"/True: fused into k7_loss over peer-mapped
        # mailboxes (distributed.PeerMailbox), set up at the first call; "auto" keeps NCCL's all_reduce when
        # the mailboxes cannot be mapped (True raises); False: always all_reduce.
        self.peer_reduce = peer_reduce
        self._peer = None
        self.cfg["peer"] = self._peer_mailbox if peer_reduce in ("auto", True) else None
        self.last_num_positives = None      # [B] int32 device tensor of the last call (for logging)
        self.last_per_image = None          # [B,2] (cls_i, reg_i)
        self.last_sums = None               # [4] {sum cls_i, sum reg_i, sum num_pos_i, images}: global when sharded
        self._last_local = (0, 1)           # (B_local, world) of the last call

    @contextlib.contextmanager
    def capture_head(self):
        """While active (in the calling thread only) forward() computes nothing and returns the
        HeadOutputs it was called with.  utils.patch_model uses it to take (classification,
        regression, anchors) out of a reference forward at the one call every detector makes after
        its conv heads (retinanet.py:294, efficientdet.py:545) - EfficientDet.forward has no
        return_head switch."""
        live = _capturing()
        live.add(id(self))
        try:
            yield self
        finally:
            live.discard(id(self))

    def _peer_mailbox(self):
        if self._peer is None:
            from .distributed import PeerMailbox
            try:
                self._peer = PeerMailbox(self.cfg.get("process_group"))
            except RuntimeError:
                if self.peer_reduce is True:
                    raise
                self._peer = False           # keep the NCCL all-reduce
        return self._peer or None

    def check_equal_shards(self):
        """Sharded calls pre-scale the gradients with B_local * world, which is the global batch only if every
        rank holds the same number of images (DataParallel's scatter, utils/utils.py:184, makes them equal).
        Reads the exchanged image count of the last call (one host sync) and raises if the shards differed."""
        if self.last_sums is None:
            return
        B, world = self._last_local
        total = float(self.last_sums[3].item())
        if total != float(B * world):
            raise RuntimeError(f"FocalLoss(sharded=True): this rank holds {B} images but the {world} ranks hold {total:g} "
                               "together; shards must be equal (gradients are scaled with B_local * world)")

    @property
    def exchange(self) -> str:
        """How the batch sums of the last sharded call were combined."""
        return "p2p-fused" if self._peer else ("nccl-allreduce" if self._peer is False or not self.cfg["peer"] else "unset")

    @classmethod
    def efficientdet(cls, **kw):
        return cls(neg_thr=0.4, pos_thr=0.5, target_std=None, box_order="yxyx", yx_targets=True, **kw)

    def start(self, classifications, regressions, anchors, annotations) -> PendingLoss:
        """Two-phase form of forward() for sharded training (one process per GPU): the loss stage runs and posts this
        rank's four sums to the peers, but NOTHING on the calling stream waits for a peer - the gradients do not need
        the global sums (they are pre-scaled with the statically known global batch, losses.py:168), only the two
        reported scalars do, and those are collected on another stream (K7PeerReduce.post_only + k7_peer_collect).
        Returns a PendingLoss: ``.backward()`` now, ``.losses()`` where the values are wanted.  The next start() waits
        for this call's collection (the mailboxes hold two calls).  Single GPU / no peer mailboxes: same interface,
        nothing deferred.  Flat [B,A,C] classifications only."""
        if isinstance(classifications, (list, tuple)):
            raise ValueError("FocalLoss.start: per-level inputs are not supported, use forward()")
        L.require_cuda(classifications, regressions, annotations)
        cls_in, reg_in = classifications, regressions
        cls = L.as_f32_contig(classifications.detach())
        reg = L.as_f32_contig(regressions.detach())
        ann = L.as_f32_contig(annotations)
        B, A = cls.shape[0], cls.shape[1]
        sp = spec_of(anchors)
        anchor_src = sp if sp is not None else anchors
        cfg = self.cfg
        want_grad = cls_in.requires_grad or reg_in.requires_grad
        dreg0 = torch.empty_like(reg) if want_grad else None
        assign, best, num_pos, dreg_zeroed = _assignment(anchor_src, ann, A, cfg, dreg0)
        world, peer, group = _exchange(cfg)
        dev = cls.device
        cur = torch.cuda.current_stream(dev)
        pr = None
        if peer is not None:
            if peer.last_collect is not None and peer.last_collect[1] == bool(torch.cuda.is_current_stream_capturing()):
                cur.wait_event(peer.last_collect[0])       # two slot sets: collect k before post k+2
            pr = peer.post_only()
        per_image, sums, dcls, dreg = loss_stage(
            anchor_src, cls, reg, ann, assign, num_pos, alpha=cfg["alpha"], beta=cfg["beta"], target_div=cfg["target_div"],
            yx_targets=cfg["yx_targets"], mean_divisor=cfg["divisor"](B, world), want_grad=want_grad,
            box_order=cfg["box_order"], dreg_zeroed=dreg_zeroed, best=best, from_logits=cfg.get("from_logits", False),
            peer_reduce=pr, spare_sms=cfg.get("spare_sms", 0))
        event = None
        if peer is not None:
            posted = torch.cuda.Event()
            posted.record(cur)
            sums = peer.collect(posted)
            event = peer.last_collect[0]
        elif world > 1:
            import torch.distributed as dist
            dist.all_reduce(sums, op=dist.ReduceOp.SUM, group=group)
        self._remember(num_pos, per_image, sums, B)
        return PendingLoss((cls_in, reg_in), (dcls, dreg), sums, event, None, self)

    def forward(self, classifications, regressions, anchors, annotations=None, **kwargs):
        if id(self) in _capturing():
            return HeadOutputs(classifications, regressions, anchors)
        if annotations is None:
            raise ValueError("FocalLoss needs annotations [B,M,5] (a model called with return_loss=True and no "
                             "annotations, e.g. model(img) on architectures/retinanet.py whose default is return_loss=True)")
        if isinstance(classifications, (list, tuple)):
            return self._forward_levels(list(classifications), regressions, anchors, annotations)
        L.require_cuda(classifications, regressions, annotations)
        if classifications.dim() != 3 or regressions.dim() != 3 or regressions.shape[2] != 4:
            raise ValueError("classifications must be [B,A,C] and regressions [B,A,4]")
        if annotations.dim() != 3 or annotations.shape[2] != 5:
            raise ValueError("annotations must be [B,M,5]")
        A = classifications.shape[1]
        sp = spec_of(anchors)
        n_anchor = sp.total if sp is not None else anchors.reshape(-1, 4).shape[0]
        if n_anchor != A or regressions.shape[1] != A:
            raise ValueError(f"anchor count mismatch: anchors {n_anchor}, classifications {A}, "
                             f"regressions {regressions.shape[1]}")
        cls = L.as_f32_contig(classifications)
        reg = L.as_f32_contig(regressions)
        ann = L.as_f32_contig(annotations)
        # pass the recipe itself when there is one (a plain ctypes struct, not a tensor)
        anchor_src = sp if sp is not None else anchors
        lc, lr, num_pos, per_image, sums = _FocalLossFn.apply(cls, reg, anchor_src, ann, self.cfg)
        self._remember(num_pos, per_image, sums, cls.shape[0])
        return lc, lr

    def _remember(self, num_pos, per_image, sums, B):
        self.last_num_positives = num_pos
        self.last_per_image = per_image
        self.last_sums = sums
        world = 1
        if self.cfg.get("process_group") is not None or self.cfg.get("sharded"):
            import torch.distributed as dist
            world = dist.get_world_size(self.cfg.get("process_group"))
        self._last_local = (int(B), world)

    def _forward_levels(self, levels, regressions, anchors, annotations):
        """`classifications` given as the head's raw per-level outputs [B, anchors*classes, H_l, W_l]
        (what the last convolution writes, before permute/view/cat - SURVEY 8f row 1)."""
        L.require_cuda(regressions, annotations)
        sp = spec_of(anchors)
        B, _ = L.check_levels(levels, sp)
        if regressions.dim() != 3 or regressions.shape[0] != B or regressions.shape[1] != sp.total or regressions.shape[2] != 4:
            raise ValueError("regressions must be [B,A,4] with A matching the anchor recipe")
        if annotations.dim() != 3 or annotations.shape[2] != 5:
            raise ValueError("annotations must be [B,M,5]")
        reg = L.as_f32_contig(regressions)
        ann = L.as_f32_contig(annotations)
        lc, lr, num_pos, per_image, sums = _FocalLossLevelsFn.apply(reg, sp, ann, self.cfg, *levels)
        self._remember(num_pos, per_image, sums, B)
        return lc, lr
