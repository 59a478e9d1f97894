"""b200det.patch_model on stand-ins for the detectors `north_star` names.

The stubs repeat the reference forwards in SIGNATURE and CONTROL FLOW (conv-free: the heads
return seeded tensors):
    RetinaStub    architectures/retinanet.py:272-318      forward(img_batch, annotations=None, return_head=False, return_loss=True)
    KSevenStub    ksevendet/architecture/ksevendet.py:222-301   the same with return_loss=False, + the convert_onnx exit
    EffDetStub    architectures/efficientdet.py:528-576   forward(images, annotations=None, return_loss=True)   (no return_head)
Their stock modules are the oracle restatements of architectures/{anchors,losses,utils}.py, so
"stock vs patched" is the reference's own tail (torch ops + torchvision.ops.nms) against the fused one.
CPU tests cover the wrapper (signatures, defaults, the capture seam, DataParallel-style replication);
the GPU tests run stock vs patched end to end.
"""
import copy
import inspect

import numpy as np
import pytest
import torch
import torch.nn as nn
from torchvision.ops import nms

import b200det
from b200det import utils as BU
from oracle import restate as R

H, W, C, M = 96, 128, 5, 6


class StockAnchors(nn.Module):                       # architectures/anchors.py:9-60
    def __init__(self, pyramid_levels=None, ratios=None):
        super().__init__()
        self.pyramid_levels = [3, 4, 5, 6, 7] if pyramid_levels is None else pyramid_levels
        self.strides = [2 ** x for x in self.pyramid_levels]
        self.sizes = [2 ** (x + 2) for x in self.pyramid_levels]
        self.ratios = np.array([0.5, 1, 2]) if ratios is None else np.array(ratios)
        self.scales = np.array([2 ** 0, 2 ** (1.0 / 3.0), 2 ** (2.0 / 3.0)])

    def forward(self, image):
        a = R.retina_anchors(int(image.shape[2]), int(image.shape[3]), pyramid_levels=self.pyramid_levels,
                             ratios=self.ratios)
        return torch.from_numpy(a)[None].to(image.device)


class StockFocalLoss(nn.Module):                     # architectures/losses.py:25-168 (runs on the host)
    def forward(self, classifications, regressions, anchors, annotations):
        out = R.focal_loss(classifications.cpu(), regressions.cpu(), anchors.cpu(), annotations.cpu())
        return out.cls_mean, out.reg_mean


class StockBBoxTransform(nn.Module):                 # architectures/utils.py:6-53
    def forward(self, boxes, deltas):
        return R.decode_boxes(boxes.cpu(), deltas.cpu()).to(deltas.device)


class StockClipBoxes(nn.Module):                     # architectures/utils.py:56-71
    def forward(self, boxes, img):
        return R.clip_boxes_(boxes, int(img.shape[2]), int(img.shape[3]))


def _eval_tail(classification, transformed_anchors):
    """retinanet.py:300-318 (= efficientdet.py:558-576, ksevendet.py:321-339), op for op."""
    scores = torch.max(classification, dim=2, keepdim=True)[0]
    scores_over_thresh = (scores > 0.05)[0, :, 0]
    if scores_over_thresh.sum() == 0:
        dev = classification.device                  # the reference says .cuda(0)
        return [torch.zeros([1]).to(dev), torch.zeros([1]).to(dev), torch.zeros([1, 4]).to(dev)]
    classification = classification[:, scores_over_thresh, :]
    transformed_anchors = transformed_anchors[:, scores_over_thresh, :]
    scores = scores[:, scores_over_thresh, :]
    anchors_nms_idx = nms(transformed_anchors[0, :, :], scores[0, :, 0], 0.5)
    nms_scores, nms_class = classification[0, anchors_nms_idx, :].max(dim=1)
    return [nms_scores, nms_class, transformed_anchors[0, anchors_nms_idx, :]]


class _StubBase(nn.Module):
    def __init__(self, ratios=None):
        super().__init__()
        self.scale = nn.Parameter(torch.ones(()))    # something for autograd / state_dict to hold
        self.anchors = StockAnchors(ratios=ratios)
        self.regressBoxes = StockBBoxTransform()
        self.clipBoxes = StockClipBoxes()
        self.focalLoss = StockFocalLoss()
        self.cls_out = None                          # seeded "head outputs", set by the test
        self.reg_out = None

    def heads(self, img_batch):
        B = img_batch.shape[0]
        return self.cls_out[:B] * self.scale, self.reg_out[:B] * self.scale


class RetinaStub(_StubBase):
    def forward(self, img_batch, annotations=None, return_head=False, return_loss=True):
        classification, regression = self.heads(img_batch)
        if return_head:
            return [classification, regression]
        anchors = self.anchors(img_batch)
        if return_loss:
            return self.focalLoss(classification, regression, anchors, annotations)
        else:
            transformed_anchors = self.regressBoxes(anchors, regression)
            transformed_anchors = self.clipBoxes(transformed_anchors, img_batch)
            return _eval_tail(classification, transformed_anchors)


class KSevenStub(_StubBase):
    def __init__(self):
        super().__init__(ratios=[1, 2])              # config/model_3s-thermal-face: ratios [1, 2]
        self.convert_onnx = False

    def forward(self, img_batch, annotations=None, return_head=False, return_loss=False):
        classification, regression = self.heads(img_batch)
        if return_head:
            return [classification, regression]
        anchors = self.anchors(img_batch)
        if return_loss:
            return self.focalLoss(classification, regression, anchors, annotations)
        else:
            if self.convert_onnx:
                return [classification, regression]
            transformed_anchors = self.regressBoxes(anchors, regression)
            transformed_anchors = self.clipBoxes(transformed_anchors, img_batch)
            return _eval_tail(classification, transformed_anchors)


class EffDetStub(_StubBase):
    def forward(self, images, annotations=None, return_loss=True):
        classification, regression = self.heads(images)
        anchors = self.anchors(images)
        if return_loss:
            return self.focalLoss(classification, regression, anchors, annotations)
        else:
            transformed_anchors = self.regressBoxes(anchors, regression)
            transformed_anchors = self.clipBoxes(transformed_anchors, images)
            return _eval_tail(classification, transformed_anchors)


STUBS = [RetinaStub, KSevenStub, EffDetStub]


def seeded(model, B=2, device="cpu", seed=3, logit_mean=-3.0):
    gen = torch.Generator().manual_seed(seed)
    A = model.anchors(torch.empty(1, 3, H, W)).shape[1]
    cls, reg = b200det.synth.make_head_outputs(B, A, C, gen, logit_mean=logit_mean, logit_std=1.5)
    ann = b200det.synth.make_annotations(B, M, C, H, W, gen, gt_min=8.0, gt_max=60.0)
    model.cls_out, model.reg_out = cls.to(device), reg.to(device)
    return ann.to(device)


# --------------------------------------------------------------------------------------------- CPU: the wrapper
@pytest.mark.parametrize("cls", STUBS)
def test_patch_swaps_modules_and_keeps_the_forward_signature(cls):
    model = cls()
    before = inspect.signature(cls.forward)
    keys = set(model.state_dict())
    assert b200det.patch_model(model) is model
    assert isinstance(model, cls) and type(model).__name__ == cls.__name__
    assert isinstance(model.anchors, b200det.Anchors) and isinstance(model.focalLoss, b200det.FocalLoss)
    assert isinstance(model.regressBoxes, b200det.BBoxTransform) and isinstance(model.clipBoxes, b200det.ClipBoxes)
    assert inspect.signature(type(model).forward) == before
    assert set(model.state_dict()) == keys                      # the swapped modules hold no parameters / buffers
    assert list(model.anchors.ratios) == list(cls().anchors.ratios)
    b200det.patch_model(model)                                   # patching twice does not stack wrappers
    assert type(model).__mro__[1] is cls


@pytest.mark.parametrize("cls", STUBS)
def test_eval_call_reaches_the_fused_tail_with_the_head_outputs(cls, monkeypatch):
    """model(img) / model(img, return_loss=False) with the model's own defaults: the reference forward runs up
    to its focalLoss call, the fused tail gets (classification, regression, anchors).  No TypeError on
    EfficientDet (no return_head), no FocalLoss call for KSevenDet's return_loss=False default."""
    model = cls()
    seeded(model)
    b200det.patch_model(model)
    stock_anchors = StockAnchors(ratios=model.anchors.ratios)
    model.anchors.forward = stock_anchors.forward               # CPU: b200det.Anchors itself needs a GPU
    seen = {}

    def fake_tail(classification, regression, anchors, img_batch, regressBoxes=None, score_thr=0.05, iou_thr=0.5,
                  from_logits=False):
        seen.update(cls=classification, reg=regression, anc=anchors, img=img_batch, rb=regressBoxes,
                    thr=(score_thr, iou_thr))
        return ["scores", "labels", "boxes"]

    monkeypatch.setattr(BU, "eval_tail", fake_tail)
    img = torch.empty(2, 3, H, W)
    out = model(img, return_loss=False)
    assert out == ["scores", "labels", "boxes"]
    assert torch.equal(seen["cls"], model.cls_out) and torch.equal(seen["reg"], model.reg_out)
    assert torch.equal(seen["anc"], stock_anchors(img)) and seen["img"] is img
    assert seen["rb"] is model.regressBoxes and seen["thr"] == (0.05, 0.5)
    if cls is KSevenStub:                                        # ksevendet.py:222: return_loss defaults to False
        seen.clear()
        assert model(img) == ["scores", "labels", "boxes"] and seen
        model.convert_onnx = True                                # ksevendet.py:276-278: the ONNX exit is the model's own
        out = model(img)
        assert len(out) == 2 and torch.equal(out[0], model.cls_out)
    else:                                                        # retinanet.py:272 / efficientdet.py:528: default is the loss
        with pytest.raises((ValueError, RuntimeError)):          # FocalLoss without annotations / without a GPU, never TypeError
            model(img)
    if cls is not EffDetStub:
        head = model(img, return_head=True)
        assert torch.equal(head[0], model.cls_out) and torch.equal(head[1], model.reg_out)
    from b200det import losses as BL
    assert id(model.focalLoss) not in BL._capturing()            # the seam is closed again


@pytest.mark.parametrize("cls", STUBS)
def test_training_call_goes_to_focal_loss_and_replicas_keep_the_override(cls):
    model = cls()
    ann = seeded(model)
    b200det.patch_model(model)
    model.anchors.forward = StockAnchors(ratios=model.anchors.ratios).forward
    img = torch.empty(2, 3, H, W)
    with pytest.raises(RuntimeError, match="CUDA"):             # b200det.FocalLoss: no CPU path, and it was reached
        model(img, ann, return_loss=True)
    # nn.DataParallel replicates with __new__(type(module)) + a __dict__ copy (torch/nn/modules/module.py
    # _replicate_for_data_parallel): the override must live on the class, not on the instance
    replica = model._replicate_for_data_parallel()
    assert type(replica) is type(model) and "forward" not in replica.__dict__
    assert type(copy.deepcopy(model)) is type(model)


def test_capture_is_thread_local():
    import threading
    fl = b200det.FocalLoss()
    got = {}

    def other():
        try:
            fl(torch.zeros(1, 4, 2), torch.zeros(1, 4, 4), torch.zeros(1, 4, 4), torch.zeros(1, 1, 5))
        except RuntimeError as e:                               # the normal path: needs CUDA
            got["err"] = str(e)

    with fl.capture_head():
        t = threading.Thread(target=other)
        t.start()
        t.join()
        cap = fl("c", "r", "a", None)
    assert "CUDA" in got["err"] and cap.classification == "c" and cap.anchors == "a"


# --------------------------------------------------------------------------------------------- GPU: stock vs patched
def _gpu_models(cls, seed, logit_mean=-3.0, B=2):
    stock = cls().cuda()
    ann = seeded(stock, B=B, device="cuda", seed=seed, logit_mean=logit_mean)
    patched = cls().cuda()
    patched.cls_out, patched.reg_out = stock.cls_out, stock.reg_out
    b200det.patch_model(patched)
    return stock, patched, ann


@pytest.mark.gpu
@pytest.mark.parametrize("cls", STUBS)
def test_patched_eval_equals_the_reference_tail(cls):
    stock, patched, _ = _gpu_models(cls, seed=11)
    img = torch.empty(2, 3, H, W, device="cuda")
    want = stock(img, return_loss=False)
    got = patched(img, return_loss=False)
    assert len(got) == 3 and want[0].numel() > 4                 # a real NMS problem, not the empty triple
    assert torch.equal(got[0], want[0])                          # scores are copies of the class maxima
    assert got[1].dtype == torch.int64 and torch.equal(got[1], want[1])
    np.testing.assert_allclose(got[2].detach().cpu().numpy(), want[2].detach().cpu().numpy(), rtol=1e-5, atol=1e-4)
    if cls is KSevenStub:
        again = patched(img)                                     # its default IS the eval branch
        assert torch.equal(again[0], want[0]) and torch.equal(again[1], want[1])
    # the oracle's own detect (image 0) agrees on the kept anchors
    _, dets = R.detect_batch(stock.cls_out.cpu(), stock.reg_out.cpu(), stock.anchors(img).cpu(), H, W)
    assert torch.equal(got[1].cpu(), dets[0].labels)


@pytest.mark.gpu
@pytest.mark.parametrize("cls", STUBS)
def test_patched_eval_empty_triple(cls):
    stock, patched, _ = _gpu_models(cls, seed=12, logit_mean=-9.0)
    img = torch.empty(1, 3, H, W, device="cuda")
    want = stock(img, return_loss=False)
    got = patched(img, return_loss=False)
    assert [tuple(t.shape) for t in got] == [(1,), (1,), (1, 4)] == [tuple(t.shape) for t in want]
    assert all(t.is_cuda and float(t.abs().sum()) == 0.0 for t in got)


@pytest.mark.gpu
@pytest.mark.parametrize("cls", STUBS)
def test_patched_training_equals_the_reference_loss(cls):
    stock, patched, ann = _gpu_models(cls, seed=13)
    img = torch.empty(2, 3, H, W, device="cuda")
    wc, wr = stock(img, ann, return_loss=True)
    gc, gr = patched(img, ann, return_loss=True)
    np.testing.assert_allclose(gc.item(), wc.item(), rtol=1e-5)
    np.testing.assert_allclose(gr.item(), wr.item(), rtol=1e-5)
    (gc + gr).sum().backward()                                   # gradients flow back to the model's parameters
    (wc + wr).sum().backward()
    np.testing.assert_allclose(patched.scale.grad.item(), stock.scale.grad.item(), rtol=2e-5)


@pytest.mark.gpu
def test_patched_fuse_sigmoid_keeps_results():
    """patch_model(fuse_sigmoid=True): the head's Sigmoid becomes the identity and both branches run on logits."""
    class WithAct(RetinaStub):
        def __init__(self):
            super().__init__()
            self.output_act = nn.Sigmoid()                       # retinanet.py:140,159

        def heads(self, img_batch):
            B = img_batch.shape[0]
            return self.output_act(self.cls_out[:B]), self.reg_out[:B]

    stock = WithAct().cuda()
    ann = seeded(stock, device="cuda", seed=14)
    stock.cls_out = torch.logit(stock.cls_out)                   # the conv output: logits
    patched = WithAct().cuda()
    patched.cls_out, patched.reg_out = stock.cls_out, stock.reg_out
    b200det.patch_model(patched, fuse_sigmoid=True)
    assert isinstance(patched.output_act, nn.Identity)
    img = torch.empty(2, 3, H, W, device="cuda")
    want, got = stock(img, return_loss=False), patched(img, return_loss=False)
    assert torch.equal(got[0], want[0]) and torch.equal(got[1], want[1])
    wc, wr = stock(img, ann)
    gc, gr = patched(img, ann)
    np.testing.assert_allclose([gc.item(), gr.item()], [wc.item(), wr.item()], rtol=1e-5)
