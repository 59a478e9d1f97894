"""Generate golden fixtures by running the UNMODIFIED reference on seeded inputs.

Run in the build container only (needs /root/reference, which does not exist on
the GPU box):

    python tests/golden/make_golden.py

It imports the reference's own modules
  architectures.anchors.Anchors, ksevendet.architecture.anchors.Anchors,
  architectures.losses.{calc_iou,FocalLoss}, architectures.loss_efficientdet.{calc_iou,FocalLoss},
  architectures.utils.{BBoxTransform,ClipBoxes}, architectures.utils_efficientdet.{Anchors,BBoxTransform,ClipBoxes},
  utils.utils.postprocess (batched eval tail) and torchvision.ops.nms
and writes
  tests/golden/anchors.json        sha256 + sampled rows of Anchors()(image) per recipe
  tests/golden/loss_*.npz          inputs, reference losses, autograd grads, IoU max/argmax/state
  tests/golden/detect_*.npz        inputs, decoded+clipped boxes, per-image keep / scores / labels / boxes
  tests/golden/head_retina.npz     per-level NCHW logits of the reference's ClassificationModel (captured in
                                   front of its output_act) and the [B,A,C] probabilities it returns
  tests/golden/collate.npz         per-image annotation rows and the padded block the reference's collater makes of them
  tests/golden/batched_nms.npz     boxes / scores / categories and what torchvision 0.26.0 batched_nms (both
                                   strategies) keeps - SURVEY 8f row 4; the reference has no per-class NMS
  tests/golden/eval_tail.npz       detections of 6 images and what datasettool/{coco_eval,ksevendata_eval,csv_eval}.py
                                   make of them (result rows / per-class detection arrays)
The reference's hidden intermediates (IoU_max, IoU_argmax, masks) are obtained by
calling the reference's own calc_iou and repeating losses.py:66,79-82 on its output.
"""
import hashlib
import io
import json
import os
import sys
from contextlib import redirect_stdout

import numpy as np
import torch

REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, REF)
sys.path.insert(0, ROOT)

from architectures import anchors as ref_anchors            # noqa: E402
from architectures import losses as ref_losses              # noqa: E402
from architectures import loss_efficientdet as ref_losses_yx  # noqa: E402
from architectures import utils as ref_utils                # noqa: E402
from architectures import utils_efficientdet as ref_utils_yx  # noqa: E402
from ksevendet.architecture import anchors as ref_k7_anchors  # noqa: E402
from utils.utils import postprocess as ref_postprocess      # noqa: E402
from torchvision.ops import nms as tv_nms                   # noqa: E402

import b200det.synth as synth                               # noqa: E402

torch.set_num_threads(4)


def sha(a: np.ndarray) -> str:
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


# ------------------------------------------------------------------ anchors
ANCHOR_RECIPES = {
    # name: (module, ctor kwargs, (H, W))
    "default_512": ("retina", {}, (512, 512)),
    "default_896": ("retina", {}, (896, 896)),
    "default_1536": ("retina", {}, (1536, 1536)),
    "default_odd_500x333": ("retina", {}, (500, 333)),
    "default_odd_97x211": ("retina", {}, (97, 211)),
    "effdet_coded_512": ("retina", dict(pyramid_levels=[3, 4, 5, 6, 7], strides=[8, 16, 32, 64, 128],
                                        sizes=[2 ** (x + 1) * 1.25 for x in [3, 4, 5, 6, 7]],
                                        ratios=[1, 2], scales=[2 ** 0, 2 ** (1.0 / 3.0), 2 ** (2.0 / 3.0)]), (512, 512)),
    "p45p6_640x512": ("retina", dict(pyramid_levels=[4, 5, 6], strides=[16, 32, 96], sizes=[32, 64, 160],
                                     ratios=[1, 1.5, 2]), (512, 640)),
    "k7_thermal_1.5_512x640": ("k7", dict(pyramid_levels=[3, 4, 5, 6, 7],
                                          sizes_mapper='lambda level: 2**(level + 1) * 1.5', ratios=[1, 2],
                                          scales='[2 ** 0, 2 ** (1.0 / 3.0), 2 ** (2.0 / 3.0)]'), (512, 640)),
    "k7_thermal_1.25_512x640": ("k7", dict(pyramid_levels=[3, 4, 5, 6, 7],
                                           sizes_mapper='lambda level: 2**(level + 1) * 1.25', ratios=[1, 2],
                                           scales='[2 ** 0, 2 ** (1.0 / 3.0), 2 ** (2.0 / 3.0)]'), (512, 640)),
    "k7_thermal_2_480x640": ("k7", dict(pyramid_levels=[3, 4, 5, 6, 7],
                                        sizes_mapper='lambda level: 2**(level + 2)', ratios=[0.5, 1, 2],
                                        scales='[2 ** 0, 2 ** (1.0 / 3.0), 2 ** (2.0 / 3.0)]'), (480, 640)),
    "yx_default_512": ("yx", {}, (512, 512)),
    "yx_default_384x640": ("yx", {}, (384, 640)),
}


def ref_anchor_tensor(kind, kwargs, hw):
    img = torch.zeros(1, 3, hw[0], hw[1])
    if kind == "retina":
        mod = ref_anchors.Anchors(**kwargs)
    elif kind == "k7":
        mod = ref_k7_anchors.Anchors(**kwargs)
    else:
        mod = ref_utils_yx.Anchors(**kwargs)
    return mod(img)


def gen_anchors():
    out = {}
    for name, (kind, kwargs, hw) in ANCHOR_RECIPES.items():
        a = ref_anchor_tensor(kind, kwargs, hw).numpy()
        assert a.dtype == np.float32 and a.shape[0] == 1
        idx = np.linspace(0, a.shape[1] - 1, 7).astype(np.int64)
        out[name] = dict(kind=kind, kwargs={k: (v if isinstance(v, str) else list(map(float, v)))
                                            for k, v in kwargs.items()},
                         height=hw[0], width=hw[1], count=int(a.shape[1]), sha256=sha(a[0]),
                         sample_idx=idx.tolist(), sample_rows=a[0, idx].astype(np.float64).tolist())
    with open(os.path.join(HERE, "anchors.json"), "w") as f:
        json.dump(out, f, indent=1)
    print("anchors.json:", {k: v["count"] for k, v in out.items()})


# ------------------------------------------------------------------ loss
def adversarial_annotations(ann: torch.Tensor, anchors_xyxy: torch.Tensor, H, W):
    """Edit a [B,M,5] annotation block in place to exercise the reference's edge
    cases: exact-duplicate GT rows (IoU ties -> first index), a GT equal to an
    anchor (IoU == 1), a tiny GT matching no anchor, a pad row in the MIDDLE,
    a sub-pixel GT (w,h < 1: clamp at losses.py:136-137) and an all-pad image."""
    B, M, _ = ann.shape
    if B >= 1 and M >= 4:
        ann[0, 1] = ann[0, 0]                      # duplicate -> tie, argmax must be row 0
        ann[0, 2, :4] = anchors_xyxy[anchors_xyxy.shape[0] // 3]   # exact anchor
        ann[0, 2, 4] = 0
        ann[0, 3] = torch.tensor([5.25, 7.5, 5.75, 8.0, 1.0])      # sub-pixel
    if B >= 2 and M >= 3:
        ann[1, 1] = -1.0                           # pad row in the middle
        ann[1, 2] = torch.tensor([1.0, 1.0, 3.5, 3.0, 0.0])        # matches nothing
    if B >= 3:
        ann[2] = -1.0                              # no GT at all
    return ann


def ref_intermediates(calc_iou, anchors, ann_img, neg_thr, pos_thr):
    rows = torch.nonzero(ann_img[:, 4] != -1).flatten()
    A = anchors.shape[0]
    if rows.numel() == 0:
        return (np.zeros(A, np.float32), np.full(A, -1, np.int64), np.full(A, -1, np.int64),
                np.full(A, 1, np.int8))
    gt = ann_img[rows]
    iou = calc_iou(anchors, gt[:, :4])
    iou_max, iou_arg = torch.max(iou, dim=1)                       # losses.py:66
    state = np.full(A, 1, np.int8)
    state[torch.lt(iou_max, neg_thr).numpy()] = 0                  # losses.py:79
    state[torch.ge(iou_max, pos_thr).numpy()] = 2                  # losses.py:80
    return iou_max.numpy(), iou_arg.numpy(), rows[iou_arg].numpy(), state


LOSS_CASES = {
    # name: (variant, H, W, B, C, M, anchor kwargs, seed, logit_mean)
    "retina_small": ("retina", 128, 160, 3, 8, 6, {}, 1, -3.0),
    "retina_c1": ("retina", 96, 96, 2, 1, 5, dict(ratios=[1, 2], sizes=[2 ** (x + 1) * 1.5 for x in [3, 4, 5, 6, 7]]), 2, -2.0),
    "retina_odd": ("retina", 75, 131, 4, 3, 9, {}, 3, -4.0),
    "retina_c90": ("retina", 96, 96, 2, 90, 12, {}, 4, -5.0),
    "effdet_small": ("effdet", 128, 128, 3, 5, 7, {}, 5, -3.0),
}


def gen_loss():
    for name, (variant, H, W, B, C, M, akw, seed, mu) in LOSS_CASES.items():
        img = torch.zeros(B, 3, H, W)
        if variant == "retina":
            anc = ref_anchors.Anchors(**akw)(img)
            loss_mod, calc, thr = ref_losses.FocalLoss(), ref_losses.calc_iou, (0.35, 0.45)
            anc_xyxy = anc[0]
        else:
            anc = ref_utils_yx.Anchors(**akw)(img)
            loss_mod, calc, thr = ref_losses_yx.FocalLoss(), ref_losses_yx.calc_iou, (0.4, 0.5)
            anc_xyxy = anc[0][:, [1, 0, 3, 2]]
        A = anc.shape[1]
        gen = torch.Generator().manual_seed(777 + seed)
        ann = synth.make_annotations(B, M, C, H, W, gen, mode="uniform", gt_min=8.0,
                                     gt_max=0.6 * min(H, W))
        ann = adversarial_annotations(ann, anc_xyxy, H, W)
        cls, reg = synth.make_head_outputs(B, A, C, gen, logit_mean=mu, logit_std=2.0)
        # force a few probabilities outside the clamp window (losses.py:62)
        cls.view(-1)[::97] = 1e-6
        cls.view(-1)[5::193] = 1.0 - 1e-7
        cls = cls.clone().requires_grad_(True)
        reg = reg.clone().requires_grad_(True)
        with redirect_stdout(io.StringIO()):
            lc, lr = loss_mod(cls, reg, anc, ann)
        gc_c, gr_c = torch.autograd.grad(lc.sum(), [cls, reg], retain_graph=True, allow_unused=True)
        gc_r, gr_r = torch.autograd.grad(lr.sum(), [cls, reg], allow_unused=True)
        z = lambda g, like: (torch.zeros_like(like) if g is None else g)
        inter = [ref_intermediates(calc, anc[0], ann[b], *thr) for b in range(B)]
        np.savez_compressed(
            os.path.join(HERE, f"loss_{name}.npz"),
            variant=variant, height=H, width=W, anchor_kwargs=json.dumps(akw),
            anchors=anc.numpy(), annotations=ann.numpy(),
            classifications=cls.detach().numpy(), regressions=reg.detach().numpy(),
            cls_loss=lc.detach().numpy(), reg_loss=lr.detach().numpy(),
            dcls=(z(gc_c, cls) + z(gc_r, cls)).numpy(), dreg=(z(gr_c, reg) + z(gr_r, reg)).numpy(),
            iou_max=np.stack([i[0] for i in inter]), iou_argmax=np.stack([i[1] for i in inter]),
            gt_row=np.stack([i[2] for i in inter]), state=np.stack([i[3] for i in inter]))
        print(f"loss_{name}: A={A} cls={lc.item():.6f} reg={lr.item():.6f} "
              f"pos={[int((i[3] == 2).sum()) for i in inter]}")


# ------------------------------------------------------------------ detect
DETECT_CASES = {
    # name: (variant, H, W, B, C, anchor kwargs, seed, logit_mean, score_thr, iou_thr)
    "retina_small": ("retina", 128, 160, 3, 8, {}, 11, -3.5, 0.05, 0.5),
    "retina_dense_c1": ("retina", 96, 128, 2, 1, dict(ratios=[1, 2], sizes=[2 ** (x + 1) * 1.5 for x in [3, 4, 5, 6, 7]]), 12, -1.0, 0.05, 0.5),
    "retina_c90": ("retina", 96, 96, 2, 90, {}, 13, -8.5, 0.05, 0.5),
    "retina_empty": ("retina", 64, 64, 2, 4, {}, 14, -9.0, 0.05, 0.5),
    "effdet_small": ("effdet", 128, 128, 2, 6, {}, 15, -3.5, 0.2, 0.3),
}


def gen_detect():
    for name, (variant, H, W, B, C, akw, seed, mu, sthr, ithr) in DETECT_CASES.items():
        img = torch.zeros(B, 3, H, W)
        if variant == "retina":
            anc = ref_anchors.Anchors(**akw)(img)
            rb, cb = ref_utils.BBoxTransform(), ref_utils.ClipBoxes()
        else:
            anc = ref_utils_yx.Anchors(**akw)(img)
            rb, cb = ref_utils_yx.BBoxTransform(), ref_utils_yx.ClipBoxes()
        A = anc.shape[1]
        gen = torch.Generator().manual_seed(999 + seed)
        cls, reg = synth.make_head_outputs(B, A, C, gen, logit_mean=mu, logit_std=1.5)
        if name == "retina_small":
            # exact score ties (stable sort must keep the lower anchor id first) and
            # a class-max tie inside one row (argmax = first index)
            cls[0, 100] = cls[0, 50]
            cls[0, 200] = cls[0, 50]
            cls[0, 300, :] = 0.0
            cls[0, 300, 2] = 0.7
            cls[0, 300, 5] = 0.7
            reg[0, 100] = reg[0, 50]        # identical boxes at different anchors are unlikely; keep deltas equal
        if name == "retina_empty":
            cls[1, 7, 1] = 0.9               # image 1 has exactly one survivor, image 0 none
        boxes = cb(rb(anc, reg), img).clone()
        with redirect_stdout(io.StringIO()):
            out = ref_postprocess(img, anc, reg, cls, rb, cb, sthr, ithr)
        # utils/utils.py:83-88 appends an extra empty dict per image when the WHOLE batch has
        # no survivor; strip those duplicates
        if len(out) == 2 * B:
            out = out[1::2]
        assert len(out) == B
        rec = dict(variant=variant, height=H, width=W, anchor_kwargs=json.dumps(akw),
                   score_thr=sthr, iou_thr=ithr, anchors=anc.numpy(), classifications=cls.numpy(),
                   regressions=reg.numpy(), boxes=boxes.numpy())
        counts = []
        for b in range(B):
            score = torch.max(cls[b], dim=1)[0]
            surv = torch.nonzero(score > sthr).flatten()
            keep = tv_nms(boxes[b][surv], score[surv], ithr) if surv.numel() else torch.zeros(0, dtype=torch.int64)
            rec[f"survivors_{b}"] = surv.numpy()
            rec[f"keep_rank_{b}"] = keep.numpy()
            rec[f"keep_anchor_{b}"] = surv[keep].numpy()
            rois = np.asarray(out[b]['rois'], dtype=np.float32).reshape(-1, 4)
            rec[f"boxes_{b}"] = rois
            rec[f"scores_{b}"] = np.asarray(out[b]['scores'], dtype=np.float32).reshape(-1)
            rec[f"labels_{b}"] = np.asarray(out[b]['class_ids'], dtype=np.int64).reshape(-1)
            assert rois.shape[0] == keep.numel()
            counts.append((int(surv.numel()), int(keep.numel())))
        np.savez_compressed(os.path.join(HERE, f"detect_{name}.npz"), **rec)
        print(f"detect_{name}: A={A} (survivors, kept)={counts}")


def gen_head():
    """architectures/retinanet.py:119-178: what the classification head hands to the hot path, and the
    logits it had right before `output_act` (the tensor the from_logits path consumes)."""
    from architectures.retinanet import ClassificationModel
    torch.manual_seed(20260411)
    A_cell, Cn, F = 9, 5, 8
    head = ClassificationModel(F, num_anchors=A_cell, num_classes=Cn, feature_size=8).eval()
    with torch.no_grad():                     # spread the logits (the default init gives p ~ 0.01 everywhere)
        head.output.weight.mul_(40.0)
        head.output.bias.uniform_(-4.0, 1.0)
    grabbed = []
    head.output.register_forward_hook(lambda m, i, o: grabbed.append(o.detach().clone()))
    feats = [torch.randn(2, F, 6, 5), torch.randn(2, F, 3, 3), torch.randn(2, F, 2, 1)]
    with torch.no_grad():
        probs = torch.cat([head(f) for f in feats], dim=1)          # retinanet.py:286
    rec = {"probs": probs.numpy(), "num_anchors": A_cell, "num_classes": Cn}
    for i, g in enumerate(grabbed):
        rec[f"logits_l{i}"] = g.numpy()
    np.savez_compressed(os.path.join(HERE, "head_retina.npz"), **rec)
    print("head_retina", probs.shape, float(probs.min()), float(probs.max()))

def gen_eval_tail():
    """datasettool/coco_eval.py:6-66, ksevendata_eval.py:8-72, csv_eval.py:64-121: what the evaluators do
    with the model's (scores, labels, boxes) per image.  The UNMODIFIED reference functions are run on a
    stub dataset and a stub model that replays fixed detections; pycocotools (absent here, and irrelevant to
    the tail) is replaced by an empty stand-in so that the modules import, and Tensor.cuda() is the identity
    on this GPU-less machine (ksevendata_eval.py calls it unconditionally)."""
    import tempfile
    import types
    stub = types.ModuleType("pycocotools")
    stub_eval = types.ModuleType("pycocotools.cocoeval")

    class COCOeval:                                   # nothing of the tail depends on it
        def __init__(self, *a, **k):
            self.params = types.SimpleNamespace()

        def evaluate(self): pass
        def accumulate(self): pass
        def summarize(self): pass
    stub_eval.COCOeval = COCOeval
    stub.cocoeval = stub_eval
    sys.modules.setdefault("pycocotools", stub)
    sys.modules.setdefault("pycocotools.cocoeval", stub_eval)
    from datasettool import coco_eval as ref_coco_eval
    from datasettool import ksevendata_eval as ref_k7_eval
    from datasettool import csv_eval as ref_csv_eval

    rng = np.random.RandomState(20260413)
    B, Cn, cap = 6, 7, 48
    counts = np.array([40, 0, 48, 1, 17, 33], dtype=np.int32)
    scores = np.zeros((B, cap), np.float32)
    labels = np.zeros((B, cap), np.int64)
    boxes = np.zeros((B, cap, 4), np.float32)
    for b in range(B):
        n = int(counts[b])
        sc = np.sort(rng.uniform(0.02, 0.99, size=n).astype(np.float32))[::-1]   # descending; a tail below 0.05 / 0.3
        scores[b, :n] = sc
        labels[b, :n] = rng.randint(0, Cn, size=n)
        xy = rng.uniform(0, 400, size=(n, 2)).astype(np.float32)
        wh = rng.uniform(1, 200, size=(n, 2)).astype(np.float32)
        boxes[b, :n, :2] = xy
        boxes[b, :n, 2:] = xy + wh
    scales = np.array([0.7314, 1.0, 1.6180339887, 0.5, 0.3333333333, 2.25])
    image_ids = [139, 285, 632, 724, 776, 581929]
    coco_label = {0: 1, 1: 2, 2: 3, 3: 5, 4: 7, 5: 11, 6: 90}

    class Coco:
        def loadRes(self, path): return None

    class DS:
        set_name = "golden"
        coco = Coco()
        kseven_data = Coco()

        def __init__(self): self.image_ids = image_ids
        def __len__(self): return B
        def __getitem__(self, i): return {"img": torch.zeros(4, 4, 3), "scale": float(scales[i])}
        def label_to_coco_label(self, label): return coco_label[label]
        def label_to_ksevendata_label(self, label): return coco_label[label]
        def num_classes(self): return Cn

    class Model:
        """Replays the detections: the zeros triple when an image has none (retinanet.py:308)."""
        def __init__(self): self.i = 0
        def eval(self): return self
        def train(self, mode=True): return self

        def __call__(self, x, return_loss=False):
            b = self.i % B
            self.i += 1
            n = int(counts[b])
            if n == 0:
                return [torch.zeros(1), torch.zeros(1), torch.zeros(1, 4)]
            return [torch.from_numpy(scores[b, :n].copy()), torch.from_numpy(labels[b, :n].copy()),
                    torch.from_numpy(boxes[b, :n].copy())]

    rec = {"scores": scores, "labels": labels, "boxes": boxes, "counts": counts, "scales": scales,
           "image_ids": np.array(image_ids, np.int64), "category_of_label": np.array([coco_label[i] for i in range(Cn)], np.int32)}
    had_cuda = torch.Tensor.cuda
    torch.Tensor.cuda = lambda self, *a, **k: self
    cwd = os.getcwd()
    try:
        with tempfile.TemporaryDirectory() as tmp, redirect_stdout(io.StringIO()):
            os.chdir(tmp)
            for tag, thr in (("coco_t005", 0.05), ("coco_t030", 0.3)):
                ref_coco_eval.evaluate_coco(DS(), Model(), threshold=thr)
                res = json.load(open("golden_bbox_results.json"))
                rec[tag] = np.array([[r["image_id"], *r["bbox"], r["score"], r["category_id"]] for r in res], np.float64)
            for tag, thr, valid in (("k7_t005_all", 0.05, None), ("k7_t030_valid", 0.3, [0, 2, 5])):
                try:
                    ref_k7_eval.coco_evaluate_ksevendata(DS(), Model(), threshold=thr, valid_class_id=valid)
                except Exception:          # whatever follows the dump needs a real annotation file
                    pass
                res = json.load(open("tmp_bbox_results.json"))
                rec[tag] = np.array([[r["image_id"], *r["bbox"], r["score"], r["category_id"]] for r in res], np.float64)
            for tag, thr, md in (("csv_t005_m100", 0.05, 100), ("csv_t030_m10", 0.3, 10)):
                det = ref_csv_eval._get_detections(DS(), Model(), score_threshold=thr, max_detections=md)
                rows, offs = [], [0]
                for b in range(B):
                    for c in range(Cn):
                        d = np.asarray(det[b][c], np.float64).reshape(-1, 5)
                        for r in d:
                            rows.append([b, *r, c])
                        offs.append(len(rows))
                rec[tag] = np.array(rows, np.float64).reshape(-1, 7)
                rec[tag + "_class_offsets"] = np.array(offs, np.int32)
    finally:
        os.chdir(cwd)
        torch.Tensor.cuda = had_cuda
    rec["k7_valid"] = np.array([0, 2, 5], np.int32)
    np.savez_compressed(os.path.join(HERE, "eval_tail.npz"), **rec)
    print("eval_tail", {k: v.shape for k, v in rec.items() if k.startswith(("coco", "k7_t", "csv"))})

def gen_collate():
    """datasettool/dataloader.py:655-691 (collater): the annotation block it hands to FocalLoss.  skimage and
    pycocotools are absent here and irrelevant to collater; empty stand-ins let the module import."""
    import types
    for name in ("skimage", "skimage.io", "skimage.transform", "skimage.color", "pycocotools", "pycocotools.coco"):
        sys.modules.setdefault(name, types.ModuleType(name))
    if not hasattr(sys.modules["pycocotools.coco"], "COCO"):
        sys.modules["pycocotools.coco"].COCO = object
    from datasettool.dataloader import collater
    rng = np.random.RandomState(20260414)
    rec = {}
    for tag, counts in (("mixed", [3, 0, 7, 1, 0, 5]), ("all_empty", [0, 0, 0]), ("single", [4])):
        data = []
        for i, n in enumerate(counts):
            a = np.concatenate([rng.uniform(0, 300, size=(n, 4)), rng.randint(0, 9, size=(n, 1))], axis=1).astype(np.float32)
            data.append({"img": torch.zeros(8 + i, 6 + i, 3), "annot": torch.from_numpy(a), "scale": 1.0})
            rec[f"{tag}_rows{i}"] = a
        out = collater(data)
        rec[f"{tag}_padded"] = out["annot"].numpy()
        rec[f"{tag}_n"] = np.array(len(counts))
    np.savez_compressed(os.path.join(HERE, "collate.npz"), **rec)
    print("collate", {k: v.shape for k, v in rec.items() if k.endswith("_padded")})


def gen_batched_nms():
    """torchvision.ops.batched_nms: _batched_nms_vanilla and _batched_nms_coordinate_trick on clustered boxes."""
    from torchvision.ops import boxes as tvb
    g = torch.Generator().manual_seed(20260415)
    rec = {}
    for tag, n, ncls, span in (("small", 300, 5, 200.0), ("many_classes", 2500, 90, 512.0), ("one_class", 800, 1, 300.0),
                               ("dense", 1500, 3, 120.0)):
        centre = torch.rand(n // 6 + 1, 2, generator=g) * span
        pick = torch.randint(0, centre.shape[0], (n,), generator=g)
        c = centre[pick] + torch.randn(n, 2, generator=g) * 6.0
        wh = torch.exp(torch.rand(n, 2, generator=g) * 1.5 + 2.5)
        boxes = torch.cat([c - wh / 2, c + wh / 2], dim=1).float()
        scores = ((torch.randperm(n, generator=g) + 1).float() / float(n + 1)) ** 2   # distinct by construction
        assert torch.unique(scores).numel() == n
        idxs = torch.randint(0, ncls, (n,), generator=g)
        rec[f"{tag}_boxes"] = boxes.numpy()
        rec[f"{tag}_scores"] = scores.numpy()
        rec[f"{tag}_idxs"] = idxs.numpy()
        rec[f"{tag}_vanilla"] = tvb._batched_nms_vanilla(boxes, scores, idxs, 0.5).numpy()
        rec[f"{tag}_trick"] = tvb._batched_nms_coordinate_trick(boxes, scores, idxs, 0.5).numpy()
        rec[f"{tag}_public"] = tvb.batched_nms(boxes, scores, idxs, 0.5).numpy()
        rec[f"{tag}_agnostic"] = tv_nms(boxes, scores, 0.5).numpy()
    np.savez_compressed(os.path.join(HERE, "batched_nms.npz"), **rec)
    print("batched_nms", {k: v.shape for k, v in rec.items() if k.endswith("_vanilla") or k.endswith("_agnostic")})


if __name__ == "__main__":
    if len(sys.argv) > 1:                       # python make_golden.py batched_nms collate ...
        for name in sys.argv[1:]:
            globals()["gen_" + name]()
        sys.exit(0)
    gen_anchors()
    gen_loss()
    gen_detect()
    gen_head()
    gen_eval_tail()
    gen_collate()
