"""Pin the oracle (oracle/restate.py) against fixtures produced by the unmodified
reference (tests/golden/make_golden.py).  CPU only."""
import glob
import hashlib
import json
import os

import numpy as np
import pytest
import torch

from oracle import restate as R

GOLDEN = os.path.join(os.path.dirname(__file__), "golden")


def _sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def oracle_anchors(kind, kwargs, H, W):
    kw = dict(kwargs)
    if kind == "k7":
        # ksevendet/architecture/anchors.py:14-21: string recipes are eval'd
        levels = [int(x) for x in kw["pyramid_levels"]]
        mapper = eval(kw.pop("sizes_mapper"))
        kw["sizes"] = [mapper(l) for l in levels]
        kw["scales"] = eval(kw["scales"])
    if kind == "yx":
        return R.effdet_anchors(H, W, **kw)
    if "pyramid_levels" in kw:
        kw["pyramid_levels"] = [int(x) for x in kw["pyramid_levels"]]
    return R.retina_anchors(H, W, **kw)


with open(os.path.join(GOLDEN, "anchors.json")) as f:
    ANCHORS = json.load(f)


@pytest.mark.parametrize("name", sorted(ANCHORS))
def test_anchors_bit_exact(name):
    g = ANCHORS[name]
    a = oracle_anchors(g["kind"], g["kwargs"], g["height"], g["width"])
    assert a.dtype == np.float32 and a.shape == (g["count"], 4)
    assert np.array_equal(a[g["sample_idx"]].astype(np.float64), np.array(g["sample_rows"]))
    assert _sha(a) == g["sha256"]


LOSS = sorted(glob.glob(os.path.join(GOLDEN, "loss_*.npz")))


@pytest.mark.parametrize("path", LOSS, ids=[os.path.basename(p)[:-4] for p in LOSS])
def test_loss_matches_reference(path):
    g = np.load(path)
    v = R.RETINA if str(g["variant"]) == "retina" else R.EFFDET
    cls = torch.from_numpy(g["classifications"]).requires_grad_(True)
    reg = torch.from_numpy(g["regressions"]).requires_grad_(True)
    anc = torch.from_numpy(g["anchors"])
    ann = torch.from_numpy(g["annotations"])
    out = R.focal_loss(cls, reg, anc, ann, v)
    # same ATen ops in the same order; only the reduction tree of .sum() depends on the
    # thread count, hence 1e-6 instead of bit equality
    np.testing.assert_allclose(out.cls_mean.detach().numpy(), g["cls_loss"], rtol=1e-6)
    np.testing.assert_allclose(out.reg_mean.detach().numpy(), g["reg_loss"], rtol=1e-6)
    (out.cls_mean.sum() + out.reg_mean.sum()).backward()
    np.testing.assert_allclose(cls.grad.numpy(), g["dcls"], rtol=1e-6, atol=1e-12)
    np.testing.assert_allclose(reg.grad.numpy(), g["dreg"], rtol=1e-6, atol=1e-12)
    for b, m in enumerate(out.matches):
        if m is None:
            assert (g["state"][b] == 1).all() and (g["iou_argmax"][b] == -1).all()
            continue
        assert np.array_equal(m.iou_max.numpy(), g["iou_max"][b])
        assert np.array_equal(m.argmax.numpy(), g["iou_argmax"][b])
        assert np.array_equal(m.gt_row.numpy(), g["gt_row"][b])
        assert np.array_equal(m.state.numpy(), g["state"][b])
        assert m.num_pos == int((g["state"][b] == 2).sum())


DET = sorted(glob.glob(os.path.join(GOLDEN, "detect_*.npz")))


@pytest.mark.parametrize("use_tv", [True, False], ids=["torchvision", "restated_nms"])
@pytest.mark.parametrize("path", DET, ids=[os.path.basename(p)[:-4] for p in DET])
def test_detect_matches_reference(path, use_tv):
    g = np.load(path)
    yx = str(g["variant"]) == "effdet"
    cls = torch.from_numpy(g["classifications"])
    reg = torch.from_numpy(g["regressions"])
    anc = torch.from_numpy(g["anchors"])
    boxes, dets = R.detect_batch(cls, reg, anc, int(g["height"]), int(g["width"]),
                                 float(g["score_thr"]), float(g["iou_thr"]), yx=yx,
                                 use_torchvision=use_tv)
    assert np.array_equal(boxes.numpy(), g["boxes"])
    for b, d in enumerate(dets):
        assert np.array_equal(d.survivors.numpy(), g[f"survivors_{b}"])
        assert np.array_equal(d.keep_rank.numpy(), g[f"keep_rank_{b}"])
        assert np.array_equal(d.keep_anchor.numpy(), g[f"keep_anchor_{b}"])
        assert np.array_equal(d.scores.numpy(), g[f"scores_{b}"])
        assert np.array_equal(d.labels.numpy(), g[f"labels_{b}"])
        assert np.array_equal(d.boxes.numpy(), g[f"boxes_{b}"])


def test_greedy_nms_equals_torchvision_random():
    from torchvision.ops import nms
    rng = np.random.default_rng(5)
    for n in (1, 2, 63, 64, 65, 500):
        xy = rng.uniform(0, 100, (n, 2)).astype(np.float32)
        wh = rng.uniform(0, 40, (n, 2)).astype(np.float32)
        boxes = np.concatenate([xy, xy + wh], 1)
        boxes[::7, 2:] = boxes[::7, :2]                  # zero-area boxes (0/0 -> NaN -> kept)
        scores = np.round(rng.uniform(0.05, 1, n), 2).astype(np.float32)   # many ties
        for thr in (0.3, 0.5):
            want = nms(torch.from_numpy(boxes), torch.from_numpy(scores), thr).numpy()
            got = R.greedy_nms(boxes, scores, thr)
            assert np.array_equal(got, want)


def test_head_layout_and_activation():
    """The oracle's head restatement (layout + sigmoid) reproduces the reference's
    ClassificationModel output from the logits captured in front of its output_act."""
    g = np.load(os.path.join(GOLDEN, "head_retina.npz"))
    levels = [torch.from_numpy(g[f"logits_l{i}"]) for i in range(3)]
    got = R.head_probabilities(levels, int(g["num_anchors"]), int(g["num_classes"]))
    assert got.shape == g["probs"].shape
    assert np.array_equal(got.numpy(), g["probs"])
    # the shared-head order (layout, then activation: simple_head.py:307-310) agrees to the last bit or one ulp
    late = R.head_probabilities(levels, int(g["num_anchors"]), int(g["num_classes"]), act_first=False)
    np.testing.assert_allclose(late.numpy(), g["probs"], rtol=2e-7, atol=0)


def test_evaluator_tails_match_reference():
    """oracle eval_rows_coco / eval_rows_csv == what the reference's evaluate_coco,
    coco_evaluate_ksevendata and _get_detections produced from the same detections (eval_tail.npz)."""
    g = np.load(os.path.join(GOLDEN, "eval_tail.npz"))
    args = (g["scores"], g["labels"], g["boxes"], g["counts"], g["scales"])
    for tag, thr, valid in (("coco_t005", 0.05, None), ("coco_t030", 0.3, None), ("k7_t005_all", 0.05, None),
                            ("k7_t030_valid", 0.3, [int(v) for v in g["k7_valid"]])):
        rows = R.eval_rows_coco(*args, g["image_ids"], g["category_of_label"], threshold=thr, valid_class_id=valid)
        assert rows.shape == g[tag].shape, tag
        assert np.array_equal(rows, g[tag]), tag          # float32 values carried in float64: exact
    nc = len(g["category_of_label"])
    for tag, thr, md in (("csv_t005_m100", 0.05, 100), ("csv_t030_m10", 0.3, 10)):
        rows, offs = R.eval_rows_csv(*args, nc, score_threshold=thr, max_detections=md)
        assert np.array_equal(offs, g[tag + "_class_offsets"]), tag
        assert np.array_equal(rows, g[tag]), tag


def test_collate_annotations_matches_reference():
    """oracle collate_annotations == annot_padded of the reference's collater (collate.npz)."""
    g = np.load(os.path.join(GOLDEN, "collate.npz"))
    for tag in ("mixed", "all_empty", "single"):
        rows = [g[f"{tag}_rows{i}"] for i in range(int(g[f"{tag}_n"]))]
        out = R.collate_annotations(rows).numpy()
        assert out.shape == g[f"{tag}_padded"].shape, tag
        assert np.array_equal(out, g[f"{tag}_padded"]), tag


# ------------------------------------------------------------------ per-category NMS
@pytest.mark.parametrize("tag", ["small", "many_classes", "one_class", "dense"])
def test_batched_nms_oracle_vs_torchvision_fixture(tag):
    """oracle batched_nms == torchvision 0.26.0 _batched_nms_vanilla, and the coordinate-trick restatement
    == _batched_nms_coordinate_trick, on tests/golden/batched_nms.npz (make_golden.py::gen_batched_nms)."""
    g = np.load(os.path.join(GOLDEN, "batched_nms.npz"))
    boxes, scores, idxs = g[f"{tag}_boxes"], g[f"{tag}_scores"], g[f"{tag}_idxs"]
    assert np.array_equal(R.batched_nms(boxes, scores, idxs, 0.5), g[f"{tag}_vanilla"])
    assert np.array_equal(R.batched_nms_coordinate_trick(boxes, scores, idxs, 0.5), g[f"{tag}_trick"])
    assert np.array_equal(R.greedy_nms(boxes, scores, 0.5), g[f"{tag}_agnostic"])
    if tag == "one_class":                       # a single category is plain NMS
        assert np.array_equal(g[f"{tag}_vanilla"], g[f"{tag}_agnostic"])


def test_soft_nms_oracle_properties():
    """Soft-NMS has no third-party implementation to pin against (oracle header: parity unpinned); the
    restatement is checked on the cases where the paper's algorithm has a known answer."""
    rng = np.random.default_rng(5)
    n = 200
    # (1) pairwise disjoint boxes: nothing decays, order = descending score, every box kept
    xy = np.stack([np.arange(n) * 10.0, np.zeros(n)], 1)
    boxes = np.concatenate([xy, xy + 5.0], 1).astype(np.float32)
    scores = rng.uniform(0.1, 1, n).astype(np.float32)
    for method in ("linear", "gaussian"):
        keep, sc = R.soft_nms(boxes, scores, method=method)
        assert np.array_equal(keep, np.argsort(-scores, kind="stable")) and np.array_equal(sc, scores[keep])
    # (2) two identical boxes: linear weight 1 - 1 = 0 -> the second falls below the score threshold
    b2 = np.array([[0, 0, 10, 10], [0, 0, 10, 10], [50, 50, 60, 60]], np.float32)
    s2 = np.array([0.9, 0.8, 0.3], np.float32)
    keep, sc = R.soft_nms(b2, s2, method="linear", iou_thr=0.3)
    assert keep.tolist() == [0, 2] and np.allclose(sc, [0.9, 0.3])
    keep, sc = R.soft_nms(b2, s2, method="gaussian", sigma=0.5)
    assert keep.tolist() == [0, 2, 1] and np.isclose(sc[2], 0.8 * np.exp(-1 / 0.5), rtol=1e-6)
    # (3) linear with an overlap below the threshold does not decay; the first pick is always the arg-max;
    #     a high score threshold ends the loop early
    b3 = np.array([[0, 0, 10, 10], [8, 0, 18, 10]], np.float32)            # IoU = 20 / 180
    keep, sc = R.soft_nms(b3, np.array([0.5, 0.6], np.float32), method="linear", iou_thr=0.3)
    assert keep.tolist() == [1, 0] and np.array_equal(sc, np.array([0.6, 0.5], np.float32))
    keep, sc = R.soft_nms(boxes, scores, method="gaussian", score_thr=0.5)
    assert keep.size == int((scores >= 0.5).sum())
    # (4) as sigma -> 0 the gaussian form tends to "any overlap removes": greedy NMS with threshold 0
    #     (boxes on a 5-pixel lattice: an overlap is either 0 or at least 25 / 1775, so exp(-iou^2 / 1e-6) underflows)
    c = rng.integers(0, 20, (300, 2)) * 5.0
    wh = rng.integers(2, 7, (300, 2)) * 5.0
    b4 = np.concatenate([c, c + wh], 1).astype(np.float32)
    s4 = rng.permutation(300).astype(np.float32) / 300 + 0.01
    keep, _ = R.soft_nms(b4, s4, method="gaussian", sigma=1e-6, score_thr=1e-3)
    assert np.array_equal(keep, R.greedy_nms(b4, s4, 0.0))
