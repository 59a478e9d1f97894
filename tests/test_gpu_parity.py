"""GPU parity tests: the CUDA path (through the C ABI / the drop-in modules) against the golden
fixtures produced by the unmodified reference and against the oracle on seeded inputs.
Bit-exact: anchors, IoU max / arg-max / state, num_pos, survivors, sort order, NMS keep
(rank and anchor id), labels, scores.  1e-5 relative: losses, gradients, decoded boxes
(boxes additionally get 2e-4 px absolute slack: x1 = cx - w/2 cancels)."""
import glob
import hashlib
import json
import os

import numpy as np
import pytest
import torch

import b200det
from b200det import _lib as L
from b200det import losses as BL
from oracle import restate as R

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
GOLDEN = os.path.join(HERE, "golden")
DEV = "cuda:0"

with open(os.path.join(GOLDEN, "anchors.json")) as f:
    ANCHORS = json.load(f)


def anchor_module(kind, kwargs):
    kw = dict(kwargs)
    if "pyramid_levels" in kw:
        kw["pyramid_levels"] = [int(x) for x in kw["pyramid_levels"]]
    return b200det.AnchorsYX(**kw) if kind in ("yx", "effdet") else b200det.Anchors(**kw)


@pytest.mark.parametrize("name", sorted(ANCHORS))
def test_anchors_bit_exact(name):
    g = ANCHORS[name]
    img = torch.empty(1, 3, g["height"], g["width"], device=DEV)
    a = anchor_module(g["kind"], g["kwargs"])(img)
    assert a.shape == (1, g["count"], 4) and a.dtype == torch.float32 and a.is_cuda
    assert hashlib.sha256(a[0].cpu().numpy().tobytes()).hexdigest() == g["sha256"]


LOSS = sorted(glob.glob(os.path.join(GOLDEN, "loss_*.npz")))
IDS = [os.path.basename(p)[:-4] for p in LOSS]


def load_loss_case(path, from_spec):
    g = np.load(path)
    yx = str(g["variant"]) == "effdet"
    img = torch.empty(1, 3, int(g["height"]), int(g["width"]), device=DEV)
    if from_spec:
        anc = anchor_module("yx" if yx else "retina", json.loads(str(g["anchor_kwargs"])))(img)
        assert torch.equal(anc.cpu(), torch.from_numpy(g["anchors"]))
    else:
        anc = torch.from_numpy(g["anchors"]).to(DEV)     # foreign tensor: kernels read it
    return g, yx, anc


@pytest.mark.parametrize("from_spec", [True, False], ids=["spec", "tensor"])
@pytest.mark.parametrize("path", LOSS, ids=IDS)
def test_match_bit_exact(path, from_spec):
    g, yx, anc = load_loss_case(path, from_spec)
    ann = torch.from_numpy(g["annotations"]).to(DEV)
    thr = (0.4, 0.5) if yx else (0.35, 0.45)
    A = g["anchors"].shape[1]
    assign, num_pos, iou_max, argmax = BL.match(anc, ann, A, thr[0], thr[1],
                                                L.BOX_YXYX if yx else L.BOX_XYXY, debug=True)
    B = ann.shape[0]
    aw = assign[:B * A].view(B, A).cpu().numpy()
    state, row = aw & 3, aw >> 2
    assert np.array_equal(state, g["state"])
    assert np.array_equal(iou_max.cpu().numpy(), g["iou_max"])
    assert np.array_equal(argmax.cpu().numpy(), g["iou_argmax"])
    has = g["gt_row"] >= 0
    assert np.array_equal(row[has], g["gt_row"][has])
    assert np.array_equal(num_pos.cpu().numpy(), (g["state"] == 2).sum(1))


@pytest.mark.parametrize("from_spec", [True, False], ids=["spec", "tensor"])
@pytest.mark.parametrize("path", LOSS, ids=IDS)
def test_match_production_mode_bit_exact(path, from_spec):
    """Without the debug outputs the kernel culls ground-truth boxes that provably stay below neg_thr:
    states, num_pos and the GT row of every POSITIVE anchor must still be the reference's."""
    g, yx, anc = load_loss_case(path, from_spec)
    ann = torch.from_numpy(g["annotations"]).to(DEV)
    thr = (0.4, 0.5) if yx else (0.35, 0.45)
    A = g["anchors"].shape[1]
    B = ann.shape[0]
    zf = torch.full((B, A, 4), 7.0, device=DEV)
    assign, num_pos = BL.match(anc, ann, A, thr[0], thr[1], L.BOX_YXYX if yx else L.BOX_XYXY, zero_fill=zf)
    aw = assign[:B * A].view(B, A).cpu().numpy()
    state, row = aw & 3, aw >> 2
    assert np.array_equal(state, g["state"])
    pos = g["state"] == 2
    assert np.array_equal(row[pos], g["gt_row"][pos])
    assert np.array_equal(num_pos.cpu().numpy(), pos.sum(1))
    assert float(zf.abs().sum()) == 0.0          # the d(reg) buffer was zero-filled on the way


@pytest.mark.parametrize("from_spec", [True, False], ids=["spec", "tensor"])
@pytest.mark.parametrize("path", LOSS, ids=IDS)
def test_focal_loss_and_grads(path, from_spec):
    g, yx, anc = load_loss_case(path, from_spec)
    cls = torch.from_numpy(g["classifications"]).to(DEV).requires_grad_(True)
    reg = torch.from_numpy(g["regressions"]).to(DEV).requires_grad_(True)
    ann = torch.from_numpy(g["annotations"]).to(DEV)
    mod = b200det.FocalLoss.efficientdet() if yx else b200det.FocalLoss()
    lc, lr = mod(cls, reg, anc, ann)
    assert lc.shape == (1,) and lr.shape == (1,)
    np.testing.assert_allclose(lc.detach().cpu().numpy(), g["cls_loss"], rtol=1e-5)
    np.testing.assert_allclose(lr.detach().cpu().numpy(), g["reg_loss"], rtol=1e-5)
    # forward without grad: another instantiation with its own tile geometry (longer slices), so the
    # float32 partial sums associate differently - equal to rounding, and bit-stable from call to call
    with torch.no_grad():
        lc2, lr2 = mod(cls.detach(), reg.detach(), anc, ann)
        lc3, lr3 = mod(cls.detach(), reg.detach(), anc, ann)
    np.testing.assert_allclose(lc2.cpu().numpy(), lc.detach().cpu().numpy(), rtol=1e-6)
    np.testing.assert_allclose(lr2.cpu().numpy(), lr.detach().cpu().numpy(), rtol=1e-6)
    assert torch.equal(lc2, lc3) and torch.equal(lr2, lr3)
    (lc + lr).sum().backward()
    np.testing.assert_allclose(cls.grad.cpu().numpy(), g["dcls"], rtol=1e-5, atol=1e-10)
    np.testing.assert_allclose(reg.grad.cpu().numpy(), g["dreg"], rtol=1e-5, atol=1e-9)
    assert np.array_equal(mod.last_num_positives.cpu().numpy(), (g["state"] == 2).sum(1))


def test_match_ignore_bitmap():
    """k7_match's optional one-bit-per-anchor form of the IGNORE state."""
    g, yx, anc = load_loss_case(LOSS[IDS.index("loss_retina_c90")], True)
    ann = torch.from_numpy(g["annotations"]).to(DEV)
    B, A = g["state"].shape
    assign, num_pos, bits = BL.match(anc, ann, A, 0.35, 0.45, with_bits=True)
    state = (assign.view(B, A) & 3).cpu().numpy()
    got_bits = np.unpackbits(bits.cpu().numpy().view(np.uint8), axis=1, bitorder="little").astype(bool)
    assert np.array_equal(got_bits[:, :A], state == 1)


LOSS_KW = dict(alpha=0.25, beta=1.0 / 9.0, target_div=(0.1, 0.1, 0.2, 0.2), yx_targets=False)


def scatter_words(sp, ann, cls, reg, thr, **kw):
    """match_scatter -> k7_loss(best=..., assign_out=...): (words [B,A], num_pos, loss outputs, best array)."""
    B, A, _ = cls.shape
    arr, gen, consume = BL.match_scatter(sp, ann, A, thr[0], thr[1])
    words = torch.empty(B * A, dtype=torch.int32, device=DEV)
    num_pos = torch.empty(B, dtype=torch.int32, device=DEV)
    out = BL.loss_stage(sp, cls, reg, ann, None, num_pos, best=(arr, gen, consume, thr[0], thr[1]), assign_out=words,
                        mean_divisor=B, **kw)
    return words.view(B, A), num_pos, out, (arr, gen)


@pytest.mark.parametrize("path", LOSS, ids=IDS)
def test_match_scatter_bit_exact(path):
    """The ground-truth-first assignment (k7_match_scatter, the production path with recipe anchors): states,
    num_pos, the IoU maximum of every non-negative anchor and its arg-max row are the reference's."""
    g, yx, anc = load_loss_case(path, True)
    sp = b200det.spec_of(anc)
    ann = torch.from_numpy(g["annotations"]).to(DEV)
    cls = torch.from_numpy(g["classifications"]).to(DEV)
    reg = torch.from_numpy(g["regressions"]).to(DEV)
    thr = (0.4, 0.5) if yx else (0.35, 0.45)
    B, A = g["state"].shape
    kw = dict(LOSS_KW, want_grad=False, target_div=(1, 1, 1, 1) if yx else (0.1, 0.1, 0.2, 0.2), yx_targets=yx)
    words, num_pos, _, (arr, gen) = scatter_words(sp, ann, cls, reg, thr, **kw)
    w = words.cpu().numpy()
    state, row = w & 3, w >> 2
    assert np.array_equal(state, g["state"])
    assert np.array_equal(num_pos.cpu().numpy(), (g["state"] == 2).sum(1))
    has_rows = (g["annotations"][:, :, 4] != -1).any(1)          # an image without a row is skipped: all IGNORE, no entry
    live = (g["state"] != 0) & has_rows[:, None]                 # anchors whose IoU maximum reached neg_thr
    assert np.array_equal(row[live], g["gt_row"][live])
    best = arr[:B * A].view(B, A).cpu().numpy().astype(np.uint64)
    tagged = (best >> np.uint64(48)) == np.uint64(gen)
    assert np.array_equal(tagged, live)                          # an entry exactly where IoU_max >= neg_thr
    iou_bits = ((best >> np.uint64(16)) & np.uint64(0xffffffff)).astype(np.uint32)
    assert np.array_equal(iou_bits[live], g["iou_max"][live].view(np.uint32))


@pytest.mark.parametrize("path", LOSS, ids=IDS)
def test_loss_from_words_and_from_best_agree(path):
    """k7_loss fed with k7_match's words and with k7_match_scatter's array is the same computation, bit for bit;
    consume_best leaves the array clean so that a replayed CUDA graph can reuse one generation."""
    g, yx, anc = load_loss_case(path, True)
    sp = b200det.spec_of(anc)
    ann = torch.from_numpy(g["annotations"]).to(DEV)
    cls = torch.from_numpy(g["classifications"]).to(DEV)
    reg = torch.from_numpy(g["regressions"]).to(DEV)
    thr = (0.4, 0.5) if yx else (0.35, 0.45)
    B, A, _ = cls.shape
    kw = dict(LOSS_KW, want_grad=True, target_div=(1, 1, 1, 1) if yx else (0.1, 0.1, 0.2, 0.2), yx_targets=yx)
    assign, num_pos = BL.match(sp, ann, A, thr[0], thr[1])
    num_pos1 = torch.empty_like(num_pos)
    a = BL.loss_stage(sp, cls, reg, ann, assign, num_pos1, mean_divisor=B, **kw)
    _, num_pos2, b, (arr, gen) = scatter_words(sp, ann, cls, reg, thr, **kw)
    assert torch.equal(num_pos, num_pos1) and torch.equal(num_pos, num_pos2)
    for x, y in zip(a, b):
        assert torch.equal(x, y)
    # replay of ONE generation (what a captured CUDA graph does): scatter -> loss(consume), three times on the same
    # array with the same generation number; the consuming call leaves nothing of that generation behind
    p = L.K7MatchParams(*thr)
    import ctypes as C
    arr, gen, _ = BL.match_scatter(sp, ann, A, thr[0], thr[1])             # a fresh generation, consumed below
    BL.loss_stage(sp, cls, reg, ann, None, best=(arr, gen, True, thr[0], thr[1]), mean_divisor=B, **kw)
    for _ in range(2):
        L.check(L.lib().k7_match_scatter(C.byref(sp), L.ptr(ann), B, A, ann.shape[1], C.byref(p), L.ptr(arr), gen,
                                         None, L.stream_ptr(DEV)), "k7_match_scatter")
        num_pos3 = torch.empty_like(num_pos)
        c = BL.loss_stage(sp, cls, reg, ann, None, num_pos3, best=(arr, gen, True, thr[0], thr[1]), mean_divisor=B, **kw)
        assert torch.equal(num_pos, num_pos3)
        for x, y in zip(a, c):
            assert torch.equal(x, y)
        assert int((arr[:B * A] >> 48 == gen).sum()) == 0        # consumed: nothing of this generation is left


def test_match_scatter_adversarial_vs_exhaustive():
    """Ground truth that stresses the candidate enumeration: duplicates (arg-max ties -> lowest row), boxes that equal
    an anchor exactly, empty / inverted / sub-pixel boxes, boxes larger than the image or outside it, extreme aspect
    ratios, a pad row in the middle, an image without any row.  Reference = the exhaustive kernel in debug mode."""
    H, W = 256, 320
    mod = b200det.Anchors()
    anc = mod(torch.empty(1, 3, H, W, device=DEV))
    sp = b200det.spec_of(anc)
    A = sp.total
    a_np = anc[0].cpu().numpy()
    rows = [list(a_np[1234]) + [1.0], list(a_np[1234]) + [2.0],                 # the same box twice, on an anchor
            list(a_np[A - 1]) + [0.0], list(a_np[0]) + [3.0],
            [10.0, 10.0, 10.0, 50.0, 1.0], [50.0, 60.0, 40.0, 90.0, 1.0],       # empty, inverted
            [100.2, 100.2, 100.9, 100.7, 2.0],                                  # sub-pixel
            [-50.0, -80.0, 400.0, 300.0, 0.0], [330.0, 10.0, 390.0, 70.0, 0.0], # larger than / outside the image
            [5.0, 120.0, 315.0, 131.0, 4.0], [150.0, 3.0, 158.0, 250.0, 4.0],   # 28:1 and 1:31
            [-1.0, -1.0, -1.0, -1.0, -1.0],                                     # pad row in the middle
            [30.5, 40.25, 95.75, 101.5, 2.0], [31.0, 41.0, 96.0, 101.0, 3.0]]
    gen = torch.Generator().manual_seed(5)
    rnd = b200det.synth.make_annotations(1, len(rows), 5, H, W, gen, mode="fixed", gt_min=4.0, gt_max=200.0)[0]
    ann = torch.stack([torch.tensor(rows), rnd, torch.full((len(rows), 5), -1.0)]).to(DEV)
    B = ann.shape[0]
    cls = torch.full((B, A, 3), 0.01, device=DEV)
    reg = torch.zeros((B, A, 4), device=DEV)
    for thr in [(0.35, 0.45), (0.4, 0.5), (0.5, 0.5), (0.05, 0.9)]:
        assign, num_pos, iou_max, argmax = BL.match(sp, ann, A, thr[0], thr[1], debug=True)
        words, num_pos2, _, (arr, g_) = scatter_words(sp, ann, cls, reg, thr, want_grad=False, **LOSS_KW)
        want = assign.view(B, A).cpu().numpy()
        got = words.cpu().numpy()
        assert np.array_equal(got & 3, want & 3), thr
        assert torch.equal(num_pos, num_pos2)
        live = (want & 3) != 0
        live[2] = False                                                         # the image without a row: all IGNORE, no entry
        assert np.array_equal((got >> 2)[live], (want >> 2)[live]), thr
        best = arr[:B * A].view(B, A).cpu().numpy().astype(np.uint64)
        bits = ((best >> np.uint64(16)) & np.uint64(0xffffffff)).astype(np.uint32)
        assert np.array_equal(bits[live], iou_max.cpu().numpy().view(np.uint32)[live]), thr


@pytest.mark.parametrize("name", ["c2_effdet_d0_512", "c3_ksevendet_thermal_512x640", "c2_effdet_d0_512_coded"])
def test_match_scatter_equals_exhaustive_on_workloads(name):
    """Full-size anchor sets of the BASELINE workloads (8 images each): scatter == exhaustive kernel."""
    wl = b200det.synth.WORKLOADS[name]
    B = 8
    _, reg, ann = b200det.synth.make_inputs(b200det.synth.Workload(**{**wl.__dict__, "classes": 1}), seed=3, device=DEV, batch=B)
    anc = b200det.Anchors(**wl.anchor_kwargs())(torch.empty(B, 3, wl.height, wl.width, device=DEV))
    sp = b200det.spec_of(anc)
    A = sp.total
    cls = torch.full((B, A, 1), 0.01, device=DEV)
    assign, num_pos = BL.match(sp, ann, A, 0.35, 0.45)
    words, num_pos2, _, _ = scatter_words(sp, ann, cls, reg, (0.35, 0.45), want_grad=False, **LOSS_KW)
    want, got = assign.view(B, A).cpu().numpy(), words.cpu().numpy()
    assert np.array_equal(got & 3, want & 3)
    assert torch.equal(num_pos, num_pos2) and int(num_pos.sum()) > 0
    pos = (want & 3) == 2
    assert np.array_equal((got >> 2)[pos], (want >> 2)[pos])


def test_second_backward_is_refused():
    """The saved gradients are scaled in place: backward(retain_graph=True) twice would scale them twice."""
    g, yx, anc = load_loss_case(LOSS[IDS.index("loss_retina_small")], True)
    ann = torch.from_numpy(g["annotations"]).to(DEV)
    cls = torch.from_numpy(g["classifications"]).to(DEV).requires_grad_(True)
    reg = torch.from_numpy(g["regressions"]).to(DEV).requires_grad_(True)
    lc, lr = b200det.FocalLoss()(cls, reg, anc, ann)
    loss = (2.0 * lc + lr).sum()
    loss.backward(retain_graph=True)
    np.testing.assert_allclose(cls.grad.cpu().numpy(), 2.0 * g["dcls"], rtol=1e-5, atol=1e-10)
    with pytest.raises(RuntimeError, match="already run"):
        loss.backward()


def test_backward_scales_by_incoming_gradient():
    g, yx, anc = load_loss_case(LOSS[IDS.index("loss_retina_small")], True)
    ann = torch.from_numpy(g["annotations"]).to(DEV)
    cls = torch.from_numpy(g["classifications"]).to(DEV).requires_grad_(True)
    reg = torch.from_numpy(g["regressions"]).to(DEV).requires_grad_(True)
    lc, lr = b200det.FocalLoss()(cls, reg, anc, ann)
    (3.0 * lc + 0.5 * lr).sum().backward()
    np.testing.assert_allclose(cls.grad.cpu().numpy(), 3.0 * g["dcls"], rtol=1e-5, atol=1e-10)
    np.testing.assert_allclose(reg.grad.cpu().numpy(), 0.5 * g["dreg"], rtol=1e-5, atol=1e-9)


def test_loss_is_bit_reproducible():
    g, yx, anc = load_loss_case(LOSS[IDS.index("loss_retina_c90")], True)
    ann = torch.from_numpy(g["annotations"]).to(DEV)
    cls = torch.from_numpy(g["classifications"]).to(DEV)
    reg = torch.from_numpy(g["regressions"]).to(DEV)
    outs = [torch.cat(b200det.FocalLoss()(cls, reg, anc, ann)).cpu() for _ in range(5)]
    assert all(torch.equal(outs[0], o) for o in outs[1:])


@pytest.mark.parametrize("shape", [(3, 1, 7), (5, 3, 4), (1, 90, 3), (2, 80, 9), (3, 33, 5), (2, 7, 1), (1, 130, 2)],
                         ids=lambda s: "B%d_C%d_M%d" % s)
def test_loss_odd_shapes_vs_oracle(shape):
    """Shapes whose element count is not a multiple of 4 (scalar tail of the bulk copies), every
    vector width, lanes-per-row 1/2/4/8, tiles straddling images; anchors as a foreign tensor."""
    B, C, M = shape
    H, W = 77, 99
    anc_np = R.retina_anchors(H, W)
    A = anc_np.shape[0] - 1 - (B % 2)          # odd/even anchor counts on purpose
    anc_np = anc_np[:A]
    gen = torch.Generator().manual_seed(B * 1000 + C * 10 + M)
    ann = b200det.synth.make_annotations(B, M, C, H, W, gen, gt_min=8.0, gt_max=60.0)
    cls, reg = b200det.synth.make_head_outputs(B, A, C, gen, logit_mean=-2.0, logit_std=2.0)
    anc = torch.from_numpy(anc_np)[None]
    c0, r0 = cls.clone().requires_grad_(True), reg.clone().requires_grad_(True)
    want = R.focal_loss(c0, r0, anc, ann)
    (want.cls_mean + want.reg_mean).sum().backward()
    c1, r1 = cls.to(DEV).requires_grad_(True), reg.to(DEV).requires_grad_(True)
    mod = b200det.FocalLoss()
    lc, lr = mod(c1, r1, anc.to(DEV), ann.to(DEV))
    (lc + lr).sum().backward()
    np.testing.assert_allclose(lc.item(), want.cls_mean.item(), rtol=1e-5)
    np.testing.assert_allclose(lr.item(), want.reg_mean.item(), rtol=1e-5)
    np.testing.assert_allclose(mod.last_per_image.cpu().numpy(),
                               torch.stack([want.cls_per_image, want.reg_per_image], 1).detach().numpy(), rtol=1e-5)
    assert mod.last_num_positives.cpu().tolist() == want.num_pos
    np.testing.assert_allclose(c1.grad.cpu().numpy(), c0.grad.numpy(), rtol=1e-5, atol=1e-10)
    want_dreg = r0.grad.numpy() if r0.grad is not None else np.zeros_like(reg.numpy())   # no positive at all
    np.testing.assert_allclose(r1.grad.cpu().numpy(), want_dreg, rtol=1e-5, atol=1e-9)


# ---------------------------------------------------------------------------------- inference
DET = sorted(glob.glob(os.path.join(GOLDEN, "detect_*.npz")))
DIDS = [os.path.basename(p)[:-4] for p in DET]


def load_det_case(path, from_spec):
    g = np.load(path)
    yx = str(g["variant"]) == "effdet"
    H, W = int(g["height"]), int(g["width"])
    img = torch.empty(g["classifications"].shape[0], 3, H, W, device=DEV)
    if from_spec:
        anc = anchor_module("yx" if yx else "retina", json.loads(str(g["anchor_kwargs"])))(img)
    else:
        anc = torch.from_numpy(g["anchors"]).to(DEV)
    return g, yx, anc, img


@pytest.mark.parametrize("from_spec", [True, False], ids=["spec", "tensor"])
@pytest.mark.parametrize("path", DET, ids=DIDS)
def test_decode_and_clip_modules(path, from_spec):
    g, yx, anc, img = load_det_case(path, from_spec)
    reg = torch.from_numpy(g["regressions"]).to(DEV)
    rb = b200det.BBoxTransformYX() if yx else b200det.BBoxTransform()
    boxes = rb(anc, reg)
    out = b200det.ClipBoxes()(boxes, img)
    assert out.data_ptr() == boxes.data_ptr()            # in place, same tensor (utils.py:71)
    np.testing.assert_allclose(out.cpu().numpy(), g["boxes"], rtol=1e-5, atol=1e-4)
    if not yx:                                           # batched anchors [B,A,4] broadcast form
        ab = torch.from_numpy(g["anchors"]).to(DEV).expand(reg.shape[0], -1, -1).contiguous()
        out2 = b200det.ClipBoxes()(rb(ab, reg), img)
        assert torch.equal(out2, out)


@pytest.mark.parametrize("path", DET, ids=DIDS)
def test_nms_stage_bit_exact_on_reference_boxes(path):
    """k7_nms on the reference's own boxes/scores must return torchvision's keep exactly."""
    g = np.load(path)
    for b in range(g["classifications"].shape[0]):
        surv = g[f"survivors_{b}"]
        boxes = torch.from_numpy(g["boxes"][b][surv]).to(DEV)
        scores = torch.from_numpy(g["classifications"][b].max(axis=1)[surv]).to(DEV)
        keep = b200det.nms(boxes, scores, float(g["iou_thr"]))
        assert keep.dtype == torch.int64
        assert np.array_equal(keep.cpu().numpy(), g[f"keep_rank_{b}"])


@pytest.mark.parametrize("n", [1, 2, 63, 64, 65, 1000, 1024, 1025, 2500, 6000])
def test_nms_random_vs_torchvision(n):
    from torchvision.ops import nms as tv_nms
    rng = np.random.default_rng(n)
    xy = rng.uniform(0, 300, (n, 2)).astype(np.float32)
    wh = rng.uniform(0, 80, (n, 2)).astype(np.float32)
    boxes = np.concatenate([xy, xy + wh], 1)
    boxes[::11, 2:] = boxes[::11, :2]                    # zero-area boxes: 0/0 -> NaN -> never suppressed
    scores = np.round(rng.uniform(-1, 1, n), 2).astype(np.float32)     # many exact ties, negatives
    for thr in (0.0, 0.3, 0.5, 0.9):
        want = tv_nms(torch.from_numpy(boxes), torch.from_numpy(scores), thr).numpy()
        got = b200det.nms(torch.from_numpy(boxes).to(DEV), torch.from_numpy(scores).to(DEV), thr).cpu().numpy()
        assert np.array_equal(got, want), (n, thr)


@pytest.mark.parametrize("n", [5000, 16384, 16385, 20000])
def test_nms_multiscale_clusters_vs_torchvision(n):
    """Boxes of very different sizes (eight area binades) piled up in clusters: long suppression
    chains for the fixed-point resolve, every cell of the spatial order populated, the 16 384-box
    cap of the multi-SM path and the single-CTA fallback above it; a few NaN / infinite boxes."""
    from torchvision.ops import nms as tv_nms
    rng = np.random.default_rng(n)
    centers = rng.uniform(0, 1500, (40, 2))
    which = rng.integers(0, 40, n)
    size = 2.0 ** rng.uniform(3, 9, n)
    c = centers[which] + rng.normal(0, 1, (n, 2)) * size[:, None] * 0.4
    w = size * rng.uniform(0.6, 1.6, n)
    h = size * rng.uniform(0.6, 1.6, n)
    boxes = np.stack([c[:, 0] - w / 2, c[:, 1] - h / 2, c[:, 0] + w / 2, c[:, 1] + h / 2], 1).astype(np.float32)
    boxes[7] = np.nan
    boxes[11, 2] = np.inf
    boxes[13] = boxes[12]                                # exact duplicate
    scores = rng.uniform(0, 1, n).astype(np.float32)
    scores[13] = scores[12]                              # tie: the lower index wins
    for thr in (0.5, 0.7):
        want = tv_nms(torch.from_numpy(boxes), torch.from_numpy(scores), thr).numpy()
        got = b200det.nms(torch.from_numpy(boxes).to(DEV), torch.from_numpy(scores).to(DEV), thr).cpu().numpy()
        assert np.array_equal(got, want), (n, thr)


def test_nms_empty():
    keep = b200det.nms(torch.zeros(0, 4, device=DEV), torch.zeros(0, device=DEV), 0.5)
    assert keep.shape == (0,) and keep.dtype == torch.int64


# ------------------------------------------------------------------ per-category NMS
@pytest.mark.parametrize("tag", ["small", "many_classes", "one_class", "dense"])
def test_batched_nms_fixture_bit_exact(tag):
    """k7_batched_nms == torchvision 0.26.0 batched_nms on the committed fixture: the per-category strategy
    and the coordinate trick, each against torchvision's own output of that strategy."""
    g = np.load(os.path.join(GOLDEN, "batched_nms.npz"))
    boxes = torch.from_numpy(g[f"{tag}_boxes"]).to(DEV)
    scores = torch.from_numpy(g[f"{tag}_scores"]).to(DEV)
    idxs = torch.from_numpy(g[f"{tag}_idxs"]).to(DEV)
    got = b200det.batched_nms(boxes, scores, idxs, 0.5)
    assert got.dtype == torch.int64 and np.array_equal(got.cpu().numpy(), g[f"{tag}_vanilla"])
    got = b200det.batched_nms(boxes, scores, idxs, 0.5, strategy="coordinate_trick")
    assert np.array_equal(got.cpu().numpy(), g[f"{tag}_trick"])
    assert np.array_equal(b200det.nms(boxes, scores, 0.5).cpu().numpy(), g[f"{tag}_agnostic"])


@pytest.mark.parametrize("n,ncls", [(1, 1), (65, 2), (1000, 7), (3000, 90), (9000, 4), (17000, 3), (20000, 90)])
def test_batched_nms_random_vs_oracle(n, ncls):
    """Clustered boxes of mixed sizes, score ties, negative category ids; n > 16 384 takes the single-CTA
    chunked kernel.  Oracle = per-category greedy NMS (restatement of torchvision's vanilla strategy);
    n > 1000 is also what torchvision.ops.batched_nms itself does on CPU."""
    rng = np.random.default_rng(n + ncls)
    centers = rng.uniform(0, 800, (max(n // 8, 1), 2))
    c = centers[rng.integers(0, centers.shape[0], n)] + rng.normal(0, 4, (n, 2))
    wh = 2.0 ** rng.uniform(3, 7, (n, 2))
    boxes = np.concatenate([c - wh / 2, c + wh / 2], 1).astype(np.float32)
    scores = np.round(rng.uniform(0, 1, n), 3).astype(np.float32)      # ties
    idxs = rng.integers(0, ncls, n).astype(np.int64) - 1               # includes -1
    for thr in (0.3, 0.5):
        want = R.batched_nms(boxes, scores, idxs, thr)
        got = b200det.batched_nms(torch.from_numpy(boxes).to(DEV), torch.from_numpy(scores).to(DEV),
                                  torch.from_numpy(idxs).to(DEV), thr).cpu().numpy()
        assert np.array_equal(got, want), (n, ncls, thr)
    if n > 1000 and np.unique(scores).size == n:                       # torchvision's own public entry point
        from torchvision.ops import batched_nms as tv_bnms
        want = tv_bnms(torch.from_numpy(boxes), torch.from_numpy(scores), torch.from_numpy(idxs), 0.5).numpy()
        assert np.array_equal(got, want)


def test_batched_nms_empty_and_errors():
    z = torch.zeros(0, 4, device=DEV)
    keep = b200det.batched_nms(z, torch.zeros(0, device=DEV), torch.zeros(0, dtype=torch.int64, device=DEV), 0.5)
    assert keep.shape == (0,) and keep.dtype == torch.int64
    with pytest.raises(ValueError):
        b200det.batched_nms(torch.zeros(3, 4, device=DEV), torch.zeros(3, device=DEV),
                            torch.zeros(2, dtype=torch.int64, device=DEV), 0.5)
    with pytest.raises(ValueError):
        b200det.batched_nms(torch.zeros(3, 4, device=DEV), torch.zeros(3, device=DEV),
                            torch.zeros(3, dtype=torch.int64, device=DEV), 0.5, strategy="fastest")


@pytest.mark.parametrize("name", ["detect_retina_small", "detect_retina_c90"])
def test_detect_class_aware_vs_oracle(name):
    """detect(class_aware=True): the eval tail with per-class suppression.  Oracle: the reference's own
    threshold / arg-max (retinanet.py:300-318) followed by the per-category NMS restatement."""
    g, yx, anc, img = load_det_case(DET[DIDS.index(name)], True)
    cls = torch.from_numpy(g["classifications"]).to(DEV)
    reg = torch.from_numpy(g["regressions"]).to(DEV)
    det = b200det.detect(cls, reg, anc, (int(g["height"]), int(g["width"])), class_aware=True)
    agn = b200det.detect(cls, reg, anc, (int(g["height"]), int(g["width"])))
    more = 0
    for b in range(cls.shape[0]):
        surv = g[f"survivors_{b}"]
        sc, lb = torch.from_numpy(g["classifications"][b][surv]).max(dim=1)
        keep = R.batched_nms(g["boxes"][b][surv], sc.numpy(), lb.numpy(), 0.5)
        k = int(det.count[b])
        assert k == keep.size
        assert np.array_equal(det.survivor_rank[b, :k].cpu().numpy(), keep)
        assert np.array_equal(det.anchor_ids[b, :k].cpu().numpy(), surv[keep])
        assert np.array_equal(det.labels[b, :k].cpu().numpy(), lb.numpy()[keep])
        assert np.array_equal(det.scores[b, :k].cpu().numpy(), sc.numpy()[keep])
        more += k - int(agn.count[b])
    assert more >= 0            # suppression inside a class only can never keep fewer boxes


@pytest.mark.parametrize("class_aware", [False, True], ids=["agnostic", "per-class"])
def test_detect_image_above_the_multi_sm_cap(class_aware):
    """An image with more survivors than the multi-SM NMS is sized for (16 384) takes the chunked
    single-CTA path inside the resolve kernel, next to an ordinary image in the same batch."""
    wl = b200det.synth.WORKLOADS["c1_retinanet_r50_512"]
    gen = torch.Generator().manual_seed(11)
    A, Cn = wl.anchors, 2
    cls, reg = b200det.synth.make_head_outputs(2, A, Cn, gen, logit_mean=-6.0, logit_std=1.0)
    z = torch.randn((A, Cn), generator=gen) * 1.5 - 3.0                # image 0: about 40 % of the anchors survive
    cls[0] = torch.sigmoid(z)
    anc = torch.from_numpy(R.retina_anchors(wl.height, wl.width))
    ref_boxes = R.clip_boxes_(R.decode_boxes(anc[None], reg), wl.height, wl.width)
    img = torch.empty(2, 3, wl.height, wl.width, device=DEV)
    anc_dev = b200det.Anchors()(img)
    # the NMS oracle runs on the device's own decoded boxes (40 % of 49 104 boxes: with millions of overlapping
    # pairs an IoU within one ulp of the threshold is likely, and expf differs between CPU and GPU by an ulp)
    boxes = b200det.ClipBoxes()(b200det.BBoxTransform()(anc_dev, reg.to(DEV)), img).cpu()
    np.testing.assert_allclose(boxes.numpy(), ref_boxes.numpy(), rtol=1e-5, atol=1e-4)
    det = b200det.detect(cls.to(DEV), reg.to(DEV), anc_dev, (wl.height, wl.width), class_aware=class_aware, cap=A)
    from torchvision.ops import nms as tv_nms
    for b in range(2):
        sc, lb = cls[b].max(dim=1)
        surv = torch.nonzero(sc > 0.05).flatten()
        assert int(det.survivors[b]) == surv.numel()
        assert (surv.numel() > 16384) == (b == 0)
        k = int(det.count[b])
        dev_boxes = det.boxes[b, :k].cpu()
        # NMS on the device's own boxes of the survivors (bit-exact stage test; decode has its own tests)
        ids = det.anchor_ids[b, :k].cpu().long()
        assert torch.equal(dev_boxes, boxes[b][ids])
        if class_aware:
            keep = torch.from_numpy(R.batched_nms(boxes[b][surv].numpy(), sc[surv].numpy(), lb[surv].numpy(), 0.5))
        else:
            keep = tv_nms(boxes[b][surv], sc[surv], 0.5)
        assert k == keep.numel()
        assert np.array_equal(det.survivor_rank[b, :k].cpu().numpy(), keep.numpy())
        assert np.array_equal(ids.numpy(), surv[keep].numpy())
        assert np.array_equal(det.labels[b, :k].cpu().numpy(), lb[surv][keep].numpy())


@pytest.mark.parametrize("n", [1, 50, 1000, 3000])
def test_soft_nms_vs_oracle(n):
    """Soft-NMS (soft variants).  Linear decay is plain float32 arithmetic: picks and scores are
    bit-exact against the restatement of the paper's algorithm.  The gaussian form goes through expf, which differs
    between host and device by an ulp: same picks, scores to 1e-5."""
    rng = np.random.default_rng(100 + n)
    centers = rng.uniform(0, 400, (max(n // 10, 1), 2))
    c = centers[rng.integers(0, centers.shape[0], n)] + rng.normal(0, 5, (n, 2))
    wh = 2.0 ** rng.uniform(3, 6, (n, 2))
    boxes = np.concatenate([c - wh / 2, c + wh / 2], 1).astype(np.float32)
    boxes[::17, 2:] = boxes[::17, :2]                                  # zero-area boxes: no decay either way
    scores = (rng.permutation(n).astype(np.float32) + 1) / np.float32(n + 1)
    tb, ts = torch.from_numpy(boxes).to(DEV), torch.from_numpy(scores).to(DEV)
    for thr in (0.3, 0.5):
        want_k, want_s = R.soft_nms(boxes, scores, method="linear", iou_thr=thr, score_thr=0.05)
        got_k, got_s = b200det.soft_nms(tb, ts, method="linear", iou_threshold=thr, score_threshold=0.05)
        assert got_k.dtype == torch.int64 and np.array_equal(got_k.cpu().numpy(), want_k)
        assert np.array_equal(got_s.cpu().numpy(), want_s)
    want_k, want_s = R.soft_nms(boxes, scores, method="gaussian", sigma=0.5, score_thr=0.01)
    got_k, got_s = b200det.soft_nms(tb, ts, method="gaussian", sigma=0.5, score_threshold=0.01)
    gk, gs = got_k.cpu().numpy(), got_s.cpu().numpy()
    assert gk.size == want_k.size and set(gk.tolist()) == set(want_k.tolist())
    order_w = np.argsort(want_k)
    order_g = np.argsort(gk)
    np.testing.assert_allclose(gs[order_g], want_s[order_w], rtol=1e-5)      # every box: the same decayed score
    assert np.mean(gk == want_k) > 0.99                                       # and the same pick order but for ulp ties


def test_soft_nms_empty_and_errors():
    k, s = b200det.soft_nms(torch.zeros(0, 4, device=DEV), torch.zeros(0, device=DEV))
    assert k.shape == (0,) and s.shape == (0,) and k.dtype == torch.int64
    with pytest.raises(ValueError):
        b200det.soft_nms(torch.zeros(3, 4, device=DEV), torch.zeros(3, device=DEV), method="hard")
    with pytest.raises(ValueError):
        b200det.soft_nms(torch.zeros(3, 4, device=DEV), torch.zeros(2, device=DEV))
    with pytest.raises(ValueError):
        b200det.soft_nms(torch.zeros(3, 4, device=DEV), torch.zeros(3, device=DEV), method="gaussian", sigma=0.0)


def check_detect(det, g, exact_boxes_from=None):
    B = g["classifications"].shape[0]
    for b in range(B):
        k = int(det.count[b])
        assert int(det.survivors[b]) == g[f"survivors_{b}"].size
        assert np.array_equal(det.anchor_ids[b, :k].cpu().numpy(), g[f"keep_anchor_{b}"])
        assert np.array_equal(det.survivor_rank[b, :k].cpu().numpy(), g[f"keep_rank_{b}"])
        assert np.array_equal(det.labels[b, :k].cpu().numpy(), g[f"labels_{b}"])
        assert np.array_equal(det.scores[b, :k].cpu().numpy(), g[f"scores_{b}"])
        np.testing.assert_allclose(det.boxes[b, :k].cpu().numpy(), g[f"boxes_{b}"], rtol=1e-5, atol=1e-4)


@pytest.mark.parametrize("chain", ["auto", "per_image", "multi_sm"])
@pytest.mark.parametrize("from_spec", [True, False], ids=["spec", "tensor"])
@pytest.mark.parametrize("path", DET, ids=DIDS)
def test_detect_end_to_end(path, from_spec, chain):
    g, yx, anc, img = load_det_case(path, from_spec)
    cls = torch.from_numpy(g["classifications"]).to(DEV)
    reg = torch.from_numpy(g["regressions"]).to(DEV)
    det = b200det.detect(cls, reg, anc, (int(g["height"]), int(g["width"])), score_thr=float(g["score_thr"]),
                         iou_thr=float(g["iou_thr"]), variant=L.DECODE_EFFDET if yx else L.DECODE_RETINA, cap=8,
                         nms_chain=chain)
    check_detect(det, g)


@pytest.mark.parametrize("path", DET, ids=DIDS)
def test_postprocess_signature(path):
    g, yx, anc, img = load_det_case(path, True)
    cls = torch.from_numpy(g["classifications"]).to(DEV)
    reg = torch.from_numpy(g["regressions"]).to(DEV)
    rb = b200det.BBoxTransformYX() if yx else b200det.BBoxTransform()
    out = b200det.postprocess(img, anc, reg, cls, rb, b200det.ClipBoxes(), float(g["score_thr"]), float(g["iou_thr"]))
    assert len(out) == cls.shape[0]
    for b, o in enumerate(out):
        assert np.array_equal(np.asarray(o["class_ids"]).reshape(-1), g[f"labels_{b}"])
        assert np.array_equal(np.asarray(o["scores"], np.float32).reshape(-1), g[f"scores_{b}"])
        np.testing.assert_allclose(np.asarray(o["rois"], np.float32).reshape(-1, 4), g[f"boxes_{b}"], rtol=1e-5, atol=1e-4)


def test_eval_tail_image0_and_empty_triple():
    g, yx, anc, img = load_det_case(DET[DIDS.index("detect_retina_small")], True)
    cls = torch.from_numpy(g["classifications"]).to(DEV)
    reg = torch.from_numpy(g["regressions"]).to(DEV)
    s, l, b = b200det.eval_tail(cls, reg, anc, img)
    assert np.array_equal(s.cpu().numpy(), g["scores_0"]) and np.array_equal(l.cpu().numpy(), g["labels_0"])
    assert l.dtype == torch.int64
    g, yx, anc, img = load_det_case(DET[DIDS.index("detect_retina_empty")], True)
    cls = torch.from_numpy(g["classifications"]).to(DEV)
    reg = torch.from_numpy(g["regressions"]).to(DEV)
    s, l, b = b200det.eval_tail(cls, reg, anc, img)      # retinanet.py:308: zeros triple
    assert s.shape == (1,) and l.shape == (1,) and b.shape == (1, 4) and float(s.abs().sum() + b.abs().sum()) == 0.0


def test_detect_max_detections_and_topk():
    g, yx, anc, img = load_det_case(DET[DIDS.index("detect_retina_small")], True)
    cls = torch.from_numpy(g["classifications"]).to(DEV)
    reg = torch.from_numpy(g["regressions"]).to(DEV)
    det = b200det.detect(cls, reg, anc, (int(g["height"]), int(g["width"])), max_detections=100)
    for b in range(cls.shape[0]):       # csv_eval.py:107 top-100 of a score-sorted list == prefix
        assert int(det.count[b]) == 100
        assert np.array_equal(det.anchor_ids[b, :100].cpu().numpy(), g[f"keep_anchor_{b}"][:100])
    det = b200det.detect(cls, reg, anc, (int(g["height"]), int(g["width"])), pre_nms_topk=500)
    _, dets = R.detect_batch(cls.cpu(), reg.cpu(), torch.from_numpy(g["anchors"]), int(g["height"]), int(g["width"]))
    for b in range(cls.shape[0]):
        sc = cls[b].max(1)[0].cpu()
        surv = torch.from_numpy(g[f"survivors_{b}"])
        order = torch.argsort(-sc[surv], stable=True)[:500]
        from torchvision.ops import nms as tv_nms
        keep = tv_nms(torch.from_numpy(g["boxes"][b])[surv][order], sc[surv][order], 0.5)
        assert np.array_equal(det.anchor_ids[b, :int(det.count[b])].cpu().numpy(), surv[order][keep].numpy())


def test_detect_large_vs_oracle():
    """BASELINE config C1 shape (512x512, 80 classes) with dense scores: tens of chunks of NMS."""
    wl = b200det.synth.WORKLOADS["c1_retinanet_r50_512"]
    gen = torch.Generator().manual_seed(4)
    cls, reg = b200det.synth.make_head_outputs(2, wl.anchors, wl.classes, gen, logit_mean=-6.0, logit_std=1.5)
    anc_np = R.retina_anchors(wl.height, wl.width)
    boxes, dets = R.detect_batch(cls, reg, torch.from_numpy(anc_np), wl.height, wl.width)
    img = torch.empty(2, 3, wl.height, wl.width, device=DEV)
    det = b200det.detect(cls.to(DEV), reg.to(DEV), b200det.Anchors()(img), (wl.height, wl.width))
    for b, d in enumerate(dets):
        k = int(det.count[b])
        assert int(det.survivors[b]) == d.survivors.numel() and d.survivors.numel() > 2048
        assert np.array_equal(det.anchor_ids[b, :k].cpu().numpy(), d.keep_anchor.numpy())
        assert np.array_equal(det.labels[b, :k].cpu().numpy(), d.labels.numpy())
        np.testing.assert_allclose(det.boxes[b, :k].cpu().numpy(), d.boxes.numpy(), rtol=1e-5, atol=1e-4)


def test_loss_full_size_properties():
    """BASELINE config C2 shape (B=8 of 64): oracle too slow for the full batch in a unit test, so
    check size-independent properties: per-image losses equal the batch-1 losses (images are
    independent), the mean is their mean, gradients of untouched (ignored) rows are zero."""
    wl = b200det.synth.WORKLOADS["c2_effdet_d0_512"]
    cls, reg, ann = b200det.synth.make_inputs(wl, seed=1, device=DEV, batch=8)
    img = torch.empty(8, 3, wl.height, wl.width, device=DEV)
    anc = b200det.Anchors()(img)
    mod = b200det.FocalLoss()
    c = cls.clone().requires_grad_(True)
    lc, lr = mod(c, reg, anc, ann)
    per = mod.last_per_image.clone()
    npos = mod.last_num_positives.clone()
    lc.backward()
    np.testing.assert_allclose(lc.item(), per[:, 0].double().mean().item(), rtol=1e-6)
    for b in (0, 5):
        lcb, lrb = mod(cls[b:b + 1], reg[b:b + 1], anc, ann[b:b + 1])
        np.testing.assert_allclose(lcb.item(), per[b, 0].item(), rtol=1e-6)
        np.testing.assert_allclose(lrb.item(), per[b, 1].item(), rtol=1e-6)
    # oracle on one image of the full-size batch
    want = R.focal_loss(cls[:1].cpu(), reg[:1].cpu(), anc.cpu(), ann[:1].cpu())
    np.testing.assert_allclose(per[0].cpu().numpy(), [want.cls_mean.item(), want.reg_mean.item()], rtol=1e-5)
    assert int(npos[0]) == want.num_pos[0]


# ---------------------------------------------------------------------------------- BASELINE shapes
@pytest.mark.parametrize("name", ["c3_ksevendet_thermal_512x640", "c4_effdet_d3_896", "c5_effdet_d7_1536"])
def test_baseline_shapes_last_image_vs_oracle(name):
    """BASELINE configs C3 (single class), C4 (A*C % 4 == 2: image bases are not 16-byte aligned)
    and C5 (442k anchors) at their real anchor counts, batch 3: the LAST image of the batch (it
    starts in the middle of a streaming tile) against the oracle run on that image alone -
    loss, num_pos, both gradients, and the full eval tail."""
    wl = b200det.synth.WORKLOADS[name]
    B, b = 3, 2
    cls, reg, ann = b200det.synth.make_inputs(wl, seed=2, device=DEV, batch=B)
    img = torch.empty(B, 3, wl.height, wl.width, device=DEV)
    anc = b200det.Anchors(**wl.anchor_kwargs())(img)
    anc_cpu = torch.from_numpy(R.retina_anchors(wl.height, wl.width, **wl.anchor_kwargs()))[None]
    assert torch.equal(anc.cpu(), anc_cpu)
    mod = b200det.FocalLoss()
    c1, r1 = cls.clone().requires_grad_(True), reg.clone().requires_grad_(True)
    lc, lr = mod(c1, r1, anc, ann)
    (lc + lr).sum().backward()

    c0 = cls[b:b + 1].cpu().requires_grad_(True)
    r0 = reg[b:b + 1].cpu().requires_grad_(True)
    want = R.focal_loss(c0, r0, anc_cpu, ann[b:b + 1].cpu())
    (want.cls_mean + want.reg_mean).sum().backward()
    assert int(mod.last_num_positives[b]) == want.num_pos[0]
    np.testing.assert_allclose(mod.last_per_image[b].cpu().numpy(),
                               [want.cls_mean.item(), want.reg_mean.item()], rtol=1e-5)
    # the module's mean runs over B images, the oracle's over one
    np.testing.assert_allclose(c1.grad[b].cpu().numpy() * B, c0.grad[0].numpy(), rtol=1e-5, atol=1e-10)
    np.testing.assert_allclose(r1.grad[b].cpu().numpy() * B, r0.grad[0].numpy(), rtol=1e-5, atol=1e-9)
    # the batch mean is the mean of the per-image losses
    np.testing.assert_allclose(lc.item(), mod.last_per_image[:, 0].double().mean().item(), rtol=1e-6)
    np.testing.assert_allclose(lr.item(), mod.last_per_image[:, 1].double().mean().item(), rtol=1e-6)

    det = b200det.detect(cls, reg, anc, (wl.height, wl.width))
    _, dets = R.detect_batch(cls[b:b + 1].cpu(), reg[b:b + 1].cpu(), anc_cpu, wl.height, wl.width)
    d = dets[0]
    k = int(det.count[b])
    assert int(det.survivors[b]) == d.survivors.numel()
    assert np.array_equal(det.anchor_ids[b, :k].cpu().numpy(), d.keep_anchor.numpy())
    assert np.array_equal(det.labels[b, :k].cpu().numpy(), d.labels.numpy())
    np.testing.assert_array_equal(det.scores[b, :k].cpu().numpy(), d.scores.numpy())
    np.testing.assert_allclose(det.boxes[b, :k].cpu().numpy(), d.boxes.numpy(), rtol=1e-5, atol=1e-4)


# ---------------------------------------------------------------------------------- fused sigmoid
@pytest.mark.parametrize("shape", [(2, 5, 6), (3, 90, 9), (2, 1, 7), (1, 33, 4)], ids=lambda s: "B%d_C%d_M%d" % s)
def test_focal_loss_from_logits_vs_oracle(shape):
    """FocalLoss(from_logits=True) on the head's pre-sigmoid logits against the oracle's
    sigmoid (head activation, retinanet.py:159) + focal loss, gradients taken with respect to the logits."""
    B, C, M = shape
    H, W = 96, 128
    anc_np = R.retina_anchors(H, W)
    A = anc_np.shape[0]
    gen = torch.Generator().manual_seed(B * 100 + C)
    ann = b200det.synth.make_annotations(B, M, C, H, W, gen, gt_min=8.0, gt_max=70.0)
    z = (torch.randn((B, A, C), generator=gen) * 2.5 - 3.0).contiguous()
    z[0, :50] = torch.linspace(-12.0, 12.0, 50)[:, None]       # both clamp ends of losses.py:62
    reg = torch.randn((B, A, 4), generator=gen) * 0.5
    anc = torch.from_numpy(anc_np)[None]
    z0, r0 = z.clone().requires_grad_(True), reg.clone().requires_grad_(True)
    want = R.focal_loss(torch.sigmoid(z0), r0, anc, ann)
    (want.cls_mean + want.reg_mean).sum().backward()
    z1, r1 = z.to(DEV).requires_grad_(True), reg.to(DEV).requires_grad_(True)
    mod = b200det.FocalLoss(from_logits=True)
    lc, lr = mod(z1, r1, anc.to(DEV), ann.to(DEV))
    (lc + lr).sum().backward()
    np.testing.assert_allclose(lc.item(), want.cls_mean.item(), rtol=1e-5)
    np.testing.assert_allclose(lr.item(), want.reg_mean.item(), rtol=1e-5)
    assert mod.last_num_positives.cpu().tolist() == want.num_pos
    # 2e-5: the reference's own float32 sigmoid backward p*(1-p) carries a rounding of this size for p -> 1
    np.testing.assert_allclose(z1.grad.cpu().numpy(), z0.grad.numpy(), rtol=2e-5, atol=1e-10)
    np.testing.assert_allclose(r1.grad.cpu().numpy(), r0.grad.numpy(), rtol=1e-5, atol=1e-9)
    # the probability path on sigmoid(z) agrees with the fused path
    with torch.no_grad():
        lc2, lr2 = b200det.FocalLoss()(torch.sigmoid(z.to(DEV)), reg.to(DEV), anc.to(DEV), ann.to(DEV))
    np.testing.assert_allclose(lc2.item(), lc.item(), rtol=1e-5)
    np.testing.assert_allclose(lr2.item(), lr.item(), rtol=1e-6)


def test_detect_from_logits_matches_probability_path():
    """detect(from_logits=True) == detect on torch.sigmoid(logits): same survivors (strict threshold on
    the float32 probability), bit-identical scores, labels decided on the probabilities (an earlier
    class whose lower logit rounds to the same probability wins, saturated logits tie), same keep."""
    H, W, B, C = 128, 160, 3, 12
    gen = torch.Generator().manual_seed(5)
    img = torch.empty(B, 3, H, W, device=DEV)
    anc = b200det.Anchors()(img)
    A = anc.shape[1]
    z = (torch.randn((B, A, C), generator=gen) * 1.5 - 4.0)
    reg = torch.randn((B, A, 4), generator=gen) * 0.5
    thr_z = float(np.log(0.05 / 0.95))
    z[0, :400, 2] = thr_z + torch.linspace(-2e-6, 2e-6, 400)     # rows sitting on the threshold
    z[1, 10, 3], z[1, 10, 7] = 9.0, 9.0001                        # plateau of the float32 sigmoid
    z[1, 11, 5], z[1, 11, 1] = 30.0, 20.0                         # saturation: both are exactly 1.0
    z[2, 12, 4], z[2, 12, 9] = 2.0, 2.0                           # equal logits
    zd, rd = z.to(DEV), reg.to(DEV)
    p = torch.sigmoid(zd)
    assert float(p[1, 10, 3]) == float(p[1, 10, 7]) and float(p[1, 11, 1]) == 1.0
    a = b200det.detect(zd, rd, anc, (H, W), from_logits=True)
    w = b200det.detect(p, rd, anc, (H, W))
    assert a.survivors.tolist() == w.survivors.tolist() and a.count.tolist() == w.count.tolist()
    assert 0 < int(w.survivors[0]) and int(w.survivors[0]) < A
    for b in range(B):
        k = int(w.count[b])
        assert torch.equal(a.anchor_ids[b, :k], w.anchor_ids[b, :k])
        assert torch.equal(a.scores[b, :k], w.scores[b, :k])
        assert torch.equal(a.labels[b, :k], w.labels[b, :k])
        assert torch.equal(a.boxes[b, :k], w.boxes[b, :k])
    # the crafted rows survive with the label torch.max over the probabilities gives
    for (b, row, lab) in ((1, 10, 3), (1, 11, 1), (2, 12, 4)):
        pos = (w.anchor_ids[b, :int(w.count[b])] == row).nonzero()
        if pos.numel():
            assert int(a.labels[b, int(pos[0])]) == lab == int(p[b, row].argmax())
    # and against the oracle on the CPU copy of those probabilities
    _, dets = R.detect_batch(p.cpu(), reg, anc.cpu(), H, W)
    for b, d in enumerate(dets):
        k = int(a.count[b])
        assert np.array_equal(a.anchor_ids[b, :k].cpu().numpy(), d.keep_anchor.numpy())
        assert np.array_equal(a.labels[b, :k].cpu().numpy(), d.labels.numpy())


def test_strip_head_sigmoid_and_head_fixture():
    """The reference head's logits (golden, captured in front of output_act) through the fused path
    equal the probability path on the reference's own head output."""
    g = np.load(os.path.join(GOLDEN, "head_retina.npz"))
    levels = [torch.from_numpy(g[f"logits_l{i}"]) for i in range(3)]
    z = R.head_logits(levels, int(g["num_anchors"]), int(g["num_classes"])).to(DEV)
    probs = torch.from_numpy(g["probs"]).to(DEV)
    B, A, C = probs.shape
    gen = torch.Generator().manual_seed(1)
    anc = torch.rand((A, 2), generator=gen) * 80
    anc = torch.cat([anc, anc + 8 + torch.rand((A, 2), generator=gen) * 40], 1)[None].to(DEV)
    ann = b200det.synth.make_annotations(B, 4, C, 128, 128, gen, gt_min=8.0, gt_max=48.0).to(DEV)
    reg = (torch.randn((B, A, 4), generator=gen) * 0.3).to(DEV)
    l1 = b200det.FocalLoss(from_logits=True)(z, reg, anc, ann)
    l0 = b200det.FocalLoss()(probs, reg, anc, ann)
    np.testing.assert_allclose(l1[0].item(), l0[0].item(), rtol=1e-5)
    np.testing.assert_allclose(l1[1].item(), l0[1].item(), rtol=1e-6)

    class Head(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.output_act = torch.nn.Sigmoid()
    m = torch.nn.Sequential(Head())
    assert b200det.strip_head_sigmoid(m) and isinstance(m[0].output_act, torch.nn.Identity)


# ---------------------------------------------------------------------------------- per-level NCHW head outputs
def _make_levels(spec, B, C, gen, mean=-3.0, std=2.0):
    K = spec.cell_anchors
    return [(torch.randn((B, K * C, spec.fh[l], spec.fw[l]), generator=gen) * std + mean).contiguous()
            for l in range(spec.num_levels)]


@pytest.mark.parametrize("from_logits", [True, False], ids=["logits", "probs"])
@pytest.mark.parametrize("shape", [(2, 5, 96, 128), (3, 90, 64, 64), (2, 1, 80, 112), (1, 7, 50, 34)],
                         ids=lambda s: "B%d_C%d_%dx%d" % s)
def test_focal_loss_on_head_levels(shape, from_logits):
    """FocalLoss fed with the head's raw per-level NCHW tensors (no permute / view / cat, and with
    from_logits no sigmoid either) against (a) the oracle on the reference's own layout + activation
    (oracle head restatement, pinned in test_oracle_golden) and (b) the [B,A,C] path of this library,
    gradients compared level by level in NCHW."""
    B, C, H, W = shape
    img = torch.empty(B, 3, H, W, device=DEV)
    anc = b200det.Anchors()(img)
    spec = b200det.spec_of(anc)
    K, A = spec.cell_anchors, spec.total
    gen = torch.Generator().manual_seed(B * 1000 + C * 7 + H)
    levels = _make_levels(spec, B, C, gen)
    if not from_logits:
        levels = [torch.sigmoid(t) for t in levels]
    ann = b200det.synth.make_annotations(B, 6, C, H, W, gen, gt_min=8.0, gt_max=60.0)
    reg = torch.randn((B, A, 4), generator=gen) * 0.5
    anc_cpu = torch.from_numpy(R.retina_anchors(H, W))[None]

    # (a) oracle: layout (+ activation) as the reference head does it, then the reference loss
    lv0 = [t.clone().requires_grad_(True) for t in levels]
    r0 = reg.clone().requires_grad_(True)
    flat = R.head_logits(lv0, K, C)
    want = R.focal_loss(torch.sigmoid(flat) if from_logits else flat, r0, anc_cpu, ann)
    (want.cls_mean + want.reg_mean).sum().backward()

    lv1 = [t.to(DEV).requires_grad_(True) for t in levels]
    r1 = reg.to(DEV).requires_grad_(True)
    mod = b200det.FocalLoss(from_logits=from_logits)
    lc, lr = mod(lv1, r1, anc, ann.to(DEV))
    (lc + lr).sum().backward()
    np.testing.assert_allclose(lc.item(), want.cls_mean.item(), rtol=1e-5)
    np.testing.assert_allclose(lr.item(), want.reg_mean.item(), rtol=1e-5)
    assert mod.last_num_positives.cpu().tolist() == want.num_pos
    for l in range(spec.num_levels):
        assert lv1[l].grad.shape == levels[l].shape
        np.testing.assert_allclose(lv1[l].grad.cpu().numpy(), lv0[l].grad.numpy(), rtol=2e-5, atol=1e-10)
    want_dreg = r0.grad.numpy() if r0.grad is not None else np.zeros_like(reg.numpy())
    np.testing.assert_allclose(r1.grad.cpu().numpy(), want_dreg, rtol=1e-5, atol=1e-9)

    # (b) the [B,A,C] path on the tensor torch's permute/view/cat would have produced
    with torch.no_grad():
        flat_gpu = R.head_logits([t.to(DEV) for t in levels], K, C)
        lc2, lr2 = b200det.FocalLoss(from_logits=from_logits)(flat_gpu, reg.to(DEV), anc, ann.to(DEV))
        lc3, lr3 = mod([t.detach() for t in lv1], reg.to(DEV), anc, ann.to(DEV))
    np.testing.assert_allclose(lc3.item(), lc2.item(), rtol=2e-6)
    np.testing.assert_allclose(lr3.item(), lr2.item(), rtol=1e-6)


@pytest.mark.parametrize("from_logits", [True, False], ids=["logits", "probs"])
@pytest.mark.parametrize("shape,mean", [((3, 11, 160, 224), -4.0), ((8, 8, 512, 512), -3.2)], ids=["small", "many-tiles"])
def test_detect_on_head_levels(shape, mean, from_logits):
    """detect() on the per-level NCHW tensors == detect() on the [B,A,C] tensor: identical survivors,
    scores, labels, boxes and keep order (bit for bit), including first-index ties between classes.
    "many-tiles": every CTA of the persistent walk takes several tiles and its survivor queue is
    flushed more than once (about a third of the rows pass the threshold)."""
    B, C, H, W = shape
    img = torch.empty(B, 3, H, W, device=DEV)
    anc = b200det.Anchors()(img)
    spec = b200det.spec_of(anc)
    K, A = spec.cell_anchors, spec.total
    gen = torch.Generator().manual_seed(77)
    levels = _make_levels(spec, B, C, gen, mean=mean, std=1.5)
    levels[0][1, 3 * C + 2, 5, 7] = levels[0][1, 3 * C + 6, 5, 7] = 3.0      # two classes tie: the first wins
    levels[1][0, 0 * C + 9, 1, 1], levels[1][0, 0 * C + 4, 1, 1] = 30.0, 20.0  # saturated logits tie at 1.0
    if not from_logits:
        levels = [torch.sigmoid(t) for t in levels]
    reg = (torch.randn((B, A, 4), generator=gen) * 0.5).to(DEV)
    lv = [t.to(DEV) for t in levels]
    flat = R.head_logits(lv, K, C)
    a = b200det.detect(lv, reg, anc, (H, W), from_logits=from_logits)
    w = b200det.detect(flat, reg, anc, (H, W), from_logits=from_logits)
    assert a.survivors.tolist() == w.survivors.tolist() and a.count.tolist() == w.count.tolist()
    assert all(0 < int(x) < A for x in w.survivors)
    for b in range(B):
        k = int(w.count[b])
        assert torch.equal(a.anchor_ids[b, :k], w.anchor_ids[b, :k])
        assert torch.equal(a.scores[b, :k], w.scores[b, :k])
        assert torch.equal(a.labels[b, :k], w.labels[b, :k])
        assert torch.equal(a.boxes[b, :k], w.boxes[b, :k])
        assert torch.equal(a.survivor_rank[b, :k], w.survivor_rank[b, :k])
    if B > 3:
        return                      # the oracle's NMS over tens of thousands of survivors takes minutes
    probs = torch.sigmoid(flat) if from_logits else flat
    _, dets = R.detect_batch(probs.cpu(), reg.cpu(), anc.cpu(), H, W)
    for b, d in enumerate(dets):
        k = int(a.count[b])
        assert np.array_equal(a.anchor_ids[b, :k].cpu().numpy(), d.keep_anchor.numpy())
        assert np.array_equal(a.labels[b, :k].cpu().numpy(), d.labels.numpy())


def test_detect_class_aware_on_head_levels_and_options():
    """Per-class suppression on the head's level tensors == on the flat tensor (logits), keeps at least as many
    boxes as the class-agnostic form, and composes with max_detections / pre_nms_topk."""
    B, C, H, W = 3, 11, 160, 224
    img = torch.empty(B, 3, H, W, device=DEV)
    anc = b200det.Anchors()(img)
    spec = b200det.spec_of(anc)
    K, A = spec.cell_anchors, spec.total
    gen = torch.Generator().manual_seed(78)
    lv = [t.to(DEV) for t in _make_levels(spec, B, C, gen, mean=-3.5, std=1.5)]
    reg = (torch.randn((B, A, 4), generator=gen) * 0.5).to(DEV)
    flat = R.head_logits(lv, K, C)
    a = b200det.detect(lv, reg, anc, (H, W), from_logits=True, class_aware=True)
    w = b200det.detect(flat, reg, anc, (H, W), from_logits=True, class_aware=True)
    g = b200det.detect(flat, reg, anc, (H, W), from_logits=True)
    assert a.count.tolist() == w.count.tolist()
    assert all(int(x) >= int(y) for x, y in zip(w.count, g.count)) and int(w.count.sum()) > int(g.count.sum())
    for b in range(B):
        k = int(w.count[b])
        assert torch.equal(a.anchor_ids[b, :k], w.anchor_ids[b, :k]) and torch.equal(a.labels[b, :k], w.labels[b, :k])
        # oracle on the device's own survivors: boxes / scores / labels by anchor id
        probs = torch.sigmoid(flat[b]).cpu()
        sc, lb = probs.max(dim=1)
        surv = torch.nonzero(sc > 0.05).flatten()
        boxes = b200det.ClipBoxes()(b200det.BBoxTransform()(anc, reg[b:b + 1]), img[b:b + 1])[0].cpu()
        keep = R.batched_nms(boxes[surv].numpy(), sc[surv].numpy(), lb[surv].numpy(), 0.5)
        assert np.array_equal(w.anchor_ids[b, :k].cpu().numpy(), surv[keep].numpy())
    m = b200det.detect(flat, reg, anc, (H, W), from_logits=True, class_aware=True, max_detections=50)
    t = b200det.detect(flat, reg, anc, (H, W), from_logits=True, class_aware=True, pre_nms_topk=300)
    for b in range(B):
        assert int(m.count[b]) == min(50, int(w.count[b]))
        assert torch.equal(m.anchor_ids[b, :int(m.count[b])], w.anchor_ids[b, :int(m.count[b])])
        # the top-300 by score, then per-class NMS among them: a prefix-closed subset of the full answer's candidates
        assert int(t.count[b]) <= 300 and int(t.survivors[b]) == int(w.survivors[b])


def test_head_levels_reject_bad_shapes():
    img = torch.empty(2, 3, 64, 64, device=DEV)
    anc = b200det.Anchors()(img)
    spec = b200det.spec_of(anc)
    gen = torch.Generator().manual_seed(0)
    lv = [t.to(DEV) for t in _make_levels(spec, 2, 4, gen)]
    reg = torch.zeros((2, spec.total, 4), device=DEV)
    ann = -torch.ones((2, 1, 5), device=DEV)
    with pytest.raises(ValueError):
        b200det.FocalLoss()(lv[:-1], reg, anc, ann)                       # a level is missing
    with pytest.raises(ValueError):
        b200det.FocalLoss()(lv, reg, anc.clone(), ann)                    # a plain tensor carries no level geometry
    bad = list(lv)
    bad[0] = bad[0][:, :, :-1].contiguous()
    with pytest.raises(ValueError):
        b200det.detect(bad, reg, anc, (64, 64))


# ------------------------------------------------------------------ evaluator tails on device
def _fixture_detections(g):
    return b200det.BatchDetections(torch.from_numpy(g["counts"]).long(), torch.from_numpy(g["counts"]).long(),
                                   torch.from_numpy(g["scores"]).to(DEV), torch.from_numpy(g["labels"]).to(DEV),
                                   torch.from_numpy(g["boxes"]).to(DEV), None, None)


def test_eval_pack_matches_reference_fixture():
    """k7_eval_pack == the rows the reference's evaluate_coco / coco_evaluate_ksevendata / _get_detections
    built from the same detections (tests/golden/eval_tail.npz), bit for bit, and the host converters
    rebuild the reference's Python objects."""
    g = np.load(os.path.join(GOLDEN, "eval_tail.npz"))
    det = _fixture_detections(g)
    ids, cat = g["image_ids"].tolist(), g["category_of_label"].tolist()
    for tag, thr, valid in (("coco_t005", 0.05, None), ("coco_t030", 0.3, None), ("k7_t005_all", 0.05, None),
                            ("k7_t030_valid", 0.3, g["k7_valid"].tolist())):
        p = b200det.eval_pack(det, g["scales"], mode="coco", image_ids=ids, category_of_label=cat, valid_class_id=valid,
                              threshold=thr)
        assert np.array_equal(p.rows.cpu().numpy(), g[tag]), tag
        res = b200det.coco_results(p)
        assert len(res) == len(g[tag])
        if len(res):
            assert res[0] == {"image_id": int(g[tag][0, 0]), "category_id": int(g[tag][0, 6]), "score": float(g[tag][0, 5]),
                              "bbox": [float(v) for v in g[tag][0, 1:5]]}
    nc = len(cat)
    for tag, thr, md in (("csv_t005_m100", 0.05, 100), ("csv_t030_m10", 0.3, 10)):
        p = b200det.eval_pack(det, g["scales"], mode="csv", num_classes=nc, threshold=thr, max_detections=md)
        assert np.array_equal(p.class_offsets.cpu().numpy(), g[tag + "_class_offsets"]), tag
        assert np.array_equal(p.rows.cpu().numpy(), g[tag]), tag
        alld = b200det.csv_all_detections(p)
        off = g[tag + "_class_offsets"]
        assert len(alld) == len(g["counts"]) and all(len(a) == nc for a in alld)
        assert np.array_equal(alld[2][3], g[tag][off[2 * nc + 3]:off[2 * nc + 4], 1:6])


@pytest.mark.parametrize("B,cap,nc,max_det", [(1, 1, 1, 100), (37, 300, 80, 100), (5, 1500, 3, 0), (300, 64, 90, 7)])
def test_eval_pack_vs_oracle_random(B, cap, nc, max_det):
    """Random batches (empty images, counts above one block, scores straddling the threshold,
    duplicated scores) against the oracle, both modes, bit for bit."""
    rng = np.random.RandomState(B * 1000 + cap)
    counts = rng.randint(0, cap + 1, size=B).astype(np.int32)
    counts[rng.randint(0, B)] = 0
    counts[rng.randint(0, B)] = cap
    scores = np.zeros((B, cap), np.float32)
    for b in range(B):
        sc = np.round(rng.uniform(0.0, 1.0, size=counts[b]), 2).astype(np.float32)   # many ties
        scores[b, :counts[b]] = np.sort(sc)[::-1]
    labels = rng.randint(0, nc, size=(B, cap)).astype(np.int64)
    xy = rng.uniform(0, 900, size=(B, cap, 2)).astype(np.float32)
    boxes = np.concatenate([xy, xy + rng.uniform(0, 300, size=(B, cap, 2)).astype(np.float32)], axis=2)
    scales = rng.uniform(0.3, 2.5, size=B)
    ids = rng.randint(1, 600000, size=B).tolist()
    cat = rng.randint(1, 91, size=nc).tolist()
    valid = sorted(set(rng.randint(0, nc, size=max(1, nc // 2)).tolist()))
    det = b200det.BatchDetections(torch.from_numpy(counts).long(), torch.from_numpy(counts).long(),
                                  torch.from_numpy(scores).to(DEV), torch.from_numpy(labels).to(DEV),
                                  torch.from_numpy(boxes).to(DEV), None, None)
    for thr, v in ((0.05, None), (0.5, valid)):
        p = b200det.eval_pack(det, scales, mode="coco", image_ids=ids, category_of_label=cat, valid_class_id=v, threshold=thr)
        want = R.eval_rows_coco(scores, labels, boxes, counts, scales, ids, cat, threshold=thr, valid_class_id=v)
        assert np.array_equal(p.rows.cpu().numpy(), want)
        off = p.row_offsets.cpu().numpy()
        assert off[0] == 0 and off[-1] == len(want) and np.all(np.diff(off) >= 0)
    for thr in (0.05, 0.5):
        p = b200det.eval_pack(det, scales, mode="csv", num_classes=nc, threshold=thr, max_detections=max_det)
        want, woff = R.eval_rows_csv(scores, labels, boxes, counts, scales, nc, score_threshold=thr, max_detections=max_det)
        assert np.array_equal(p.class_offsets.cpu().numpy(), woff)
        assert np.array_equal(p.rows.cpu().numpy(), want)


def test_eval_pack_after_detect_end_to_end():
    """detect() -> eval_pack() on device == the reference's per-image host loop on the oracle's detections."""
    B, C, H, W = 4, 6, 128, 160
    img = torch.empty(B, 3, H, W, device=DEV)
    anc = b200det.Anchors()(img)
    gen = torch.Generator().manual_seed(5)
    cls = torch.sigmoid(torch.randn((B, anc.shape[1], C), generator=gen) * 1.5 - 4.0).to(DEV)
    reg = (torch.randn((B, anc.shape[1], 4), generator=gen) * 0.5).to(DEV)
    det = b200det.detect(cls, reg, anc, (H, W))
    scales = [0.8, 1.25, 0.5, 1.0]
    p = b200det.eval_pack(det, scales, mode="coco", image_ids=[11, 12, 13, 14], category_of_label=list(range(1, C + 1)))
    _, dets = R.detect_batch(cls.cpu(), reg.cpu(), anc.cpu(), H, W)
    cap = max(1, max(len(d.scores) for d in dets))
    sc = np.zeros((B, cap), np.float32)
    lb = np.zeros((B, cap), np.int64)
    bx = np.zeros((B, cap, 4), np.float32)
    for b, d in enumerate(dets):
        k = len(d.scores)
        sc[b, :k], lb[b, :k], bx[b, :k] = d.scores.numpy(), d.labels.numpy(), d.boxes.numpy()
    want = R.eval_rows_coco(sc, lb, bx, [len(d.scores) for d in dets], scales, [11, 12, 13, 14], list(range(1, C + 1)))
    got = p.rows.cpu().numpy()
    assert got.shape == want.shape and got.shape[0] > 0
    assert np.array_equal(got[:, [0, 6]], want[:, [0, 6]])
    np.testing.assert_allclose(got[:, 1:6], want[:, 1:6], rtol=1e-5, atol=1e-3)


def test_eval_pack_rejects_bad_arguments():
    g = np.load(os.path.join(GOLDEN, "eval_tail.npz"))
    det = _fixture_detections(g)
    with pytest.raises(ValueError):
        b200det.eval_pack(det, g["scales"][:2], num_classes=7)
    with pytest.raises(ValueError):
        b200det.eval_pack(det, g["scales"], mode="voc", num_classes=7)
    with pytest.raises(ValueError):
        b200det.eval_pack(det, g["scales"])
    with pytest.raises(ValueError):
        b200det.eval_pack(det, g["scales"], mode="csv", num_classes=10000)


# ------------------------------------------------------------------ annotation packing
def test_collate_annotations_fixture_and_random():
    """collate_annotations (ragged block + k7_annotations_pad) == the reference collater's annot_padded
    (collate.npz) and == the oracle on random ragged batches; the ragged views round-trip."""
    g = np.load(os.path.join(GOLDEN, "collate.npz"))
    for tag in ("mixed", "all_empty", "single"):
        rows = [g[f"{tag}_rows{i}"] for i in range(int(g[f"{tag}_n"]))]
        out = b200det.collate_annotations(rows, DEV)
        assert out.dtype == torch.float32 and tuple(out.shape) == g[f"{tag}_padded"].shape, tag
        assert np.array_equal(out.cpu().numpy(), g[f"{tag}_padded"]), tag
    rng = np.random.RandomState(3)
    for B, nmax in ((1, 1), (17, 130), (256, 40)):
        rows = [np.concatenate([rng.uniform(0, 900, (n, 4)), rng.randint(0, 90, (n, 1))], axis=1).astype(np.float32)
                for n in rng.randint(0, nmax + 1, size=B)]
        out, ragged, off = b200det.collate_annotations([torch.from_numpy(r) for r in rows], DEV, return_ragged=True)
        assert np.array_equal(out.cpu().numpy(), R.collate_annotations(rows).numpy())
        assert off.cpu().tolist() == np.concatenate([[0], np.cumsum([len(r) for r in rows])]).tolist()
        if ragged.numel():
            assert np.array_equal(ragged.cpu().numpy(), np.concatenate(rows, axis=0))
    with pytest.raises(RuntimeError):
        b200det.collate_annotations(rows, "cpu")


# ---------------------------------------------------------------------------------- boxes: the reference ON CUDA
# On a GPU box the reference itself runs its box arithmetic on CUDA (architectures/utils.py:26-71 after .cuda(),
# retinanet.py:297-298).  The same restatement on device="cuda" tensors is therefore the exact target: every ATen
# kernel rounds once per operation and uses CUDA's expf, as k7_decode does (-fmad=false, explicit __f*_rn) - BIT
# equality, no tolerance.  The comparison with the torch-CPU fixtures keeps a small absolute slack (host libm expf
# differs from CUDA's by an ulp and x1 = cx - w/2 cancels); the worst deviation seen is part of the message.
def _cuda_reference_boxes(anc, reg, yx, H, W):
    a = anc if anc.dim() == 3 else anc[None]
    boxes = R.decode_boxes_yx(a, reg) if yx else R.decode_boxes(a, reg)
    return R.clip_boxes_(boxes.contiguous(), H, W)


def _worst(got, want):
    d = np.abs(got.astype(np.float64) - want.astype(np.float64))
    rel = d / np.maximum(np.abs(want.astype(np.float64)), 1e-30)
    return float(d.max()) if d.size else 0.0, float(rel[np.abs(want) > 1.0].max()) if (np.abs(want) > 1.0).any() else 0.0


@pytest.mark.parametrize("path", DET, ids=DIDS)
def test_boxes_bit_equal_to_the_reference_on_cuda(path):
    g, yx, anc, img = load_det_case(path, True)
    H, W = int(g["height"]), int(g["width"])
    reg = torch.from_numpy(g["regressions"]).to(DEV)
    cls = torch.from_numpy(g["classifications"]).to(DEV)
    anc_t = torch.from_numpy(g["anchors"]).to(DEV)
    want = _cuda_reference_boxes(anc_t, reg, yx, H, W)
    rb = b200det.BBoxTransformYX() if yx else b200det.BBoxTransform()
    got = b200det.ClipBoxes()(rb(anc, reg), img)
    assert torch.equal(got, want), "decode+clip differs from the reference's own ops on CUDA: worst abs %g rel %g" % _worst(
        got.cpu().numpy(), want.cpu().numpy())
    # the fused tail returns those very boxes for the anchors it keeps
    det = b200det.detect(cls, reg, anc, (H, W), score_thr=float(g["score_thr"]), iou_thr=float(g["iou_thr"]),
                         variant=L.DECODE_EFFDET if yx else L.DECODE_RETINA)
    for b in range(cls.shape[0]):
        k = int(det.count[b])
        ids = det.anchor_ids[b, :k].long()
        assert torch.equal(det.boxes[b, :k], want[b, ids])
    # against the torch-CPU fixture: the deviation is host expf vs CUDA expf, recorded here
    worst_abs, worst_rel = _worst(got.cpu().numpy(), g["boxes"])
    assert worst_abs <= 1.3e-4, f"worst deviation from the CPU fixture: abs {worst_abs:g} px, rel {worst_rel:g}"
    print(f"[{os.path.basename(path)}] vs torch-CPU fixture: worst abs {worst_abs:.3g} px, worst rel (|x|>1) {worst_rel:.3g}")


@pytest.mark.parametrize("name", ["c1_retinanet_r50_512", "c3_ksevendet_thermal_512x640", "c4_effdet_d3_896", "c5_effdet_d7_1536"])
def test_boxes_bit_equal_on_baseline_shapes(name):
    """The BASELINE shapes (two images each): decode + clip bit-equal to the reference's ops on CUDA, for the anchors of
    a recipe and for large deltas (dw, dh up to +-3: widths from 5 % to 20 x the anchor)."""
    wl = b200det.synth.WORKLOADS[name]
    B = 2
    gen = torch.Generator(device=DEV).manual_seed(11)
    img = torch.empty(B, 3, wl.height, wl.width, device=DEV)
    anc = b200det.Anchors(**wl.anchor_kwargs())(img)
    A = anc.shape[1]
    reg = torch.randn((B, A, 4), generator=gen, device=DEV) * torch.tensor([2.0, 2.0, 1.5, 1.5], device=DEV)
    want = _cuda_reference_boxes(anc.clone(), reg, False, wl.height, wl.width)
    got = b200det.ClipBoxes()(b200det.BBoxTransform()(anc, reg), img)
    assert torch.equal(got, want), "worst abs %g rel %g" % _worst(got.cpu().numpy(), want.cpu().numpy())


# ---------------------------------------------------------------------------------- streams and graphs
def _small_case(seed, B=6, C=7, H=192, W=256, M=9):
    gen = torch.Generator().manual_seed(seed)
    img = torch.empty(B, 3, H, W, device=DEV)
    anc = b200det.Anchors()(img)
    A = anc.shape[1]
    ann = b200det.synth.make_annotations(B, M, C, H, W, gen, gt_min=8.0, gt_max=90.0).to(DEV)
    cls, reg = b200det.synth.make_head_outputs(B, A, C, gen, logit_mean=-3.5, logit_std=1.5)
    return img, anc, ann, cls.to(DEV), reg.to(DEV), (H, W)


def test_two_streams_do_not_share_workspaces():
    """Two callers overlapped on two streams of one device (decode/NMS of two batches; a FocalLoss next to a detect):
    workspaces are per (device, stream), so the results are those of the calls run one after the other."""
    img, anc, ann, cls1, reg1, hw = _small_case(1)
    _, _, _, cls2, reg2, _ = _small_case(2)
    want1 = b200det.detect(cls1, reg1, anc, hw)
    want2 = b200det.detect(cls2, reg2, anc, hw)
    loss = b200det.FocalLoss()
    with torch.no_grad():
        lw = [t.clone() for t in loss(cls1, reg1, anc, ann)]
    torch.cuda.synchronize()
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    for _ in range(20):                                  # many rounds: any sharing would show as a flipped keep
        with torch.cuda.stream(s1):
            got1 = b200det.detect(cls1, reg1, anc, hw, sync=False, want_rank=True)
            with torch.no_grad():
                lg = loss(cls1, reg1, anc, ann)
        with torch.cuda.stream(s2):
            got2 = b200det.detect(cls2, reg2, anc, hw, sync=False, want_rank=True)
        torch.cuda.synchronize()
        for got, want in ((got1, want1), (got2, want2)):
            assert got.count.cpu().tolist() == want.count.tolist()
            for b in range(cls1.shape[0]):
                k = int(want.count[b])
                assert torch.equal(got.anchor_ids[b, :k], want.anchor_ids[b, :k])
                assert torch.equal(got.boxes[b, :k], want.boxes[b, :k])
        assert torch.equal(lg[0], lw[0]) and torch.equal(lg[1], lw[1])


def test_graphed_step_replays_with_new_inputs():
    """b200det.GraphedStep: FocalLoss forward+backward and detect (on a second stream) captured once, replayed on
    inputs copied into the static tensors; losses, gradients and detections equal the eager calls bit for bit."""
    img, anc, ann, cls, reg, hw = _small_case(3)
    _, _, ann2, cls2, reg2, _ = _small_case(4)
    loss = b200det.FocalLoss()
    c = cls.clone().requires_grad_(True)
    r = reg.clone().requires_grad_(True)
    a = ann.clone()
    side = torch.cuda.Stream()
    out = {}

    def step():
        cur = torch.cuda.current_stream()
        side.wait_stream(cur)
        with torch.cuda.stream(side):
            out["det"] = b200det.detect(c.detach(), r.detach(), anc, hw, sync=False, want_rank=True)
        c.grad = None
        r.grad = None
        lc, lr = loss(c, r, anc, a)
        (lc + lr).sum().backward()
        out["loss"] = (lc.detach(), lr.detach())
        cur.wait_stream(side)

    g = b200det.GraphedStep(step)
    assert g.launches >= 6
    for cc, rr, aa in ((cls, reg, ann), (cls2, reg2, ann2), (cls, reg, ann)):
        with torch.no_grad():
            c.copy_(cc)
            r.copy_(rr)
            a.copy_(aa)
        g()
        torch.cuda.synchronize()
        got_det, got_loss, got_gc, got_gr = out["det"], out["loss"], c.grad.clone(), r.grad.clone()
        c0, r0 = cc.clone().requires_grad_(True), rr.clone().requires_grad_(True)
        lc, lr = b200det.FocalLoss()(c0, r0, anc, aa)
        (lc + lr).sum().backward()
        want = b200det.detect(cc, rr, anc, hw)
        assert torch.equal(got_loss[0], lc.detach()) and torch.equal(got_loss[1], lr.detach())
        assert torch.equal(got_gc, c0.grad) and torch.equal(got_gr, r0.grad)
        assert got_det.count.cpu().tolist() == want.count.tolist()
        for b in range(cc.shape[0]):
            k = int(want.count[b])
            assert torch.equal(got_det.anchor_ids[b, :k], want.anchor_ids[b, :k])
            assert torch.equal(got_det.scores[b, :k], want.scores[b, :k])


def test_spare_sms_changes_no_result():
    """FocalLoss(spare_sms=n): the streaming kernel runs on fewer SMs (room for another stream's NMS kernels);
    gradients stay bit-identical, the losses equal to the last bit or two (the stream's partial sums are grouped
    by CTA before they enter the integer accumulators)."""
    img, anc, ann, cls, reg, hw = _small_case(5, B=5, C=11)
    outs = []
    for spare in (0, 24, 100):
        c, r = cls.clone().requires_grad_(True), reg.clone().requires_grad_(True)
        lc, lr = b200det.FocalLoss(spare_sms=spare)(c, r, anc, ann)
        (lc + lr).sum().backward()
        outs.append((lc.detach(), lr.detach(), c.grad, r.grad))
    for o in outs[1:]:
        assert torch.equal(o[2], outs[0][2]) and torch.equal(o[3], outs[0][3])
        np.testing.assert_allclose(o[0].item(), outs[0][0].item(), rtol=1e-6)
        np.testing.assert_allclose(o[1].item(), outs[0][1].item(), rtol=1e-6)


def test_start_is_forward_in_two_phases():
    """FocalLoss.start() -> PendingLoss on one GPU (nothing to defer): losses and gradients are those of forward() +
    backward(), for leaf inputs (the gradient buffer itself becomes .grad) and for non-leaf inputs (autograd)."""
    img, anc, ann, cls, reg, hw = _small_case(6, B=4, C=9)
    c0, r0 = cls.clone().requires_grad_(True), reg.clone().requires_grad_(True)
    lc0, lr0 = b200det.FocalLoss()(c0, r0, anc, ann)
    (lc0 + lr0).sum().backward()
    c1, r1 = cls.clone().requires_grad_(True), reg.clone().requires_grad_(True)
    loss = b200det.FocalLoss()
    pend = loss.start(c1, r1, anc, ann)
    pend.backward()
    lc1, lr1 = pend.losses()
    assert torch.equal(lc1, lc0.detach()) and torch.equal(lr1, lr0.detach())
    assert torch.equal(c1.grad, c0.grad) and torch.equal(r1.grad, r0.grad)
    assert torch.equal(pend.num_positives, loss.last_num_positives)
    # non-leaf inputs: the gradients flow on through autograd (here: through a scaling of the leaves)
    c2, r2 = cls.clone().requires_grad_(True), reg.clone().requires_grad_(True)
    pend = loss.start(c2 * 1.0, r2 * 2.0, anc, ann)
    pend.backward(scale=0.5)
    c3, r3 = cls.clone().requires_grad_(True), reg.clone().requires_grad_(True)
    lc3, lr3 = b200det.FocalLoss()(c3 * 1.0, r3 * 2.0, anc, ann)
    (0.5 * (lc3 + lr3)).sum().backward()
    np.testing.assert_allclose(c2.grad.cpu().numpy(), c3.grad.cpu().numpy(), rtol=1e-6, atol=1e-12)
    np.testing.assert_allclose(r2.grad.cpu().numpy(), r3.grad.cpu().numpy(), rtol=1e-6, atol=1e-12)


@pytest.mark.parametrize("opts", [dict(), dict(class_aware=True), dict(max_detections=37), dict(pre_nms_topk=500),
                                  dict(class_aware=True, max_detections=5, pre_nms_topk=900)],
                         ids=["plain", "per-class", "maxdet", "topk", "all"])
def test_both_nms_chains_agree(opts):
    """K7DetectParams.nms_chain: the per-image chain and the multi-SM chain give the same detections, bit for bit -
    dense images (thousands of survivors, clustered boxes), an empty image, options."""
    B, C, H, W = 5, 6, 384, 512
    gen = torch.Generator().manual_seed(21)
    img = torch.empty(B, 3, H, W, device=DEV)
    anc = b200det.Anchors()(img)
    A = anc.shape[1]
    cls, reg = b200det.synth.make_head_outputs(B, A, C, gen, logit_mean=-2.2, logit_std=1.3)
    reg = reg * 0.3                                          # neighbours overlap a lot: long suppression chains
    cls[3] = 0.001                                           # an image without survivors
    cls, reg = cls.to(DEV), reg.to(DEV)
    a = b200det.detect(cls, reg, anc, (H, W), nms_chain="per_image", **opts)
    b = b200det.detect(cls, reg, anc, (H, W), nms_chain="multi_sm", **opts)
    assert a.count.tolist() == b.count.tolist() and a.survivors.tolist() == b.survivors.tolist()
    assert int(a.survivors.max()) > 3000 and int(a.count[3]) == 0
    for i in range(B):
        k = int(a.count[i])
        for x, y in ((a.anchor_ids, b.anchor_ids), (a.scores, b.scores), (a.labels, b.labels), (a.boxes, b.boxes),
                     (a.survivor_rank, b.survivor_rank)):
            assert torch.equal(x[i, :k], y[i, :k])
    with pytest.raises(ValueError):
        b200det.detect(cls, reg, anc, (H, W), nms_chain="fastest")
