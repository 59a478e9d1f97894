"""CPU checks of the scalar arithmetic the CUDA kernels use (csrc/k7_math.cuh compiled for the
host by tests/hostmath) against the oracle and the golden fixtures.  This validates result BITS
(IoU, assignment, NMS order) and loss accuracy without a GPU; the kernels themselves are
checked by the -m gpu tests."""
import ctypes as C
import glob
import os
import subprocess

import numpy as np
import pytest
import torch

from oracle import restate as R

HERE = os.path.dirname(os.path.abspath(__file__))
GOLDEN = os.path.join(HERE, "golden")


@pytest.fixture(scope="session")
def hm(tmp_path_factory):
    out = str(tmp_path_factory.mktemp("hm") / "k7_hostmath.so")
    src = os.path.join(HERE, "hostmath", "k7_hostmath.cpp")
    subprocess.check_call(["g++", "-O2", "-ffp-contract=off", "-shared", "-fPIC", "-o", out, src])
    return C.CDLL(out)


def fp(a):
    return a.ctypes.data_as(C.c_void_p)


LOSS = sorted(glob.glob(os.path.join(GOLDEN, "loss_*.npz")))


@pytest.mark.parametrize("path", LOSS, ids=[os.path.basename(p)[:-4] for p in LOSS])
def test_match_bits(hm, path):
    g = np.load(path)
    yx = str(g["variant"]) == "effdet"
    thr = (0.4, 0.5) if yx else (0.35, 0.45)
    anc = g["anchors"][0]
    anc_xyxy = np.ascontiguousarray(anc[:, [1, 0, 3, 2]] if yx else anc)
    A = anc.shape[0]
    for b in range(g["annotations"].shape[0]):
        ann = g["annotations"][b]
        rows = np.nonzero(ann[:, 4] != -1)[0]
        if rows.size == 0:
            continue
        gt = np.ascontiguousarray(ann[rows, :4])
        iou = np.empty(A, np.float32); arg = np.empty(A, np.int32); st = np.empty(A, np.int32)
        hm.hm_match(fp(anc_xyxy), A, fp(gt), len(rows), C.c_float(thr[0]), C.c_float(thr[1]), fp(iou), fp(arg), fp(st))
        assert np.array_equal(iou, g["iou_max"][b])
        assert np.array_equal(arg, g["iou_argmax"][b])
        assert np.array_equal(st, g["state"][b])


def test_focal_terms_vs_torch(hm):
    rng = np.random.default_rng(0)
    x = np.concatenate([
        1.0 / (1.0 + np.exp(-rng.normal(-4.6, 2.0, 200000))),
        rng.uniform(0, 1, 50000), [0.0, 1e-6, 1e-4, 0.25, 0.2500001, 0.5, 0.9999, 0.99995, 1.0],
    ]).astype(np.float32)
    n = x.size
    outs = [np.empty(n, np.float32) for _ in range(4)]
    hm.hm_focal(fp(x), n, C.c_float(0.25), *[fp(o) for o in outs])
    neg, neg_g, pos, pos_g = outs
    # the reference's float32 expressions (losses.py:62,100-112) and their autograd
    for target, val, grad in ((0.0, neg, neg_g), (1.0, pos, pos_g)):
        xt = torch.from_numpy(x).clone().requires_grad_(True)
        p = torch.clamp(xt, 1e-4, 1.0 - 1e-4)
        t = torch.full_like(p, target)
        af = torch.where(t == 1.0, torch.full_like(p, 0.25), 1.0 - torch.full_like(p, 0.25))
        fw = torch.where(t == 1.0, 1.0 - p, p)
        el = af * torch.pow(fw, 2.0) * -(t * torch.log(p) + (1.0 - t) * torch.log(1.0 - p))
        el.sum().backward()
        np.testing.assert_allclose(val, el.detach().numpy(), rtol=2e-6, atol=1e-30)
        np.testing.assert_allclose(grad, xt.grad.numpy(), rtol=1e-5, atol=1e-12)
        # the SUM is what the 1e-5 contract is stated on
        assert abs(val.astype(np.float64).sum() / el.detach().double().sum().item() - 1) < 1e-6


def test_reg_targets_and_smooth_l1(hm):
    rng = np.random.default_rng(1)
    div = np.array([0.1, 0.1, 0.2, 0.2], np.float32)
    for yx, use_div in ((0, True), (1, False)):
        for _ in range(200):
            a = rng.uniform(0, 200, 2).astype(np.float32)
            anc = np.array([a[0], a[1], a[0] + rng.uniform(8, 90), a[1] + rng.uniform(8, 90)], np.float32)
            gt = (anc + rng.normal(0, 6, 4)).astype(np.float32)
            if rng.uniform() < 0.1:
                gt[2] = gt[0] + 0.5        # sub-pixel width: clamp to 1 (losses.py:136)
            reg = rng.normal(0, 1.5, 4).astype(np.float32)
            tg = np.empty(4, np.float32); ls = np.empty(4, np.float32); gr = np.empty(4, np.float32)
            hm.hm_reg(fp(anc), fp(gt), fp(reg), fp(div) if use_div else None, yx, C.c_float(1.0 / 9.0),
                      C.c_float(4.5), C.c_float(0.5 / 9.0), fp(tg), fp(ls), fp(gr))
            A_, G_ = torch.from_numpy(anc), torch.from_numpy(gt)
            aw, ah = A_[2] - A_[0], A_[3] - A_[1]
            acx, acy = A_[0] + 0.5 * aw, A_[1] + 0.5 * ah
            gw, gh = G_[2] - G_[0], G_[3] - G_[1]
            gcx, gcy = G_[0] + 0.5 * gw, G_[1] + 0.5 * gh
            gw, gh = gw.clamp(min=1), gh.clamp(min=1)
            dx, dy = (gcx - acx) / aw, (gcy - acy) / ah
            dw, dh = torch.log(gw / aw), torch.log(gh / ah)
            t = torch.stack((dy, dx, dh, dw) if yx else (dx, dy, dw, dh))
            if use_div:
                t = t / torch.from_numpy(div)
            r = torch.from_numpy(reg).clone().requires_grad_(True)
            d = torch.abs(t - r)
            l = torch.where(d <= 1.0 / 9.0, 0.5 * 9.0 * torch.pow(d, 2), d - 0.5 / 9.0)
            l.sum().backward()
            np.testing.assert_allclose(tg, t.numpy(), rtol=2e-6, atol=1e-7)
            np.testing.assert_allclose(ls, l.detach().numpy(), rtol=1e-5, atol=1e-6)
            np.testing.assert_allclose(gr, r.grad.numpy(), rtol=1e-5, atol=1e-6)


DET = sorted(glob.glob(os.path.join(GOLDEN, "detect_*.npz")))


@pytest.mark.parametrize("path", DET, ids=[os.path.basename(p)[:-4] for p in DET])
def test_decode_and_nms(hm, path):
    g = np.load(path)
    yx = str(g["variant"]) == "effdet"
    anc = g["anchors"][0]
    anc_xyxy = np.ascontiguousarray(anc[:, [1, 0, 3, 2]] if yx else anc)
    mean = np.zeros(4, np.float32); std = np.array([0.1, 0.1, 0.2, 0.2], np.float32)
    B, A = g["regressions"].shape[:2]
    for b in range(B):
        out = np.empty((A, 4), np.float32)
        hm.hm_decode(fp(anc_xyxy), fp(np.ascontiguousarray(g["regressions"][b])), A, 1 if yx else 0, fp(mean), fp(std),
                     1, int(g["height"]), int(g["width"]), fp(out))
        # expf differs between libm and ATen by an ulp: 1e-5 relative (+ cancellation slack in pixels)
        np.testing.assert_allclose(out, g["boxes"][b], rtol=1e-5, atol=2e-4)
        # NMS stage on the REFERENCE's boxes: keep must be bit-exact
        surv = g[f"survivors_{b}"]
        if surv.size == 0:
            continue
        boxes = np.ascontiguousarray(g["boxes"][b][surv])
        scores = np.ascontiguousarray(g["classifications"][b].max(axis=1)[surv])
        keep = np.empty(surv.size, np.int64); keys = np.empty(surv.size, np.uint64)
        k = hm.hm_nms(fp(boxes), fp(scores), surv.size, C.c_float(float(g["iou_thr"])), fp(keep), fp(keys))
        assert np.array_equal(keep[:k], g[f"keep_rank_{b}"])


def test_constants():
    assert np.float32(0.9999) == np.float32(1.0 - 1e-4)
    assert np.float32(4.5) == np.float32(0.5 * 9.0)
