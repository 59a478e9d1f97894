"""CPU tests of the host logic: anchor recipes -> K7AnchorSpec, the C-ABI library exports,
module API surface, fail-loud behaviour without a GPU.  No CUDA compute is executed."""
import ctypes as C
import hashlib
import json
import os
import re

import numpy as np
import pytest
import torch

import b200det
from b200det import _lib as L

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)

with open(os.path.join(HERE, "golden", "anchors.json")) as f:
    ANCHORS = json.load(f)


def module_for(g):
    kw = dict(g["kwargs"])
    if "pyramid_levels" in kw:
        kw["pyramid_levels"] = [int(x) for x in kw["pyramid_levels"]]
    return b200det.AnchorsYX(**kw) if g["kind"] == "yx" else b200det.Anchors(**kw)


def emulate_spec(sp) -> np.ndarray:
    """The device formula of csrc/k7_common.cuh::k7_anchor_from_spec in numpy:
    float32(base + (i + 0.5) * stride) with float64 arithmetic."""
    parts = []
    for l in range(sp.num_levels):
        fh, fw, st = sp.fh[l], sp.fw[l], sp.stride[l]
        base = np.array([[sp.base[l][k][c] for c in range(4)] for k in range(sp.cell_anchors)])
        sx = (np.arange(fw, dtype=np.float64) + 0.5) * st
        sy = (np.arange(fh, dtype=np.float64) + 0.5) * st
        gx, gy = np.meshgrid(sx, sy)
        if sp.box_order == L.BOX_XYXY:
            sh = np.stack([gx.ravel(), gy.ravel(), gx.ravel(), gy.ravel()], 1)
        else:
            sh = np.stack([gy.ravel(), gx.ravel(), gy.ravel(), gx.ravel()], 1)
        parts.append((sh[:, None, :] + base[None]).reshape(-1, 4))
    return np.concatenate(parts).astype(np.float32)


@pytest.mark.parametrize("name", sorted(ANCHORS))
def test_spec_reproduces_reference_anchors(name):
    g = ANCHORS[name]
    sp = module_for(g).spec_for(g["height"], g["width"])
    assert sp.total == g["count"]
    a = emulate_spec(sp)
    assert hashlib.sha256(a.tobytes()).hexdigest() == g["sha256"]


def test_library_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "k7det.h")).read()
    declared = set(re.findall(r"^K7_API\s+[\w\s\*]+?\b(k7_\w+)\s*\(", hdr, flags=re.M))
    assert declared == set(L.SYMBOLS), declared ^ set(L.SYMBOLS)
    lib = C.CDLL(L.LIB_PATH)
    for name in declared:
        assert hasattr(lib, name), name
    assert L.lib().k7_version() == L.ABI_VERSION


def test_struct_layout_matches_header(tmp_path):
    """sizeof of every struct as the C compiler sees include/k7det.h == the ctypes mirror
    (a mismatch would corrupt every call)."""
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    src = tmp_path / "sizes.c"
    src.write_text('#include <stdio.h>\n#include "k7det.h"\nint main(void){printf("%zu %zu %zu %zu %zu %zu %zu %d\\n",'
                   'sizeof(K7AnchorSpec),sizeof(K7MatchParams),sizeof(K7LossParams),sizeof(K7DecodeParams),'
                   'sizeof(K7DetectParams),sizeof(K7EvalParams),sizeof(K7PeerReduce)+1000*sizeof(K7SoftNmsParams),K7_ABI_VERSION);return 0;}\n')
    exe = tmp_path / "sizes"
    subprocess.run(["gcc", "-I", os.path.join(root, "include"), str(src), "-o", str(exe)], check=True)
    got = [int(x) for x in subprocess.run([str(exe)], check=True, capture_output=True, text=True).stdout.split()]
    want = [C.sizeof(L.K7AnchorSpec), C.sizeof(L.K7MatchParams), C.sizeof(L.K7LossParams), C.sizeof(L.K7DecodeParams),
            C.sizeof(L.K7DetectParams), C.sizeof(L.K7EvalParams), C.sizeof(L.K7PeerReduce) + 1000 * C.sizeof(L.K7SoftNmsParams),
            L.ABI_VERSION]
    assert got == want


def test_no_cpu_fallback():
    img = torch.zeros(1, 3, 64, 64)
    with pytest.raises(RuntimeError):
        b200det.Anchors()(img)
    with pytest.raises(RuntimeError):
        b200det.FocalLoss()(torch.rand(1, 774, 2), torch.rand(1, 774, 4), torch.rand(1, 774, 4), -torch.ones(1, 1, 5))
    with pytest.raises(RuntimeError):
        b200det.BBoxTransform()(torch.rand(1, 8, 4), torch.rand(1, 8, 4))
    with pytest.raises(RuntimeError):
        b200det.nms(torch.rand(4, 4), torch.rand(4), 0.5)
    with pytest.raises(RuntimeError):
        b200det.batched_nms(torch.rand(4, 4), torch.rand(4), torch.zeros(4, dtype=torch.int64), 0.5)
    for strategy in ("coordinate_trick", "auto"):                 # these reach nms() after a few torch ops
        with pytest.raises(RuntimeError):
            b200det.batched_nms(torch.rand(4, 4), torch.rand(4), torch.zeros(4, dtype=torch.int64), 0.5, strategy=strategy)
    with pytest.raises(RuntimeError):
        b200det.soft_nms(torch.rand(4, 4), torch.rand(4))
    with pytest.raises(RuntimeError):
        det = b200det.BatchDetections(torch.tensor([1]), torch.tensor([1]), torch.rand(1, 2), torch.zeros(1, 2, dtype=torch.int64),
                                      torch.rand(1, 2, 4), torch.zeros(1, 2, dtype=torch.int32), torch.zeros(1, 2, dtype=torch.int32))
        b200det.eval_pack(det, [1.0], num_classes=3)


def test_modules_are_stateless():
    for m in (b200det.Anchors(), b200det.FocalLoss(), b200det.BBoxTransform(), b200det.ClipBoxes()):
        assert len(m.state_dict()) == 0      # checkpoints stay interchangeable with the reference


def test_argument_validation_without_gpu():
    lib = L.lib()
    assert lib.k7_loss_workspace_bytes(0, 10, 10) == 0
    assert lib.k7_detect_workspace_bytes(2, 1000, 3) > 2 * 1000 * 56
    # argument checks of the entry points added last (no device work happens before they fail)
    assert lib.k7_batched_nms(None, None, None, 5, C.c_float(0.5), None, C.c_void_p(16), C.c_void_p(256), 1 << 20, None) == -1
    assert b"k7_batched_nms" in lib.k7_last_error()
    assert lib.k7_scale_inplace_multi(None, None, None, 1, None) == -1
    three = (C.c_void_p * 3)(16, 32, 48)
    n3 = (C.c_size_t * 3)(4, 4, 4)
    assert lib.k7_scale_inplace_multi(three, n3, three, 9, None) == -1 and b"count" in lib.k7_last_error()
    sp = b200det.Anchors().spec_for(64, 64)
    sp.total += 1
    rc = lib.k7_anchors_materialize(C.byref(sp), C.c_void_p(16), None)
    assert rc == -1 and b"anchor spec" in lib.k7_last_error()
