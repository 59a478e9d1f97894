"""ctypes binding of libk7det.so (include/k7det.h).

The library is built in-tree by ``csrc/build.sh`` (``__graft_entry__.build()``)
and loaded lazily.  There is no fallback: if the shared object is missing or a
call fails, a RuntimeError / ValueError is raised.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
import threading

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libk7det.so")
CSRC = os.path.join(_HERE, "csrc")

K7_MAX_LEVELS = 8
K7_MAX_CELL_ANCHORS = 16
BOX_XYXY, BOX_YXYX = 0, 1
DECODE_RETINA, DECODE_EFFDET = 0, 1
STATE_NEG, STATE_IGNORE, STATE_POS = 0, 1, 2


class K7AnchorSpec(C.Structure):
    _fields_ = [
        ("num_levels", C.c_int32),
        ("cell_anchors", C.c_int32),
        ("height", C.c_int32),
        ("width", C.c_int32),
        ("box_order", C.c_int32),
        ("total", C.c_int32),
        ("fh", C.c_int32 * K7_MAX_LEVELS),
        ("fw", C.c_int32 * K7_MAX_LEVELS),
        ("offset", C.c_int32 * (K7_MAX_LEVELS + 1)),
        ("stride", C.c_double * K7_MAX_LEVELS),
        ("base", ((C.c_double * 4) * K7_MAX_CELL_ANCHORS) * K7_MAX_LEVELS),
    ]


class K7MatchParams(C.Structure):
    _fields_ = [("neg_thr", C.c_float), ("pos_thr", C.c_float)]


K7_MAX_PEERS = 16


class K7PeerReduce(C.Structure):
    _fields_ = [
        ("world", C.c_int32),
        ("rank", C.c_int32),
        ("seq", C.c_uint32),
        ("timeout_ms", C.c_int32),
        ("mailbox", C.c_void_p * K7_MAX_PEERS),
        ("seq_counter", C.c_void_p),
        ("post_only", C.c_int32),
    ]


class K7LossParams(C.Structure):
    _fields_ = [
        ("alpha", C.c_float),
        ("beta", C.c_float),
        ("target_div", C.c_float * 4),
        ("yx_targets", C.c_int32),
        ("mean_divisor", C.c_int32),
        ("dreg_prezeroed", C.c_int32),
        ("from_logits", C.c_int32),
        ("peer_reduce", C.POINTER(K7PeerReduce)),
        ("match_generation", C.c_uint32),
        ("neg_thr", C.c_float),
        ("pos_thr", C.c_float),
        ("assign_out", C.c_void_p),
        ("spare_sms", C.c_int32),
        ("consume_best", C.c_int32),
    ]


class K7DecodeParams(C.Structure):
    _fields_ = [
        ("variant", C.c_int32),
        ("mean", C.c_float * 4),
        ("std", C.c_float * 4),
        ("clip", C.c_int32),
        ("height", C.c_int32),
        ("width", C.c_int32),
    ]


class K7DetectParams(C.Structure):
    _fields_ = [
        ("decode", K7DecodeParams),
        ("score_thr", C.c_float),
        ("iou_thr", C.c_float),
        ("pre_nms_topk", C.c_int32),
        ("max_detections", C.c_int32),
        ("from_logits", C.c_int32),
        ("class_aware", C.c_int32),
        ("nms_chain", C.c_int32),
    ]


class K7SoftNmsParams(C.Structure):
    _fields_ = [("method", C.c_int32), ("iou_thr", C.c_float), ("sigma", C.c_float), ("score_thr", C.c_float)]


class K7EvalParams(C.Structure):
    _fields_ = [
        ("mode", C.c_int32),
        ("max_detections", C.c_int32),
        ("num_classes", C.c_int32),
        ("reserved", C.c_int32),
        ("score_thr", C.c_double),
    ]


EVAL_COCO, EVAL_CSV = 0, 1
ABI_VERSION = 14                     # K7_ABI_VERSION of include/k7det.h


BEST_GEN_MAX = 65535                 # K7_BEST_GEN_MAX


def ignore_words(A: int) -> int:
    """K7_IGNORE_WORDS(A): uint32 words per image of the ignore bitmap."""
    return ((A + 127) >> 7) << 2


# every symbol include/k7det.h declares: name -> (restype, argtypes)
_P = C.c_void_p
SYMBOLS = {
    "k7_version": (C.c_int, []),
    "k7_last_error": (C.c_char_p, []),
    "k7_launch_count": (C.c_ulonglong, []),
    "k7_anchors_materialize": (C.c_int, [C.POINTER(K7AnchorSpec), _P, _P]),
    "k7_match": (C.c_int, [C.POINTER(K7AnchorSpec), _P, C.c_int, _P, C.c_int, C.c_int, C.c_int,
                           C.POINTER(K7MatchParams), _P, _P, _P, _P, _P, _P, _P]),
    "k7_match_scatter_workspace_bytes": (C.c_size_t, [C.c_int, C.c_int]),
    "k7_match_scatter": (C.c_int, [C.POINTER(K7AnchorSpec), _P, C.c_int, C.c_int, C.c_int, C.POINTER(K7MatchParams),
                                   _P, C.c_uint32, _P, _P]),
    "k7_peer_collect": (C.c_int, [C.POINTER(K7PeerReduce), _P, _P]),
    "k7_loss_workspace_bytes": (C.c_size_t, [C.c_int, C.c_int, C.c_int]),
    "k7_loss": (C.c_int, [C.POINTER(K7AnchorSpec), _P, C.c_int, _P, _P, _P, _P, _P, _P,
                          C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(K7LossParams),
                          _P, _P, _P, _P, _P, C.c_size_t, _P]),
    "k7_loss_levels_workspace_bytes": (C.c_size_t, [C.POINTER(K7AnchorSpec), C.c_int, C.c_int]),
    "k7_loss_levels": (C.c_int, [C.POINTER(K7AnchorSpec), _P, _P, _P, _P, _P, _P, C.c_int, C.c_int, C.c_int,
                                 C.POINTER(K7LossParams), _P, _P, _P, _P, _P, C.c_size_t, _P]),
    "k7_scale_inplace": (C.c_int, [_P, C.c_size_t, _P, _P]),
    "k7_scale_inplace_multi": (C.c_int, [_P, _P, _P, C.c_int, _P]),
    "k7_decode": (C.c_int, [C.POINTER(K7AnchorSpec), _P, C.c_int, C.c_size_t, _P, C.c_int, C.c_int,
                            C.POINTER(K7DecodeParams), _P, _P]),
    "k7_clip": (C.c_int, [_P, C.c_int, C.c_int, C.c_int, C.c_int, _P]),
    "k7_detect_workspace_bytes": (C.c_size_t, [C.c_int, C.c_int, C.c_int]),
    "k7_detect": (C.c_int, [C.POINTER(K7AnchorSpec), _P, C.c_int, _P, _P, C.c_int, C.c_int, C.c_int,
                            C.POINTER(K7DetectParams), _P, _P, C.c_size_t, _P]),
    "k7_detect_levels": (C.c_int, [C.POINTER(K7AnchorSpec), _P, _P, C.c_int, C.c_int, C.POINTER(K7DetectParams), _P, _P,
                                   C.c_size_t, _P]),
    "k7_detect_gather": (C.c_int, [_P, C.c_size_t, C.c_int, C.c_int, C.c_int, C.c_int,
                                   _P, _P, _P, _P, _P, _P]),
    "k7_nms_workspace_bytes": (C.c_size_t, [C.c_int]),
    "k7_nms": (C.c_int, [_P, _P, C.c_int, C.c_float, _P, _P, _P, C.c_size_t, _P]),
    "k7_batched_nms": (C.c_int, [_P, _P, _P, C.c_int, C.c_float, _P, _P, _P, C.c_size_t, _P]),
    "k7_nms_static_max_n": (C.c_int, []),
    "k7_nms_static": (C.c_int, [_P, _P, C.c_int, C.c_float, _P, _P, _P]),
    "k7_soft_nms": (C.c_int, [_P, _P, C.c_int, C.POINTER(K7SoftNmsParams), _P, _P, _P, _P]),
    "k7_peer_mailbox_create": (C.c_int, [C.c_int, C.POINTER(C.c_void_p), _P]),
    "k7_peer_mailbox_open": (C.c_int, [_P, C.POINTER(C.c_void_p)]),
    "k7_peer_mailbox_close": (C.c_int, [_P]),
    "k7_peer_mailbox_destroy": (C.c_int, [_P]),
    "k7_annotations_pad": (C.c_int, [_P, _P, C.c_int, C.c_int, _P, _P]),
    "k7_eval_pack_workspace_bytes": (C.c_size_t, [C.c_int, C.c_int]),
    "k7_eval_pack": (C.c_int, [_P, _P, _P, C.c_int, _P, C.c_int, C.c_int, _P, _P, _P, _P, C.POINTER(K7EvalParams),
                               _P, _P, _P, _P, C.c_size_t, _P]),
}

_lib = None
_lock = threading.Lock()


def build(verbose: bool = False) -> str:
    """Compile csrc/*.cu into libk7det.so for sm_100a (nvcc cross-compiles without a GPU)."""
    cmd = ["bash", os.path.join(CSRC, "build.sh")]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("building libk7det.so failed:\n" + r.stdout + r.stderr)
    if verbose:
        print(r.stdout.strip())
    return LIB_PATH


def lib():
    """The loaded library; raises if it has not been built (no CPU fallback exists)."""
    global _lib
    if _lib is None:
        with _lock:
            if _lib is None:
                if not os.path.exists(LIB_PATH):
                    raise RuntimeError(
                        f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                        "(b200det has no CPU or PyTorch fallback)")
                l = C.CDLL(LIB_PATH)
                for name, (res, args) in SYMBOLS.items():
                    fn = getattr(l, name)
                    fn.restype = res
                    fn.argtypes = args
                if l.k7_version() != ABI_VERSION:
                    raise RuntimeError("libk7det.so ABI version mismatch; rebuild")
                _lib = l
    return _lib


def check(rc: int, what: str):
    if rc == 0:
        return
    msg = lib().k7_last_error().decode(errors="replace")
    if rc == -1:
        raise ValueError(f"{what}: {msg}")
    raise RuntimeError(f"{what}: {msg} (code {rc})")


def ptr(t):
    return None if t is None else C.c_void_p(t.data_ptr())


def stream_ptr(device=None):
    return C.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def require_cuda(*tensors):
    for t in tensors:
        if t is not None and not t.is_cuda:
            raise RuntimeError("b200det operators run on CUDA tensors only (there is no CPU path)")


def as_f32_contig(t: torch.Tensor) -> torch.Tensor:
    """float32, contiguous, 16-byte aligned view (copy only when needed)."""
    if t.dtype != torch.float32:
        t = t.float()
    if not t.is_contiguous():
        t = t.contiguous()
    if t.data_ptr() % 16:
        t = t.clone(memory_format=torch.contiguous_format)
    return t


def ptr_array(tensors):
    """Host array of device pointers (const float* const*) for the per-level entry points; None stays NULL."""
    if tensors is None:
        return None
    return (C.c_void_p * len(tensors))(*[t.data_ptr() if t is not None else None for t in tensors])


def check_levels(levels, spec, B=None):
    """Per-level head outputs [B, cell_anchors*C, fh, fw] (NCHW, float32, contiguous, CUDA).  Returns (B, C)."""
    if spec is None:
        raise ValueError("per-level head outputs need anchors built by b200det.Anchors (the level geometry comes from them)")
    if len(levels) != spec.num_levels:
        raise ValueError(f"{len(levels)} level tensors for an anchor recipe of {spec.num_levels} levels")
    Cn = None
    for l, t in enumerate(levels):
        require_cuda(t)
        if t.dim() != 4 or t.dtype != torch.float32 or not t.is_contiguous():
            raise ValueError("level tensors must be contiguous float32 [B, anchors*classes, H, W]")
        if B is None:
            B = t.shape[0]
        if t.shape[0] != B or t.shape[2] != spec.fh[l] or t.shape[3] != spec.fw[l] or t.shape[1] % spec.cell_anchors:
            raise ValueError(f"level {l}: shape {tuple(t.shape)} does not match the anchor recipe "
                             f"({spec.cell_anchors} anchors per cell on a {spec.fh[l]}x{spec.fw[l]} grid)")
        c = t.shape[1] // spec.cell_anchors
        if Cn is None:
            Cn = c
        if c != Cn:
            raise ValueError("levels disagree on the number of classes")
    return B, Cn
