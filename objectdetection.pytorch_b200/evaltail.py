"""The two data formats either side of the hot path: evaluator tails on
device, and the collater's annotation padding.

The reference's evaluators copy every image's (scores, labels, boxes) to the host and loop over the
detections in Python (datasettool/coco_eval.py:25-57, ksevendata_eval.py:24-61, csv_eval.py:93-121).
`eval_pack` does the same arithmetic for a whole batch on the device (k7_eval_pack) and returns ONE
packed float64 [N,7] array; `coco_results` / `csv_all_detections` turn it into the exact Python
objects the reference builds (the list of dicts it json.dumps, the nested per-class arrays), so
`COCOeval.loadRes` / `csv_eval.evaluate` can be fed unchanged.

No CPU fallback: the tensors must live on a CUDA device.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass
from typing import List, Optional, Sequence

import numpy as np
import torch

from . import _lib as L
from .losses import _WS


@dataclass
class PackedDetections:
    """rows[N,7] float64 on the device.
    coco: (image_id, x, y, w, h, score, category_id), images in order, score descending per image.
    csv:  (image index, x1, y1, x2, y2, score, label), ordered by (image, label, rank)."""
    mode: str
    rows: torch.Tensor                      # [N,7] f64
    row_offsets: torch.Tensor               # [B+1] i32 (device)
    class_offsets: Optional[torch.Tensor]   # csv: [B*num_classes+1] i32 (device)
    num_classes: int
    batch: int


def eval_pack(det, scales: Sequence[float], *, mode: str = "coco", image_ids: Optional[Sequence[int]] = None,
              category_of_label: Optional[Sequence[int]] = None, valid_class_id: Optional[Sequence[int]] = None,
              num_classes: Optional[int] = None, threshold: float = 0.05, max_detections: int = 100) -> PackedDetections:
    """det: a b200det.BatchDetections (or any object with .scores [B,cap] f32 descending, .labels [B,cap] i64,
    .boxes [B,cap,4] f32, .count [B]).  scales: data['scale'] of every image (coco_eval.py:18).
    mode "coco": evaluate_coco / coco_evaluate_ksevendata (valid_class_id, category_of_label =
    dataset.label_to_coco_label); mode "csv": csv_eval._get_detections (max_detections)."""
    if mode not in ("coco", "csv"):
        raise ValueError("mode must be 'coco' or 'csv'")
    scores, labels, boxes = det.scores, det.labels, det.boxes
    L.require_cuda(scores, labels, boxes)
    dev = scores.device
    scores = L.as_f32_contig(scores)
    boxes = L.as_f32_contig(boxes)
    labels = labels.to(torch.int64).contiguous()
    B, cap = scores.shape
    if labels.shape != (B, cap) or boxes.shape != (B, cap, 4):
        raise ValueError("scores [B,cap], labels [B,cap], boxes [B,cap,4] expected")
    if len(scales) != B:
        raise ValueError("one scale per image expected")
    counts = torch.as_tensor(det.count).to(device=dev, dtype=torch.int32).contiguous()
    if num_classes is None:
        if category_of_label is None:
            raise ValueError("num_classes or category_of_label is needed")
        num_classes = len(category_of_label)
    nc = int(num_classes)
    # boxes /= scale divides float32 values by float32(scale) (torch and numpy both cast the Python float down)
    scale_t = torch.from_numpy(np.asarray(scales, dtype=np.float64).astype(np.float32)).to(dev)
    ids_t = torch.as_tensor(list(image_ids), dtype=torch.int64).to(dev) if image_ids is not None else None
    cat_t = None
    if category_of_label is not None:
        if len(category_of_label) != nc:
            raise ValueError("category_of_label must have num_classes entries")
        cat_t = torch.as_tensor(list(category_of_label), dtype=torch.int32).to(dev)
    valid_t = None
    if valid_class_id:                                   # ksevendata_eval.py:46: `if valid_class_id and ...`
        v = np.zeros(nc, np.uint8)
        v[[int(c) for c in valid_class_id if 0 <= int(c) < nc]] = 1
        valid_t = torch.from_numpy(v).to(dev)
    p = L.K7EvalParams()
    p.mode = L.EVAL_COCO if mode == "coco" else L.EVAL_CSV
    p.max_detections, p.num_classes, p.score_thr = int(max_detections), nc, float(threshold)
    lib = L.lib()
    ws = _WS.get("eval", lib.k7_eval_pack_workspace_bytes(B, nc), dev)
    rows = torch.empty((B * cap, 7), dtype=torch.float64, device=dev)
    row_off = torch.empty(B + 1, dtype=torch.int32, device=dev)
    cls_off = torch.empty(B * nc + 1, dtype=torch.int32, device=dev) if mode == "csv" else None
    with torch.cuda.device(dev):
        L.check(lib.k7_eval_pack(L.ptr(scores), L.ptr(labels), L.ptr(boxes), cap, L.ptr(counts), 1, B, L.ptr(scale_t),
                                 L.ptr(ids_t), L.ptr(cat_t), L.ptr(valid_t), C.byref(p), L.ptr(rows), L.ptr(row_off),
                                 L.ptr(cls_off), L.ptr(ws), ws.numel(), L.stream_ptr(dev)), "k7_eval_pack")
    n = int(row_off[B].item())                           # the one synchronisation
    return PackedDetections(mode, rows[:n], row_off, cls_off, nc, B)


def coco_results(packed: PackedDetections) -> List[dict]:
    """The `results` list evaluate_coco json.dumps (coco_eval.py:49-57): one device->host copy."""
    if packed.mode != "coco":
        raise ValueError("coco_results needs mode='coco'")
    r = packed.rows.cpu().numpy()
    return [{"image_id": int(x[0]), "category_id": int(x[6]), "score": float(x[5]), "bbox": [float(v) for v in x[1:5]]}
            for x in r]


def csv_all_detections(packed: PackedDetections) -> List[List[np.ndarray]]:
    """all_detections[image][label] = float64 [n,5] (x1,y1,x2,y2,score) as csv_eval._get_detections returns
    (csv_eval.py:113-118); empty buckets are zeros((0,5))."""
    if packed.mode != "csv":
        raise ValueError("csv_all_detections needs mode='csv'")
    r = packed.rows.cpu().numpy()
    off = packed.class_offsets.cpu().numpy()
    nc = packed.num_classes
    return [[r[off[b * nc + c]:off[b * nc + c + 1], 1:6].copy() for c in range(nc)] for b in range(packed.batch)]


# ------------------------------------------------------------------ annotation packing
def collate_annotations(annots, device, *, return_ragged: bool = False):
    """The annotation half of the reference's collater (datasettool/dataloader.py:672-688): a list of
    per-image [n_i,5] arrays/tensors -> annot_padded [B, max(max n_i, 1), 5] float32 with -1 rows, on `device`.
    The batch crosses PCIe once as a pinned ragged block (rows + offsets); the padding is written on the
    device (k7_annotations_pad) instead of in a Python loop over a host tensor."""
    dev = torch.device(device)
    if dev.type != "cuda":
        raise RuntimeError("b200det operators run on CUDA tensors only (there is no CPU path)")
    B = len(annots)
    if B == 0:
        raise ValueError("empty batch")
    arrs = [np.asarray(a.detach().cpu() if isinstance(a, torch.Tensor) else a, dtype=np.float32).reshape(-1, 5) for a in annots]
    counts = np.array([a.shape[0] for a in arrs], dtype=np.int64)
    offsets = np.zeros(B + 1, dtype=np.int32)
    offsets[1:] = np.cumsum(counts)
    N, M = int(offsets[-1]), max(int(counts.max()), 1)
    # one pinned block: [offsets as int32 | rows as float32]
    block = torch.empty(B + 1 + 5 * max(N, 1), dtype=torch.int32).pin_memory()
    block[:B + 1] = torch.from_numpy(offsets)
    if N:
        block[B + 1:B + 1 + 5 * N] = torch.from_numpy(np.concatenate(arrs, axis=0).reshape(-1)).view(torch.int32)
    d_block = block.to(dev, non_blocking=True)
    d_off = d_block[:B + 1]
    d_rows = d_block[B + 1:].view(torch.float32)
    out = torch.empty((B, M, 5), dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        L.check(L.lib().k7_annotations_pad(L.ptr(d_rows), L.ptr(d_off), B, M, L.ptr(out), L.stream_ptr(dev)),
                "k7_annotations_pad")
    if return_ragged:
        return out, d_rows[:5 * N].view(-1, 5), d_off
    return out
