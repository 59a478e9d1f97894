"""Anchors — drop-in for the reference's anchor modules.

  Anchors      architectures/anchors.py:8-60 and ksevendet/architecture/anchors.py:8-60
               (same ctor: pyramid_levels, strides, sizes, ratios, scales, **kwargs with
               `sizes_mapper`, string `scales`, `logger`)
  AnchorsYX    architectures/utils_efficientdet.py:54-137 (anchor_scale, (y1,x1,y2,x2) layout)

The reference rebuilds the whole [A,4] table in numpy float64 and copies it to the
device on every forward.  Here the recipe is reduced once to a ``K7AnchorSpec`` (per
level: grid size, stride and the <=16 float64 in-cell boxes) and the hot kernels
generate anchors in registers from it.  ``forward(image)`` still returns the
``[1, A, 4]`` float32 tensor the reference returns (materialised by a CUDA kernel,
cached per image shape and device); that tensor carries the spec as
``tensor.k7_spec`` so FocalLoss / BBoxTransform / detect() can skip reading it.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Sequence

import numpy as np
import torch
import torch.nn as nn

from . import _lib as L


def _cell_table_xyxy(base_size, ratios: np.ndarray, scales: np.ndarray) -> np.ndarray:
    """In-cell boxes around the origin, float64, ratio-major/scale-minor — the arithmetic of
    generate_anchors (architectures/anchors.py:62-97), expression for expression, so the
    float64 values (and therefore the float32 roundings) are identical."""
    k = len(ratios) * len(scales)
    t = np.zeros((k, 4))
    t[:, 2:] = base_size * np.tile(scales, (2, len(ratios))).T
    areas = t[:, 2] * t[:, 3]
    t[:, 2] = np.sqrt(areas / np.repeat(ratios, len(scales)))
    t[:, 3] = t[:, 2] * np.repeat(ratios, len(scales))
    t[:, 0::2] -= np.tile(t[:, 2] * 0.5, (2, 1)).T
    t[:, 1::2] -= np.tile(t[:, 3] * 0.5, (2, 1)).T
    return t


def make_spec(height: int, width: int, strides: Sequence, cell_tables: Sequence[np.ndarray],
              box_order: int) -> L.K7AnchorSpec:
    nl = len(strides)
    if not 1 <= nl <= L.K7_MAX_LEVELS:
        raise ValueError(f"1..{L.K7_MAX_LEVELS} pyramid levels supported, got {nl}")
    k = cell_tables[0].shape[0]
    if not 1 <= k <= L.K7_MAX_CELL_ANCHORS:
        raise ValueError(f"1..{L.K7_MAX_CELL_ANCHORS} anchors per cell supported, got {k}")
    s = L.K7AnchorSpec()
    s.num_levels, s.cell_anchors = nl, k
    s.height, s.width, s.box_order = int(height), int(width), box_order
    off = 0
    for l, (st, tab) in enumerate(zip(strides, cell_tables)):
        # (image_shape + stride - 1) // stride, architectures/anchors.py:43
        fh = int((np.array(height) + st - 1) // st)
        fw = int((np.array(width) + st - 1) // st)
        s.fh[l], s.fw[l], s.offset[l] = fh, fw, off
        s.stride[l] = float(st)
        assert tab.shape == (k, 4)
        for a in range(k):
            for c in range(4):
                s.base[l][a][c] = float(tab[a, c])
        off += fh * fw * k
    s.offset[nl] = off
    s.total = off
    return s


class _AnchorBase(nn.Module):
    """Shared: spec cache per image shape, materialised tensor cache per (shape, device)."""

    def __init__(self):
        super().__init__()
        self._specs = {}
        self._tensors = {}

    def _build_spec(self, height: int, width: int) -> L.K7AnchorSpec:
        raise NotImplementedError

    def spec_for(self, height: int, width: int) -> L.K7AnchorSpec:
        key = (int(height), int(width))
        sp = self._specs.get(key)
        if sp is None:
            sp = self._specs[key] = self._build_spec(*key)
        return sp

    def num_anchors(self, height: int, width: int) -> int:
        return self.spec_for(height, width).total

    def forward(self, image: torch.Tensor) -> torch.Tensor:
        h, w = int(image.shape[2]), int(image.shape[3])
        dev = image.device
        if dev.type != "cuda":
            raise RuntimeError("b200det.Anchors runs on CUDA tensors only (there is no CPU path)")
        key = (h, w, dev)
        t = self._tensors.get(key)
        if t is None:
            sp = self.spec_for(h, w)
            t = torch.empty((1, sp.total, 4), dtype=torch.float32, device=dev)
            with torch.cuda.device(dev):
                L.check(L.lib().k7_anchors_materialize(C.byref(sp), L.ptr(t), L.stream_ptr(dev)),
                        "k7_anchors_materialize")
            t.k7_spec = sp
            self._tensors[key] = t
        return t


class Anchors(_AnchorBase):
    def __init__(self, pyramid_levels=None, strides=None, sizes=None, ratios=None, scales=None, **kwargs):
        super().__init__()
        self.pyramid_levels = [3, 4, 5, 6, 7] if pyramid_levels is None else pyramid_levels
        self.strides = [2 ** x for x in self.pyramid_levels] if strides is None else strides
        if 'sizes_mapper' in kwargs:
            # ksevendet/architecture/anchors.py:14-16: yaml gives a lambda as a string
            mapper = kwargs['sizes_mapper']
            mapper = eval(mapper) if isinstance(mapper, str) else mapper
            self.sizes = [mapper(x) for x in self.pyramid_levels]
        else:
            self.sizes = [2 ** (x + 2) for x in self.pyramid_levels] if sizes is None else sizes
        self.ratios = np.array([0.5, 1, 2]) if ratios is None else np.array(ratios)
        if isinstance(scales, str):
            scales = eval(scales)                      # ksevendet/architecture/anchors.py:20-21
        self.scales = (np.array([2 ** 0, 2 ** (1.0 / 3.0), 2 ** (2.0 / 3.0)])
                       if scales is None else np.array(scales))
        if len(self.strides) != len(self.sizes):
            raise ValueError("strides and sizes must have one entry per pyramid level")
        logger = kwargs.get('logger', None)
        if logger:
            logger.info('Anchors Sizes   : {}'.format(str(self.sizes)))
            logger.info('Anchors Scales  : {}'.format(str(self.scales)))
            logger.info('Anchors Ratios  : {}'.format(str(self.ratios)))
            logger.info('Anchors Num     : {}'.format(len(self.ratios) * len(self.scales)))

    def _build_spec(self, height, width):
        tabs = [_cell_table_xyxy(sz, self.ratios, self.scales) for sz in self.sizes]
        return make_spec(height, width, self.strides, tabs, L.BOX_XYXY)


class AnchorsYX(_AnchorBase):
    """utils_efficientdet.Anchors: (y1,x1,y2,x2) boxes, scale-major / ratio-minor in a cell."""

    def __init__(self, anchor_scale=4., pyramid_levels=None, **kwargs):
        super().__init__()
        self.anchor_scale = anchor_scale
        self.pyramid_levels = [3, 4, 5, 6, 7] if pyramid_levels is None else pyramid_levels
        self.strides = kwargs.get('strides', [2 ** x for x in self.pyramid_levels])
        self.scales = np.array(kwargs.get('scales', [2 ** 0, 2 ** (1.0 / 3.0), 2 ** (2.0 / 3.0)]))
        self.ratios = kwargs.get('ratios', [(1.0, 1.0), (1.4, 0.7), (0.7, 1.4)])

    def _build_spec(self, height, width):
        tabs = []
        for stride in self.strides:
            if width % stride != 0:                    # utils_efficientdet.py:109-110
                raise ValueError('input size must be divided by the stride.')
            rows = []
            for scale in self.scales:
                for ratio in self.ratios:
                    base = self.anchor_scale * stride * scale
                    hx = base * ratio[0] / 2.0
                    hy = base * ratio[1] / 2.0
                    # y - hy, x - hx, y + hy, x + hx  (utils_efficientdet.py:121-122)
                    rows.append([0.0 - hy, 0.0 - hx, hy, hx])
            tabs.append(np.array(rows, dtype=np.float64))
        return make_spec(height, width, self.strides, tabs, L.BOX_YXYX)


def spec_of(anchors) -> Optional[L.K7AnchorSpec]:
    """The recipe attached to a tensor returned by Anchors.forward (None for foreign tensors)."""
    if isinstance(anchors, L.K7AnchorSpec):
        return anchors
    return getattr(anchors, "k7_spec", None)
