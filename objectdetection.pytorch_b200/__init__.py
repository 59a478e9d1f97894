"""b200det — B200-native detection-head hot path (drop-in for the reference's
Anchors / FocalLoss / BBoxTransform / ClipBoxes / eval tail).

Imported as ``b200det`` (see ../b200det/__init__.py).  The CUDA library is
loaded lazily on first use and there is NO CPU fallback: every operator raises
if ``libk7det.so`` is missing or the tensors are not on a CUDA device.
"""
__version__ = "0.1.0"

from .anchors import Anchors, AnchorsYX, spec_of            # noqa: F401
from .losses import FocalLoss                                # noqa: F401
from .utils import (BBoxTransform, BBoxTransformYX, ClipBoxes, BatchDetections, detect, nms, batched_nms, soft_nms,  # noqa: F401
                    postprocess, eval_tail, patch_model, strip_head_sigmoid)
from .evaltail import (PackedDetections, eval_pack, coco_results, csv_all_detections,  # noqa: F401
                       collate_annotations)
from .graphs import GraphedStep                              # noqa: F401
from . import synth                                          # noqa: F401
from . import distributed                                    # noqa: F401
