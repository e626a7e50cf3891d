"""detbox — B200-native (sm_100a) box-and-RoI hot path of CharlesShang/Detectron-PYTORCH.

Layout
  csrc/        hand-written CUDA kernels + the C ABI (include/detbox.h) -> libdetbox.so
  _lib.py      ctypes binding (no fallback: raises if the library is missing)
  ops*.py      device-tensor operators (torch used for memory + streams only)
  libs/        mirror of the reference's module paths and call signatures (the drop-in)

The directory name carries a hyphen, so import it as ``import detbox`` (alias module at the
repository root) or ``importlib.import_module("detectron-pytorch_b200")``.
"""
from . import _lib  # noqa: F401
from ._lib import DetboxError, lib  # noqa: F401
from . import ops, ops_post, ops_roi, ops_target, parallel, pipeline, fused  # noqa: F401
from . import libs  # noqa: F401  (reference-shaped mirror: detbox.libs.install() aliases it as `libs`)
from .ops import (anchor_pyramid, anchor_target, base_anchors, bbox_exclude_ignored_areas,  # noqa: F401
                  bbox_intersections, bbox_overlaps, box_decode, box_encode, clip_boxes_, expand_targets, launches,
                  make_pyramid_desc, make_target_cfg, pyramid_desc, NMS_GE_F32, NMS_GE_F64, NMS_GT_F32, ROI_BIN_CENTRE, ROI_CROP_NORM, ROI_GRID_POINT)
from .ops_post import dets_cluster, nms, sort_desc, threshold_segments, to_dets, topk_decode  # noqa: F401
from .ops_roi import assign_levels, roi_align_backward, roi_align_forward  # noqa: F401
from .ops_target import match_boxes, roi_target  # noqa: F401
