"""ctypes binding of libdetbox.so (the C ABI declared in include/detbox.h).

There is no fallback: if the shared library is missing or a symbol cannot be resolved the
import of any op fails loudly.  Build it with ``python detectron-pytorch_b200/build.py``
(``__graft_entry__.build()`` does).
"""
from __future__ import annotations

import ctypes
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libdetbox.so")

c_void = ctypes.c_void_p
c_int, c_long, c_size, c_double, c_float = ctypes.c_int, ctypes.c_long, ctypes.c_size_t, ctypes.c_double, ctypes.c_float
P_int = ctypes.POINTER(ctypes.c_int)
P_double = ctypes.POINTER(ctypes.c_double)
P_float = ctypes.POINTER(ctypes.c_float)

MAX_LEVELS = 8


class TargetCfg(ctypes.Structure):
    """detbox_target_cfg (include/detbox.h)."""
    _fields_ = [("bg_threshold", c_double), ("fg_threshold", c_double), ("ignore_fraction", c_double),
                ("clobber_positives", c_int), ("box_encoding", c_int), ("bbweights", c_float * 4)]


class PyramidDesc(ctypes.Structure):
    """detbox_pyramid_desc (include/detbox.h)."""
    _fields_ = [("n_levels", c_int), ("n_cell", c_int * MAX_LEVELS), ("height", c_int * MAX_LEVELS), ("width", c_int * MAX_LEVELS)]


class RoiCfg(ctypes.Structure):
    """detbox_roi_cfg (include/detbox.h)."""
    _fields_ = [("n_levels", c_int),
                ("feat", c_void * MAX_LEVELS), ("grad_feat", c_void * MAX_LEVELS),
                ("height", c_int * MAX_LEVELS), ("width", c_int * MAX_LEVELS),
                ("spatial_scale", c_float * MAX_LEVELS), ("ref_area", c_float * MAX_LEVELS),
                ("batch", c_int), ("channels", c_int), ("layout", c_int), ("out_layout", c_int),
                ("pooled_h", c_int), ("pooled_w", c_int), ("mode", c_int), ("pool", c_int),
                ("all_levels", c_int), ("extrapolation", c_float), ("roi_order", c_void)]


class MatchCfg(ctypes.Structure):
    """detbox_match_cfg (include/detbox.h)."""
    _fields_ = [("bg_thresh", c_double), ("fg_thresh", c_double), ("ignore_thresh", c_double), ("min_size", c_float)]


# name -> (restype, argtypes); every symbol include/detbox.h declares
SIGNATURES = {
    "detbox_error_string": (ctypes.c_char_p, [c_int]),
    "detbox_version": (c_int, []),
    "detbox_last_cuda_error": (c_int, []),
    "detbox_trace_kernel": (c_int, [c_int, c_void, c_void]),
    "detbox_selftest_rcp": (c_int, [ctypes.c_uint32, ctypes.c_uint32, c_void, c_void]),
    "detbox_base_anchors_host": (c_int, [P_double, c_int, P_double, c_int, c_long, P_double]),
    "detbox_anchor_pyramid": (c_int, [c_void, c_int, P_int, P_int, P_int, P_int, c_void, c_int, c_void]),
    "detbox_bbox_overlaps_f64": (c_int, [c_void, c_long, c_void, c_long, c_void, c_void]),
    "detbox_bbox_intersections_f64": (c_int, [c_void, c_long, c_void, c_long, c_void, c_void]),
    "detbox_bbox_exclude_ignored_f64": (c_int, [c_void, c_long, c_void, c_long, c_double, c_void, c_void]),
    "detbox_box_encode": (c_int, [c_void, c_void, c_long, c_int, c_void, c_void]),
    "detbox_box_decode": (c_int, [c_void, c_void, c_long, c_int, c_int, c_int, c_void, c_void]),
    "detbox_clip_boxes": (c_int, [c_void, c_long, c_int, c_int, c_int, c_void]),
    "detbox_anchor_target_workspace_bytes": (c_size, [c_long, c_int, c_long]),
    "detbox_anchor_target": (c_int, [c_void, c_long, c_void, c_void, c_int, c_long, ctypes.POINTER(TargetCfg),
                                     c_void, c_void, c_void, c_void, c_void, c_void, c_size, c_void]),
    "detbox_anchor_target_ex_workspace_bytes": (c_size, [c_long, c_int, c_long, ctypes.POINTER(PyramidDesc)]),
    "detbox_anchor_target_ex": (c_int, [c_void, c_long, ctypes.POINTER(PyramidDesc), c_void, c_void, c_int, c_long,
                                        ctypes.POINTER(TargetCfg), c_int, c_void, c_void, c_void, c_void, c_void, c_void,
                                        c_void, c_size, c_void]),
    "detbox_pack_padded_gt": (c_int, [c_void, c_void, c_int, c_int, c_void, c_void, c_void]),
    "detbox_expand_targets_host": (c_int, [c_void, c_void, c_long, c_int, P_float, c_void, c_void, c_void, c_int]),
    "detbox_topk_decode_workspace_bytes": (c_size, [c_int, c_long, c_long, c_int]),
    "detbox_topk_decode": (c_int, [c_void, c_int, c_void, c_void, c_int, c_long, c_int, c_long, c_int, c_int,
                                   c_void, c_void, c_void, c_void, c_void, c_void, c_void, c_size, c_void]),
    "detbox_sort_desc_workspace_bytes": (c_size, [c_long]),
    "detbox_sort_desc": (c_int, [c_void, c_long, c_void, c_void, c_size, c_void]),
    "detbox_nms_workspace_bytes": (c_size, [c_long, c_int]),
    "detbox_nms": (c_int, [c_void, c_int, c_void, c_long, c_void, c_int, c_double, c_int, c_void, c_void,
                           c_void, c_size, c_void]),
    "detbox_nms_static": (c_int, [c_void, c_int, c_void, c_long, c_void, c_int, c_double, c_int, c_void, c_void, c_void]),
    "detbox_nms_variant": (c_int, [c_void, c_int, c_void, c_long, c_void, c_int, c_double, c_int, c_int, c_void, c_void,
                                   c_void, c_void, c_size, c_void]),
    "detbox_roi_align_forward": (c_int, [ctypes.POINTER(RoiCfg), c_void, c_long, c_void, c_void, c_void]),
    "detbox_roi_align_backward": (c_int, [ctypes.POINTER(RoiCfg), c_void, c_long, c_void, c_void, c_int, c_void]),
    "detbox_roi_align_backward_tiled_workspace_bytes": (c_size, [ctypes.POINTER(RoiCfg), c_long]),
    "detbox_roi_align_backward_tiled": (c_int, [ctypes.POINTER(RoiCfg), c_void, c_long, c_void, c_void, c_void, c_size, c_void]),
    "detbox_roi_locality_order_workspace_bytes": (c_size, [c_long]),
    "detbox_roi_locality_order": (c_int, [c_void, c_long, c_void, P_float, c_int, c_float, c_float, c_void, c_void, c_size, c_void]),
    "detbox_assign_levels": (c_int, [c_void, c_long, P_float, c_int, c_void, c_void]),
    "detbox_match_boxes": (c_int, [c_void, c_int, c_void, c_long, c_void, c_void, c_int, ctypes.POINTER(MatchCfg),
                                   c_void, c_void, c_void, c_void]),
    "detbox_roi_target": (c_int, [c_void, c_int, c_void, c_long, c_void, c_void, c_long, c_float, c_float,
                                  c_void, c_void, c_void, c_void]),
    "detbox_sample_compact": (c_int, [c_void, c_int, c_void, c_void, c_long, c_void, c_void, c_void, c_int,
                                      c_void, c_void, c_void, c_void, c_void, c_void, c_void]),
    "detbox_sample_gather": (c_int, [c_void, c_void, c_void, c_void, c_void, c_void, c_int, c_void, c_int,
                                     c_void, c_void, c_void, c_void]),
    "detbox_threshold_segments": (c_int, [c_void, c_void, c_long, c_int, c_float, c_void, c_void, c_void]),
    "detbox_to_dets": (c_int, [c_void, c_void, c_int, c_void, c_void, c_void, c_long, c_int, c_int, c_int, c_float,
                               c_void, c_long, c_void, c_void]),
    "detbox_pack_detections": (c_int, [c_void, c_void, c_void, c_void, c_void, c_int, c_int, c_void, c_void, c_void]),
    "detbox_dets_cluster": (c_int, [c_void, c_long, c_void, c_void, c_long, c_void, c_void, c_void]),
    "detbox_host_device_pointer": (c_int, [c_void, ctypes.POINTER(c_void)]),
    "_nms": (None, [c_void, c_void, c_void, c_int, c_int, c_float, c_int]),
    "detbox_nms_host": (c_int, [c_void, c_int, c_double, c_int, c_void, c_void, c_int]),
    "detbox_anchor_target_host": (c_int, [c_void, c_long, c_void, c_int, ctypes.POINTER(TargetCfg),
                                          c_void, c_void, c_void, c_void, c_int]),
}

_lib = None


class DetboxError(RuntimeError):
    pass


def lib():
    """Load libdetbox.so once and attach the prototypes.  Raises if it is absent."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise DetboxError("libdetbox.so is missing (%s): build it with "
                              "`python detectron-pytorch_b200/build.py`; there is no CPU fallback" % LIB_PATH)
        L = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)          # AttributeError if the symbol is not exported
            fn.restype, fn.argtypes = res, args
        _lib = L
    return _lib


def check(code: int, what: str = ""):
    if code != 0:
        msg = lib().detbox_error_string(code).decode()
        raise DetboxError("%s failed: %s (code %d)" % (what or "detbox call", msg, code))
