"""Loader for the compiled, UNMODIFIED reference modules under ``oracle/_ref`` (built by
``oracle/build_ref.py``).  TEST INFRASTRUCTURE ONLY — never imported by the product path.

``load()`` returns a namespace with the reference's own callables:

  cython_bbox      libs/boxes/bbox.pyx            bbox_overlaps / bbox_intersections / ...
  cython_nms       libs/boxes/nms.pyx             nms / nms_new / nms_multianchor
  cpu_nms          libs/nms/cpu_nms.pyx           cpu_nms
  py_cpu_nms       libs/nms/py_cpu_nms.py         py_cpu_nms            ('>' comparator)
  boxes_anchor     libs/boxes/anchor.py           anchor_pyramid / anchors / ...
  bbox_transform   libs/boxes/bbox_transform.py   bbox_transform(_inv) / clip_boxes / *_linear
  layers_anchor    libs/layers/anchor.py          encode / decode
  layers_box       libs/layers/box.py             matching_box / sample_rois (CPU torch tensors in and out)
  layers_roi       libs/layers/roi.py             encode / decode (RoI targets, NumPy)
  cfg              libs/configs/config.py         the module-global config the above read
  crop_and_resize_forward / crop_and_resize_backward
                   libs/layers/roi_align_tf/src/crop_and_resize.c through the TH stub

The only run-time shims are the NumPy aliases the 2018 code expects (``np.float``, ``np.int``):
attribute aliases, no arithmetic.
"""
from __future__ import annotations

import ctypes
import os
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.path.join(HERE, "_ref")

_cache = None


def available() -> bool:
    return os.path.exists(os.path.join(REF, "BUILT_FROM"))


class _THFloat(ctypes.Structure):
    _fields_ = [("data", ctypes.POINTER(ctypes.c_float)), ("size", ctypes.c_long * 4),
                ("nDimension", ctypes.c_int)]


class _THInt(ctypes.Structure):
    _fields_ = [("data", ctypes.POINTER(ctypes.c_int)), ("size", ctypes.c_long * 4),
                ("nDimension", ctypes.c_int)]


def _thf(a: np.ndarray) -> _THFloat:
    assert a.dtype == np.float32 and a.flags.c_contiguous
    t = _THFloat()
    t.data = a.ctypes.data_as(ctypes.POINTER(ctypes.c_float))
    for i, s in enumerate(a.shape):
        t.size[i] = s
    t.nDimension = a.ndim
    return t


def _thi(a: np.ndarray) -> _THInt:
    assert a.dtype == np.int32 and a.flags.c_contiguous
    t = _THInt()
    t.data = a.ctypes.data_as(ctypes.POINTER(ctypes.c_int))
    for i, s in enumerate(a.shape):
        t.size[i] = s
    t.nDimension = a.ndim
    return t


def load():
    """Import the compiled reference; raises if oracle/_ref was not built."""
    global _cache
    if _cache is not None:
        return _cache
    if not available():
        raise RuntimeError("oracle/_ref not built: run `python oracle/build_ref.py` where "
                           "/root/reference exists")
    # aliases removed from NumPy >= 1.24 that the reference spells out
    if not hasattr(np, "float"):
        np.float = np.float64  # type: ignore[attr-defined]
    if not hasattr(np, "int"):
        np.int = np.int64  # type: ignore[attr-defined]
    if REF not in sys.path:
        sys.path.insert(0, REF)
    for m in [k for k in sys.modules if k == "libs" or k.startswith("libs.")]:
        if not getattr(sys.modules[m], "__file__", "").startswith(REF):
            raise RuntimeError("a different `libs` package is already imported: %s" % m)
    import libs.configs.config as cfg  # noqa
    import libs.boxes.cython_bbox as cython_bbox
    import libs.boxes.cython_nms as cython_nms
    import libs.boxes.cython_anchor as cython_anchor  # noqa: F401  (imported by boxes.anchor)
    import libs.boxes.anchor as boxes_anchor
    import libs.boxes.bbox_transform as bbox_transform
    import libs.nms.cpu_nms as cpu_nms
    import libs.nms.py_cpu_nms as py_cpu_nms
    import libs.layers.anchor as layers_anchor
    try:                                    # needs torch (CPU tensors) — matching_box / sample_rois / roi.encode
        import libs.layers.box as layers_box
        import libs.layers.roi as layers_roi
    except ImportError:
        layers_box = layers_roi = None

    ns = types.SimpleNamespace(
        cfg=cfg, cython_bbox=cython_bbox, cython_nms=cython_nms, boxes_anchor=boxes_anchor,
        bbox_transform=bbox_transform, cpu_nms=cpu_nms, py_cpu_nms=py_cpu_nms,
        layers_anchor=layers_anchor, layers_box=layers_box, layers_roi=layers_roi)

    lib = ctypes.CDLL(os.path.join(REF, "libcrop_and_resize_ref.so"))
    lib.crop_and_resize_forward.argtypes = [
        ctypes.POINTER(_THFloat), ctypes.POINTER(_THFloat), ctypes.POINTER(_THInt),
        ctypes.c_float, ctypes.c_int, ctypes.c_int, ctypes.POINTER(_THFloat)]
    lib.crop_and_resize_forward.restype = None
    lib.crop_and_resize_backward.argtypes = [
        ctypes.POINTER(_THFloat), ctypes.POINTER(_THFloat), ctypes.POINTER(_THInt),
        ctypes.POINTER(_THFloat)]
    lib.crop_and_resize_backward.restype = None

    def crop_and_resize_forward(image, boxes, box_ind, extrapolation_value, crop_h, crop_w):
        """libs/layers/roi_align_tf/src/crop_and_resize.c:115 — image [B,C,H,W] f32,
        boxes [N,4] normalised (y1,x1,y2,x2) f32, box_ind [N] i32 -> crops [N,C,ch,cw]."""
        image = np.ascontiguousarray(image, np.float32)
        boxes = np.ascontiguousarray(boxes, np.float32)
        box_ind = np.ascontiguousarray(box_ind, np.int32)
        crops = np.zeros((boxes.shape[0], image.shape[1], crop_h, crop_w), np.float32)
        ti, tb, tx, tc = _thf(image), _thf(boxes), _thi(box_ind), _thf(crops)
        lib.crop_and_resize_forward(ctypes.byref(ti), ctypes.byref(tb), ctypes.byref(tx),
                                    float(extrapolation_value), crop_h, crop_w, ctypes.byref(tc))
        return crops

    def crop_and_resize_backward(grads, boxes, box_ind, image_shape):
        """crop_and_resize.c:164 — grads [N,C,ch,cw] -> grads_image [B,C,H,W] (serial)."""
        grads = np.ascontiguousarray(grads, np.float32)
        boxes = np.ascontiguousarray(boxes, np.float32)
        box_ind = np.ascontiguousarray(box_ind, np.int32)
        gi = np.zeros(tuple(image_shape), np.float32)
        tg, tb, tx, to = _thf(grads), _thf(boxes), _thi(box_ind), _thf(gi)
        lib.crop_and_resize_backward(ctypes.byref(tg), ctypes.byref(tb), ctypes.byref(tx),
                                     ctypes.byref(to))
        return gi

    ns.crop_and_resize_forward = crop_and_resize_forward
    ns.crop_and_resize_backward = crop_and_resize_backward
    _cache = ns
    return ns
