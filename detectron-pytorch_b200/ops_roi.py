"""RoIAlign operators over FPN levels (device tensors in/out).  See ops.py for the conventions."""
from __future__ import annotations

import ctypes

import numpy as np
import torch

from . import _lib
from ._lib import check, lib
from .ops import _ws, POOL, ROI_BIN_CENTRE, ROI_CROP_NORM, ROI_GRID_POINT, _count, _dev, _ptr, _stream  # noqa: F401
import os

# The tile-centric dense backward (detbox_roi_align_backward_tiled: every gradient map written once, deterministic summation
# order, 1.6 GB of DRAM traffic instead of 2.9 GB at BASELINE configs[2], 0.53 ms against 0.77 ms for memset + scatter on B200) is
# the default where it applies; roi_align_backward(..., deterministic=False) or DETBOX_ROI_BWD_NOTILED=1 selects the scatter kernels.
_TILED_BWD = os.environ.get("DETBOX_ROI_BWD_NOTILED") is None
tiled_backward_calls = 0          # how often the tiled entry point ran (tests check that it really took a configuration)


def _is_channels_last(t: torch.Tensor) -> bool:
    return t.dim() == 4 and t.shape[1] > 1 and t.is_contiguous(memory_format=torch.channels_last) \
        and not t.is_contiguous()


def _make_cfg(feats, spatial_scales, ref_areas, ph, pw, mode, pool, all_levels, grads=None, extrapolation=0.0):
    L = len(feats)
    if not 1 <= L <= _lib.MAX_LEVELS:
        raise _lib.DetboxError("1..%d pyramid levels supported" % _lib.MAX_LEVELS)
    nhwc = all(_is_channels_last(f) for f in feats)
    c = _lib.RoiCfg()
    c.n_levels = L
    keep = []
    B, C = feats[0].shape[0], feats[0].shape[1]
    for l, f in enumerate(feats):
        if not f.is_cuda or f.dtype != torch.float32:
            raise _lib.DetboxError("feature maps must be float32 CUDA tensors: detbox has no CPU path")
        if f.shape[0] != B or f.shape[1] != C:
            raise _lib.DetboxError("all levels must share batch and channel sizes")
        if not nhwc:
            f = f.contiguous()
        keep.append(f)
        c.feat[l] = f.data_ptr()
        c.height[l], c.width[l] = f.shape[2], f.shape[3]
        c.spatial_scale[l] = float(spatial_scales[l])
        c.ref_area[l] = float(ref_areas[l]) if ref_areas is not None else 1.0
        if grads is not None:
            c.grad_feat[l] = grads[l].data_ptr()
    c.batch, c.channels = B, C
    c.layout = c.out_layout = 1 if nhwc else 0
    c.pooled_h, c.pooled_w = int(ph), int(pw)
    c.mode, c.pool, c.all_levels = int(mode), POOL[pool], int(bool(all_levels))
    c.extrapolation = float(extrapolation)
    return c, keep, nhwc


ORDER_MIN_ROIS = 512        # below this the feature working set fits L2 anyway and the sort is not worth two launches


def roi_locality_order(rois, spatial_scales, feat_shapes, levels=None, ref_areas=None):
    """detbox_roi_locality_order: int32 [R] processing order that keeps overlapping RoIs close together in time
    ((level, image, row band, column band) of the RoI centre), for detbox_roi_cfg.roi_order.  The canvas size is taken
    from level 0 (H_0 / scale_0)."""
    r = _dev(rois).view(-1, 5)
    R = r.shape[0]
    L = len(feat_shapes)
    lv = _dev(levels, torch.int32, "levels").view(-1) if levels is not None else None
    ra = np.ascontiguousarray(ref_areas, np.float32) if ref_areas is not None else None
    im_h, im_w = feat_shapes[0][2] / float(spatial_scales[0]), feat_shapes[0][3] / float(spatial_scales[0])
    order = torch.empty((R,), dtype=torch.int32, device=r.device)
    ws = _ws.get(lib().detbox_roi_locality_order_workspace_bytes(R), r.device)
    check(lib().detbox_roi_locality_order(r.data_ptr(), R, _ptr(lv), ra.ctypes.data_as(_lib.P_float) if ra is not None else None,
                                          L if (lv is not None or ra is not None) else 1, float(im_h), float(im_w), order.data_ptr(),
                                          ws.data_ptr(), ws.numel(), _stream()), "roi_locality_order")
    _count(2 if R <= 4096 else 14)
    return order


def _auto_order(order, nhwc, all_levels, r, spatial_scales, feat_shapes, levels, ref_areas):
    """order: None / False -> RoI order (block i handles RoI i); "auto" -> roi_locality_order for NCHW inputs with enough
    RoIs; a tensor -> used as is.  Off by default: measured on B200 at BASELINE configs[2] the order leaves the staged
    forward where it is (0.243 ms; ~600 resident blocks x a 180 KB window over 256 channels already fill the 126 MB L2, so
    overlapping RoIs rarely find each other's rows resident) and gains 3.5 % on the backward scatter — less than its
    two launches cost a forward (profiles/r02_roi_staged_vs_generic.txt)."""
    if order is None or order is False:
        return None
    if torch.is_tensor(order):
        return _dev(order, torch.int32, "order").view(-1)
    if order != "auto":
        raise _lib.DetboxError("order must be None, False, 'auto' or an int32 permutation")
    if nhwc or all_levels or r.shape[0] < ORDER_MIN_ROIS:
        return None
    return roi_locality_order(r, spatial_scales, feat_shapes, levels, ref_areas)


def _out_tensor(R, C, ph, pw, nhwc, device):
    if nhwc:
        return torch.empty((R, ph, pw, C), dtype=torch.float32, device=device).permute(0, 3, 1, 2)
    return torch.empty((R, C, ph, pw), dtype=torch.float32, device=device)


def roi_align_forward(feats, rois, spatial_scales, ph, pw, mode=ROI_BIN_CENTRE, pool=None,
                      levels=None, ref_areas=None, all_levels=False, extrapolation=0.0, order=None):
    """One launch over all pyramid levels.  feats: list of [B,C,H_l,W_l] (all NCHW-contiguous or all
    channels_last); rois [R,5] = (batch idx, x1, y1, x2, y2) in image coordinates.  Level of a RoI:
    `levels` (int32 [R]) if given, else computed in-kernel from ref_areas when there are several
    levels (pyramid_roi_align.py:33-37).  Output [R, C (or L*C), ph, pw] in the input's memory format.
    order: the block -> RoI assignment (results do not depend on it).  None / False: RoI order; "auto": for NCHW
    inputs with at least ORDER_MIN_ROIS RoIs an L2-friendly order is computed (roi_locality_order: two small launches);
    an int32 [R] permutation: used as is (share one between forward and backward)."""
    r = _dev(rois).view(-1, 5)
    cfg, keep, nhwc = _make_cfg(feats, spatial_scales, ref_areas, ph, pw, mode, pool, all_levels, extrapolation=extrapolation)
    if len(feats) > 1 and not all_levels and levels is None and ref_areas is None:
        raise _lib.DetboxError("several levels need `levels`, `ref_areas` or all_levels=True")
    lv = _dev(levels, torch.int32, "levels").view(-1) if levels is not None else None
    Ctot = cfg.channels * (len(feats) if all_levels else 1)
    od = _auto_order(order, nhwc, all_levels, r, spatial_scales, [tuple(f.shape) for f in feats], lv, ref_areas)
    cfg.roi_order = _ptr(od) or None
    out = _out_tensor(r.shape[0], Ctot, ph, pw, nhwc, r.device)
    check(lib().detbox_roi_align_forward(ctypes.byref(cfg), r.data_ptr(), r.shape[0], _ptr(lv), out.data_ptr(), _stream()),
          "roi_align_forward")
    _count(1)
    return out


def roi_align_backward(grad_out, feat_shapes, rois, spatial_scales, mode=ROI_BIN_CENTRE, levels=None,
                       ref_areas=None, all_levels=False, channels_last=False, grads=None, pool=None, feats=None, order=None,
                       deterministic=None):
    """Gradient w.r.t. the feature maps (dense, zero-filled here unless `grads` is given, in which
    case it is accumulated into).  grad_out [R, C(or L*C), ph, pw] in the same memory format as the
    features.  pool='avg'|'max' differentiates the fused 2x2 pool of RoIAlignAvg/Max as well
    ('max' needs `feats`, the forward's feature maps, to recompute the arg-max).  Returns the list
    of gradient maps.  The tile-centric dense kernel runs where it applies (NCHW, bin-centre / crop mode, no pool, pooled
    size <= 8 x 8, `grads` not given) — bit-identical results from run to run, unlike the atomics of the reference and of
    the scatter kernels; deterministic=False forces the scatter path, deterministic=True requires nothing more than the default."""
    r = _dev(rois).view(-1, 5)
    dev = r.device
    g = grad_out
    if not g.is_cuda or g.dtype != torch.float32:
        raise _lib.DetboxError("grad_out must be a float32 CUDA tensor")
    g = g.contiguous(memory_format=torch.channels_last) if channels_last else g.contiguous()
    zero = grads is None
    if grads is None:
        grads = []
        for (B, C, H, W) in feat_shapes:
            t = torch.empty((B, H, W, C), dtype=torch.float32, device=dev).permute(0, 3, 1, 2) if channels_last \
                else torch.empty((B, C, H, W), dtype=torch.float32, device=dev)
            grads.append(t)
    L = len(grads)
    c = _lib.RoiCfg()
    c.n_levels = L
    for l, t in enumerate(grads):
        c.grad_feat[l] = t.data_ptr()
        c.height[l], c.width[l] = t.shape[2], t.shape[3]
        c.spatial_scale[l] = float(spatial_scales[l])
        c.ref_area[l] = float(ref_areas[l]) if ref_areas is not None else 1.0
    c.batch, c.channels = grads[0].shape[0], grads[0].shape[1]
    c.layout = c.out_layout = 1 if channels_last else 0
    c.pooled_h, c.pooled_w = g.shape[2], g.shape[3]
    c.mode, c.pool, c.all_levels = int(mode), POOL[pool], int(bool(all_levels))
    if POOL[pool] == 2:
        if feats is None:
            raise _lib.DetboxError("roi_align_backward(pool='max') needs the forward's feature maps")
        # saved maps are converted to the chosen layout explicitly: a mixed-format pyramid ran (and saved) NCHW copies or not,
        # and the arg-max must be recomputed on memory that really has the layout the kernel is told
        feats = [f.contiguous(memory_format=torch.channels_last) if channels_last else f.contiguous() for f in feats]
        for l, f in enumerate(feats):
            c.feat[l] = f.data_ptr()
    lv = _dev(levels, torch.int32, "levels").view(-1) if levels is not None else None
    if zero and not channels_last and POOL[pool] == 0 and not all_levels and int(mode) != ROI_GRID_POINT and \
            (_TILED_BWD if deterministic is None else deterministic):
        # the dense gradient written once, tile by tile (no memset, no atomics); configurations it does not cover fall through
        nbytes = lib().detbox_roi_align_backward_tiled_workspace_bytes(ctypes.byref(c), r.shape[0])
        ws = _ws.get(nbytes, dev)
        rc = lib().detbox_roi_align_backward_tiled(ctypes.byref(c), r.data_ptr(), r.shape[0], _ptr(lv), g.data_ptr(),
                                                   ws.data_ptr(), ws.numel(), _stream())
        if rc != -4:                                  # DETBOX_ERR_UNSUPPORTED: not covered, use the scatter path
            check(rc, "roi_align_backward_tiled")
            global tiled_backward_calls
            tiled_backward_calls += 1
            _count(2 if r.shape[0] else 1)
            return grads
    od = _auto_order(order, channels_last, all_levels, r, spatial_scales, [tuple(t.shape) for t in grads], lv, ref_areas)
    c.roi_order = _ptr(od) or None
    check(lib().detbox_roi_align_backward(ctypes.byref(c), r.data_ptr(), r.shape[0], _ptr(lv), g.data_ptr(), int(zero), _stream()),
          "roi_align_backward")
    _count(1 + (L if zero else 0))
    return grads


def assign_levels(bboxes, ref_areas):
    """pyramid_roi_align.py:33-37 -> int32 [R] FPN level per box."""
    b = _dev(bboxes)[:, :4].contiguous()
    ra = np.ascontiguousarray(ref_areas, np.float32)
    out = torch.empty((b.shape[0],), dtype=torch.int32, device=b.device)
    check(lib().detbox_assign_levels(b.data_ptr(), b.shape[0], ra.ctypes.data_as(_lib.P_float), int(ra.size), out.data_ptr(), _stream()),
          "assign_levels")
    _count(1)
    return out
