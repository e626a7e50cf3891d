"""libs.layers.roi_align_tf (reference libs/layers/roi_align_tf/{crop_and_resize,roi_align_tf,
pyramid_roi_align}.py): TF-style crop_and_resize and the bin-centre RoIAlign built on it, with
autograd, on the hand-written kernels of roi_align.cu.  Gradients flow to the image only, as
upstream (crop_and_resize.py:55)."""
import numpy as np
import torch
import torch.nn as nn
from torch.autograd import Function

from ... import ops, ops_roi
from ..configs import config as cfg


class _RoiFn(Function):
    """One launch over L pyramid levels.  meta = dict(mode, ph, pw, scales, pool, ref_areas,
    all_levels, extrapolation); rois [R,5]; levels int32 [R] or None; *feats the L feature maps."""

    @staticmethod
    def forward(ctx, rois, levels, meta, *feats):
        ctx.meta, ctx.shapes = meta, [tuple(f.shape) for f in feats]
        ctx.nhwc = all(ops_roi._is_channels_last(f) for f in feats)     # the layout _make_cfg chose: NHWC only if EVERY level is
        # an L2-friendly block -> RoI order (meta["order"] = "auto" or a permutation; off by default) is computed once and
        # shared by forward and backward
        od = ops_roi._auto_order(meta.get("order"), ctx.nhwc, meta.get("all_levels", False), rois.view(-1, 5), meta["scales"], ctx.shapes,
                                 levels, meta.get("ref_areas"))
        ctx.order = od if od is not None else False
        out = ops_roi.roi_align_forward(list(feats), rois, meta["scales"], meta["ph"], meta["pw"], meta["mode"],
                                        pool=meta.get("pool"), levels=levels, ref_areas=meta.get("ref_areas"),
                                        all_levels=meta.get("all_levels", False), extrapolation=meta.get("extrapolation", 0.0),
                                        order=ctx.order)
        ctx.save_for_backward(rois, levels if levels is not None else rois.new_zeros(0), *(feats if meta.get("pool") == "max" else ()))
        ctx.has_levels = levels is not None
        return out

    @staticmethod
    def backward(ctx, grad_out):
        saved = ctx.saved_tensors
        rois, levels, feats = saved[0], (saved[1] if ctx.has_levels else None), list(saved[2:])
        m = ctx.meta
        grads = ops_roi.roi_align_backward(grad_out, ctx.shapes, rois, m["scales"], m["mode"], levels=levels,
                                           ref_areas=m.get("ref_areas"), all_levels=m.get("all_levels", False),
                                           channels_last=ctx.nhwc, pool=m.get("pool"), feats=feats or None, order=ctx.order)
        return (None, None, None) + tuple(grads)


class CropAndResizeFunction(object):
    """crop_and_resize.py:13-55.  Old-style call shape kept: CropAndResizeFunction(ch, cw, ext)(image,
    boxes, box_ind) with image [B,C,H,W], boxes [N,4] NORMALISED (y1,x1,y2,x2), box_ind int32 [N]."""

    def __init__(self, crop_height, crop_width, extrapolation_value=0):
        self.crop_height, self.crop_width, self.extrapolation_value = crop_height, crop_width, extrapolation_value

    def __call__(self, image, boxes, box_ind):
        rois = torch.cat((box_ind.reshape(-1, 1).to(boxes.dtype), boxes), 1)          # (b, y1, x1, y2, x2)
        meta = dict(mode=ops.ROI_CROP_NORM, ph=self.crop_height, pw=self.crop_width, scales=[1.0],
                    extrapolation=float(self.extrapolation_value))
        return _RoiFn.apply(rois.detach(), None, meta, image)


class CropAndResize(nn.Module):
    """crop_and_resize.py:58-72."""

    def __init__(self, crop_height, crop_width, extrapolation_value=0):
        super(CropAndResize, self).__init__()
        self.crop_height, self.crop_width, self.extrapolation_value = crop_height, crop_width, extrapolation_value

    def forward(self, image, boxes, box_ind):
        return CropAndResizeFunction(self.crop_height, self.crop_width, self.extrapolation_value)(image, boxes, box_ind)


class RoIAlign(nn.Module):
    """roi_align_tf.py:7-42: featuremap [N,C,H,W], rois [M,5] = (id,x1,y1,x2,y2) in image coordinates,
    spatial_scale float -> [M,C,oH,oW].  The normalise / de-normalise round trip of upstream
    (:24-40 then crop_and_resize.c:44-56) is evaluated inside the kernel in the same float32 order."""

    def __init__(self, crop_height, crop_width, extrapolation_value=0):
        super(RoIAlign, self).__init__()
        self.crop_height, self.crop_width, self.extrapolation_value = crop_height, crop_width, extrapolation_value

    def forward(self, featuremap, rois, spatial_scale):
        meta = dict(mode=ops.ROI_BIN_CENTRE, ph=self.crop_height, pw=self.crop_width, scales=[float(spatial_scale)])
        return _RoiFn.apply(rois.detach(), None, meta, featuremap)


class PyramidRoIAlign(nn.Module):
    """pyramid_roi_align.py:9-60: find the level of every box (argmin |log2 sqrt(area/ref_l)|), crop
    there, return [R,C,h,w] in the ORIGINAL RoI order.  Upstream: per level nonzero + gather + kernel,
    then cat + inverse-permutation scatter; here the level is picked inside ONE launch."""

    def __init__(self, aligned_height, aligned_width):
        super(PyramidRoIAlign, self).__init__()
        self.spatial_scales = np.asarray([1.0 / i for i in cfg.strides], dtype=np.float32)
        self.ref_areas_np = (np.asarray(cfg.anchor_scales, dtype=np.float32).mean(axis=1) * cfg.anchor_base) ** 2
        assert self.ref_areas_np.size == self.spatial_scales.size
        self.num_levels = self.ref_areas_np.shape[0]
        self.aligned_height, self.aligned_width = aligned_height, aligned_width

    def forward(self, pyramids, bboxes, batch_inds):
        rois = torch.cat((batch_inds.reshape(-1, 1).float(), bboxes), dim=1).detach()
        meta = dict(mode=ops.ROI_BIN_CENTRE, ph=self.aligned_height, pw=self.aligned_width,
                    scales=[float(s) for s in self.spatial_scales], ref_areas=[float(a) for a in self.ref_areas_np])
        return _RoiFn.apply(rois, None, meta, *pyramids[:self.num_levels])


class PyramidRoIAlign2(nn.Module):
    """pyramid_roi_align.py:62-94: every RoI on every level, concatenated along channels (one launch),
    then the 1x1 conv + ReLU `reduce` (a dense conv: plain torch, outside this path's scope)."""

    def __init__(self, aligned_height, aligned_width, num_channels=256):
        super(PyramidRoIAlign2, self).__init__()
        self.num_levels = len(cfg.strides)
        self.aligned_height, self.aligned_width = aligned_height, aligned_width
        self.spatial_scales = np.asarray([1.0 / i for i in cfg.strides], dtype=np.float32)
        self.reduce = nn.Sequential(nn.Conv2d(self.num_levels * num_channels, num_channels, 1, 1, padding=0, bias=True),
                                    nn.ReLU(inplace=True))

    def crop(self, pyramids, bboxes, batch_inds):
        rois = torch.cat((batch_inds.reshape(-1, 1).float(), bboxes), dim=1).detach()
        meta = dict(mode=ops.ROI_BIN_CENTRE, ph=self.aligned_height, pw=self.aligned_width,
                    scales=[float(s) for s in self.spatial_scales], all_levels=True)
        return _RoiFn.apply(rois, None, meta, *pyramids[:self.num_levels])

    def forward(self, pyramids, bboxes, batch_inds):
        return self.reduce(self.crop(pyramids, bboxes, batch_inds))
