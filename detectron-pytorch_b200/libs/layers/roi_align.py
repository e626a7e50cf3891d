"""libs.layers.roi_align (reference libs/layers/roi_align/{functions,modules}/*.py): the grid-point
RoIAlign (one bilinear sample per cell of a grid that spans the RoI inclusive of both ends),
RoIAlignAvg / RoIAlignMax with the 2x2 stride-1 pool FUSED into the kernel (forward and backward),
and the level-assigning pyramid wrapper."""
import numpy as np
import torch
import torch.nn as nn
from torch.nn.modules.module import Module

from ... import ops
from ..configs import config as cfg
from .roi_align_tf import _RoiFn


class RoIAlignFunction(object):
    """functions/roi_align.py:12-51, old-style call shape: RoIAlignFunction(h, w, scale)(features, rois);
    rois [N,5] = (batch idx, x1, y1, x2, y2).  CUDA only, like upstream (:34 raises on CPU)."""

    def __init__(self, aligned_height, aligned_width, spatial_scale):
        self.aligned_width, self.aligned_height = int(aligned_width), int(aligned_height)
        self.spatial_scale = float(spatial_scale)

    def __call__(self, features, rois, pool=None):
        if not features.is_cuda:
            raise NotImplementedError
        meta = dict(mode=ops.ROI_GRID_POINT, ph=self.aligned_height, pw=self.aligned_width, scales=[self.spatial_scale], pool=pool)
        return _RoiFn.apply(rois.detach(), None, meta, features)


class RoIAlign(Module):
    """modules/roi_align.py:11-20."""

    def __init__(self, aligned_height, aligned_width):
        super(RoIAlign, self).__init__()
        self.aligned_width, self.aligned_height = int(aligned_width), int(aligned_height)

    def forward(self, features, rois, spatial_scale):
        return RoIAlignFunction(self.aligned_height, self.aligned_width, spatial_scale)(features, rois)


class RoIAlignAvg(Module):
    """modules/roi_align.py:23-33: (h+1)x(w+1) grid -> avg_pool2d(2, 1), in one kernel."""

    def __init__(self, aligned_height, aligned_width):
        super(RoIAlignAvg, self).__init__()
        self.aligned_width, self.aligned_height = int(aligned_width), int(aligned_height)

    def forward(self, features, rois, spatial_scale):
        return RoIAlignFunction(self.aligned_height, self.aligned_width, spatial_scale)(features, rois, pool="avg")


class RoIAlignMax(Module):
    """modules/roi_align.py:36-46: (h+1)x(w+1) grid -> max_pool2d(2, 1), in one kernel (the reference's
    stand-in for RoIPool)."""

    def __init__(self, aligned_height, aligned_width):
        super(RoIAlignMax, self).__init__()
        self.aligned_width, self.aligned_height = int(aligned_width), int(aligned_height)

    def forward(self, features, rois, spatial_scale):
        return RoIAlignFunction(self.aligned_height, self.aligned_width, spatial_scale)(features, rois, pool="max")


class PyramidRoIAlign(nn.Module):
    """modules/pyramid_roi_align.py:13-63 (uses RoIAlignAvg, :10)."""

    def __init__(self, aligned_height, aligned_width):
        super(PyramidRoIAlign, self).__init__()
        self.spatial_scales = np.asarray([1.0 / i for i in cfg.strides], dtype=np.float32)
        self.ref_areas_np = (np.asarray(cfg.anchor_scales, dtype=np.float32).mean(axis=1) * cfg.anchor_base) ** 2
        assert self.ref_areas_np.size == self.spatial_scales.size
        self.num_levels = self.ref_areas_np.shape[0]
        self.aligned_height, self.aligned_width = aligned_height, aligned_width

    def forward(self, pyramids, bboxes, batch_inds):
        rois = torch.cat((batch_inds.reshape(-1, 1).float(), bboxes), dim=1).detach()
        meta = dict(mode=ops.ROI_GRID_POINT, ph=self.aligned_height, pw=self.aligned_width, pool="avg",
                    scales=[float(s) for s in self.spatial_scales], ref_areas=[float(a) for a in self.ref_areas_np])
        return _RoiFn.apply(rois, None, meta, *pyramids[:self.num_levels])
