"""libs.layers.roi_target.modules.roi_target.RoITarget / functions.roi_target.RoITargetFunction (reference
libs/layers/roi_target/modules/roi_target.py:13-30, functions/roi_target.py:11-31): labels, box deltas and
box weights of the RoIs that survive sampling, from the ground truth of their own image.  The kernel keeps
ROITargetForward's arithmetic (roi_target_kernel.cu:49-90: float32 IoU, inter/area(roi) against ignore boxes,
last maximum wins) and is checked against that kernel compiled unmodified (tests/test_gpu_ref_cuda.py)."""
import torch
from torch.nn import Module

from ... import ops_target
from ..._lib import DetboxError


class RoITargetFunction(object):
    """Old-style Function upstream (stateful __init__, called once); no gradient flows through it."""

    def __init__(self, fg_overlap):
        self.fg_overlap = fg_overlap

    def forward(self, rois, roi_batch_inds, gt_boxes, gt_batch_inds):
        if not rois.is_cuda:
            raise NotImplementedError                                    # functions/roi_target.py:26
        with torch.no_grad():
            return ops_target.roi_target(rois, roi_batch_inds, gt_boxes, gt_batch_inds, self.fg_overlap, 0.)

    __call__ = forward

    def backward(self, grad_output):
        pass


class RoITarget(Module):
    def __init__(self, fg_overlap, box_encoding='fastrcnn'):
        super(RoITarget, self).__init__()
        self.fg_overlap = fg_overlap
        self.box_encoding = box_encoding

    def forward(self, rois, roi_batch_inds, gt_boxes, gt_batch_inds):
        rois = rois[:, :4].contiguous()
        labels, deltas, bbwgts = RoITargetFunction(self.fg_overlap)(rois, roi_batch_inds, gt_boxes.contiguous(),
                                                                    gt_batch_inds.contiguous())
        return labels, deltas, bbwgts


__all__ = ["RoITarget", "RoITargetFunction", "DetboxError"]
