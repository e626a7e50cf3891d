"""libs.layers.roi (reference libs/layers/roi.py): the NumPy RoI sampler `encode` (:16-100) and `decode` (:102-125).

The two computations — the float64 IoU matrix (bbox.pyx:15-55) and bbox_transform / bbox_transform_inv — run as
libdetbox.so kernels; the sampling logic around them is the reference's own host logic, kept call for call so that its
np.random.choice draws happen in the reference's order (a seeded run reproduces the reference's sample)."""
import numpy as np
import torch

from ... import ops
from ..._lib import DetboxError
from ..boxes.bbox_transform import bbox_transform, bbox_transform_inv, clip_boxes
from ..configs import config as cfg


def _compute_targets(ex_rois, gt_rois, labels, num_classes):
    """roi.py:127-153: 4-of-4K targets — only the row's own class carries its (dx,dy,dw,dh) and weight 1."""
    assert ex_rois.shape[0] == gt_rois.shape[0] and ex_rois.shape[1] == 4 and gt_rois.shape[1] == 4
    targets = bbox_transform(ex_rois, gt_rois)
    bbox_targets = np.zeros((labels.size, 4 * num_classes), dtype=np.float32)
    bbox_inside_weights = np.zeros(bbox_targets.shape, dtype=np.float32)
    for ind in np.where(labels > 0)[0]:
        start = 4 * int(labels[ind])
        bbox_targets[ind, start:start + 4] = targets[ind, 0:4]
        bbox_inside_weights[ind, start:start + 4] = 1
    return bbox_targets, bbox_inside_weights


def _unmap(data, count, inds, fill=0):
    """roi.py:155-167."""
    ret = np.empty((count,) + data.shape[1:], dtype=np.float32)
    ret.fill(fill)
    ret[inds] = data
    return ret


def encode(gt_boxes, rois, num_classes):
    """roi.py:16-100.  gt_boxes [G,5], rois [R,>=4] -> (labels [R] float32 in {-1, 0, class}, bbox_targets [R,4K],
    bbox_inside_weights [R,4K]).  fg: max overlap >= cfg.fg_threshold, subsampled to rois_per_image * roi_fg_fraction;
    bg: max overlap < cfg.roi_bg_threshold, subsampled to max(min(rois_per_image - fg, 3 fg), 64); the rest -1."""
    if not torch.cuda.is_available():
        raise DetboxError("detbox has no CPU path: a CUDA device is required")
    gt_boxes, rois = np.asarray(gt_boxes), np.asarray(rois)
    num_rois = rois.shape[0]
    if gt_boxes.size > 0:
        dev = lambda a: torch.from_numpy(np.ascontiguousarray(a, dtype=np.float64)).cuda()
        overlaps = ops.bbox_overlaps(dev(rois[:, 0:4]), dev(gt_boxes[:, :4])).cpu().numpy()      # R x G, float64
        gt_assignment = overlaps.argmax(axis=1)
        max_overlaps = overlaps[np.arange(num_rois), gt_assignment]
        labels = np.zeros([num_rois], dtype=np.float32)
        labels[:] = -1
        fg_inds = np.where(max_overlaps >= cfg.fg_threshold)[0]
        fg_rois = int(min(fg_inds.size, cfg.rois_per_image * cfg.roi_fg_fraction))
        if fg_inds.size > 0 and fg_rois < fg_inds.size:
            fg_inds = np.random.choice(fg_inds, size=fg_rois, replace=False)
        labels[fg_inds] = gt_boxes[gt_assignment[fg_inds], 4]
        bg_inds = np.where(max_overlaps < cfg.roi_bg_threshold)[0]
        bg_rois = max(min(cfg.rois_per_image - fg_rois, fg_rois * 3), 64)
        if bg_inds.size > 0 and bg_rois < bg_inds.size:
            bg_inds = np.random.choice(bg_inds, size=bg_rois, replace=False)
        labels[bg_inds] = 0
        ignore_inds = np.where((max_overlaps > cfg.roi_bg_threshold) & (max_overlaps < cfg.fg_threshold))[0]
        labels[ignore_inds] = -1
        keep_inds = np.append(fg_inds, bg_inds)
        bbox_targets, bbox_inside_weights = _compute_targets(rois[keep_inds, 0:4], gt_boxes[gt_assignment[keep_inds], :4],
                                                             labels[keep_inds], num_classes)
        bbox_targets = _unmap(bbox_targets, num_rois, keep_inds, 0)
        bbox_inside_weights = _unmap(bbox_inside_weights, num_rois, keep_inds, 0)
    else:                                                        # roi.py:89-98: no GT
        labels = np.zeros((num_rois,), np.float32)
        bbox_targets = np.zeros((num_rois, 4 * num_classes), np.float32)
        bbox_inside_weights = np.zeros((num_rois, 4 * num_classes), np.float32)
        bg_rois = min(int(cfg.rois_per_image * (1 - cfg.roi_fg_fraction)), 64)
        if bg_rois < num_rois:
            ignore_inds = np.random.choice(np.arange(num_rois), size=num_rois - bg_rois, replace=False)
            labels[ignore_inds] = -1
    return labels, bbox_targets, bbox_inside_weights


def decode(boxes, scores, rois, ih, iw):
    """roi.py:102-125: boxes [R,4K] deltas, scores [R,K], rois [R,4] -> (final_boxes [R,4] of the arg-max class, clipped,
    classes int32 [R], scores [R])."""
    boxes = bbox_transform_inv(rois, deltas=boxes)
    classes = np.argmax(scores, axis=1).astype(np.int32)
    scores = np.max(scores, axis=1)
    final_boxes = np.zeros((boxes.shape[0], 4), dtype=np.float32)
    for i in np.arange(0, boxes.shape[0]):
        ind = classes[i] * 4
        final_boxes[i, 0:4] = boxes[i, ind:ind + 4]
    final_boxes = clip_boxes(final_boxes, (ih, iw))
    return final_boxes, classes, scores
