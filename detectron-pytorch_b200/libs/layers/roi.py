"""libs.layers.roi (reference libs/layers/roi.py): the NumPy RoI sampler `encode` (:16-100) and `decode` (:102-125).

The two computations — the float64 IoU matrix (bbox.pyx:15-55) and bbox_transform / bbox_transform_inv — run as
libdetbox.so kernels.  The sampling around them is host logic; what must match the reference for a seeded run to reproduce
its sample is only WHICH index arrays are handed to np.random.choice, with which sizes, in which order — `_draw` below
makes exactly the reference's calls (:55-56 foreground, :62-63 background, :96-98 the no-GT branch)."""
import numpy as np
import torch

from ... import ops
from ..._lib import DetboxError
from ..boxes.bbox_transform import bbox_transform, bbox_transform_inv, clip_boxes
from ..configs import config as cfg


def _draw(candidates, quota):
    """All candidates if the quota covers them, else `quota` of them drawn without replacement by np.random.choice —
    the reference's own call and condition (`inds.size > 0 and quota < inds.size`)."""
    if candidates.size > 0 and quota < candidates.size:
        return np.random.choice(candidates, size=quota, replace=False)
    return candidates


def _iou_rows(rois, gt):
    """Row-wise best match: (arg-max GT, its float64 IoU) from the device IoU matrix (first maximum, like ndarray.argmax)."""
    up = lambda a: torch.from_numpy(np.ascontiguousarray(a, dtype=np.float64)).cuda()
    iou = ops.bbox_overlaps(up(rois[:, 0:4]), up(gt[:, :4])).cpu().numpy()
    match = iou.argmax(axis=1)
    return match, iou[np.arange(rois.shape[0]), match]


def encode(gt_boxes, rois, num_classes):
    """roi.py:16-100.  gt_boxes [G,5], rois [R,>=4] -> (labels [R] float32 in {-1, 0, class}, bbox_targets [R,4K],
    bbox_inside_weights [R,4K]).  Foreground: best IoU >= cfg.fg_threshold, at most rois_per_image * roi_fg_fraction of
    them; background: best IoU < cfg.roi_bg_threshold, at most max(min(rois_per_image - fg quota, 3 * fg quota), 64);
    everything else -1.  A foreground row carries its regression target and weight 1 in the four columns of its class."""
    if not torch.cuda.is_available():
        raise DetboxError("detbox has no CPU path: a CUDA device is required")
    gt, rois = np.asarray(gt_boxes), np.asarray(rois)
    R, width = rois.shape[0], 4 * num_classes
    targets = np.zeros((R, width), np.float32)
    weights = np.zeros((R, width), np.float32)
    if gt.size == 0:                                             # :89-98: everything background, all but 64 rows ignored
        labels = np.zeros((R,), np.float32)
        n_bg = min(int(cfg.rois_per_image * (1 - cfg.roi_fg_fraction)), 64)
        if n_bg < R:
            labels[np.random.choice(np.arange(R), size=R - n_bg, replace=False)] = -1
        return labels, targets, weights
    match, best = _iou_rows(rois, gt)
    fg_all = np.flatnonzero(best >= cfg.fg_threshold)
    fg_quota = int(min(fg_all.size, cfg.rois_per_image * cfg.roi_fg_fraction))
    fg = _draw(fg_all, fg_quota)
    bg_quota = max(min(cfg.rois_per_image - fg_quota, fg_quota * 3), 64)
    bg = _draw(np.flatnonzero(best < cfg.roi_bg_threshold), bg_quota)
    # rows between the two thresholds are in neither set and keep the initial -1 (the reference re-assigns it, :66-68)
    labels = np.full((R,), -1, np.float32)
    labels[fg] = gt[match[fg], 4]
    labels[bg] = 0
    pos = fg[labels[fg] > 0]                                     # only rows with a positive class get targets (:145-150)
    if pos.size:
        cols = 4 * labels[pos].astype(np.int64)[:, None] + np.arange(4)[None, :]
        targets[pos[:, None], cols] = bbox_transform(rois[pos, 0:4], gt[match[pos], :4])
        weights[pos[:, None], cols] = 1
    return labels, targets, weights


def decode(boxes, scores, rois, ih, iw):
    """roi.py:102-125: boxes [R,4K] deltas, scores [R,K], rois [R,4] -> (the arg-max class's decoded box [R,4] clipped to
    the image, classes int32 [R], their scores [R])."""
    decoded = bbox_transform_inv(rois, deltas=boxes)
    classes = np.argmax(scores, axis=1).astype(np.int32)
    R = decoded.shape[0]
    picked = decoded.reshape(R, -1, 4)[np.arange(R), classes].astype(np.float32)
    return clip_boxes(np.ascontiguousarray(picked), (ih, iw)), classes, np.max(scores, axis=1)
