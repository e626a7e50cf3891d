"""libs.layers.anchor (reference libs/layers/anchor.py): anchor-target `encode` and `decode`,
NumPy in / NumPy out, computed by the fused CUDA kernels (anchor_target.cu, box_transform.cu)."""
import ctypes

import numpy as np
import torch

from ... import ops
from ..._lib import DetboxError, check, lib
from ..boxes import bbox_transform
from ..configs import config as cfg


def _target_cfg():
    if cfg.rpn_sample_strategy not in ('simple', 'traditional'):
        raise ValueError('RPN sample strategy %s has not been implemented yet' % cfg.rpn_sample_strategy)     # anchor.py:124
    return ops.make_target_cfg(cfg.rpn_bg_threshold, cfg.rpn_fg_threshold, cfg.rpn_clobber_positives,
                               cfg.ignored_area_intersection_fraction, cfg.bbweights, cfg.rpn_box_encoding)


def sample_traditional_(labels, label_weights):
    """libs/layers/anchor.py:95-109 on one image's labels, IN PLACE: keep at most rpn_fg_fraction * rpn_batch_size
    foreground and max(min(rpn_batch_size - fg, 5 fg), 128) background anchors, the rest become -1; label weights
    stay zero for this strategy (:93, only 'simple' fills them).  The two np.random.choice draws happen in the
    reference's order and under its conditions, so a seeded run reproduces the reference's sample."""
    label_weights[...] = 0
    num_fg = int(cfg.rpn_fg_fraction * cfg.rpn_batch_size)
    fg_inds = np.where(labels >= 1)[0]
    if len(fg_inds) > num_fg:
        labels[np.random.choice(fg_inds, size=(len(fg_inds) - num_fg), replace=False)] = -1
    else:
        num_fg = len(fg_inds)
    num_bg = max(min(cfg.rpn_batch_size - num_fg, num_fg * 5), 128)
    bg_inds = np.where(labels == 0)[0]
    if len(bg_inds) > num_bg:
        labels[np.random.choice(bg_inds, size=(len(bg_inds) - num_bg), replace=False)] = -1


def encode(gt_boxes, all_anchors):
    """libs/layers/anchor.py:17-145.  gt_boxes [G,5] = (x1,y1,x2,y2,cls), cls < 0 marks an ignore
    region; all_anchors [A,4].  Returns (labels [A], label_weights [A], bbox_targets [A,4],
    bbox_inside_weights [A,4]) float32.  An empty gt_boxes gives all-background (:89-91)."""
    if not torch.cuda.is_available():
        raise DetboxError("detbox has no CPU path: a CUDA device is required")
    anc = np.ascontiguousarray(np.asarray(all_anchors).reshape(-1, 4), np.float32)
    gt = np.ascontiguousarray(np.asarray(gt_boxes, np.float32).reshape(-1, 5))
    A, G = anc.shape[0], gt.shape[0]
    labels, lw = np.empty((A,), np.float32), np.empty((A,), np.float32)
    bt, bw = np.empty((A, 4), np.float32), np.empty((A, 4), np.float32)
    tc = _target_cfg()
    check(lib().detbox_anchor_target_host(anc.ctypes.data, A, gt.ctypes.data if G else None, G, ctypes.byref(tc),
                                          labels.ctypes.data, lw.ctypes.data, bt.ctypes.data, bw.ctypes.data,
                                          torch.cuda.current_device()), "anchor.encode")
    ops._count(4 if G else 2)
    if cfg.rpn_sample_strategy == 'traditional':
        sample_traditional_(labels, lw)
    return labels, lw, bt, bw


def decode(boxes, scores, all_anchors, ih, iw, num_classes=None):
    """libs/layers/anchor.py:148-176 -> (final_boxes [R,4] clipped to the image, classes int32 [R], scores [R,K])."""
    num_classes = cfg.num_classes if num_classes is None else num_classes
    all_anchors = all_anchors.reshape((-1, 4))
    boxes = boxes.reshape((-1, 4))
    scores = scores.reshape((-1, num_classes))
    assert scores.shape[0] == boxes.shape[0] == all_anchors.shape[0], \
        'Anchor layer shape error %d vs %d vs %d' % (scores.shape[0], boxes.shape[0], all_anchors.shape[0])
    if boxes.shape[0] == 0:
        return np.zeros((0, 4), boxes.dtype), np.zeros((0,), np.int32), scores
    enc = cfg.rpn_box_encoding
    out = ops.box_decode(torch.from_numpy(np.ascontiguousarray(all_anchors, np.float32)).cuda(),
                         torch.from_numpy(np.ascontiguousarray(boxes, np.float32)).cuda(), enc, im_shape=(ih, iw))
    classes = np.argmax(scores, axis=1).astype(np.int32)
    return out.cpu().numpy().astype(boxes.dtype if boxes.dtype in (np.float32, np.float64) else np.float32, copy=False), classes, scores
