"""libs.layers.data_layer — only the hot-path hook compute_rpn_targets_in_batch
(reference libs/layers/data_layer.py:228-239), batched into ONE set of kernel launches."""
import numpy as np
import torch

from ... import ops
from ..._lib import DetboxError
from . import anchor as _anchor


def compute_rpn_targets_in_batch(gt_boxes_list, anchors):
    """gt_boxes_list: B arrays [G_i,5]; anchors: [A,4] array (or a list of B identical arrays, as
    upstream allows).  Returns [labels [B,A], label_weights [B,A], bbox_targets [B,A,4],
    bbox_inside_weights [B,A,4]] float32 NumPy arrays (np.stack of the per-image results upstream)."""
    if not torch.cuda.is_available():
        raise DetboxError("detbox has no CPU path: a CUDA device is required")
    B = len(gt_boxes_list)
    if isinstance(anchors, list):
        if any(a is not anchors[0] and not np.array_equal(a, anchors[0]) for a in anchors[1:]):   # per-image anchor sets
            res = [_anchor.encode(g, a) for g, a in zip(gt_boxes_list, anchors)]
            return [np.stack([r[i] for r in res]) for i in range(4)]
        anchors = anchors[0]
    gts = [np.asarray(g, np.float32).reshape(-1, 5) for g in gt_boxes_list]
    off = np.zeros(B + 1, np.int32)
    off[1:] = np.cumsum([g.shape[0] for g in gts])
    gt = np.concatenate(gts, 0) if off[-1] else np.zeros((0, 5), np.float32)
    dev = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    out = ops.anchor_target(dev(np.asarray(anchors, np.float32).reshape(-1, 4)), dev(gt), dev(off), _anchor._target_cfg())
    res = [t.cpu().numpy() for t in out]
    if _anchor.cfg.rpn_sample_strategy == 'traditional':         # per image, in batch order: upstream's loop calls encode() image by image
        for b in range(B):
            _anchor.sample_traditional_(res[0][b], res[1][b])
    return res


def compute_rpn_targets_in_batch_cuda(gt_boxes_list, anchors, pyramid=None):
    """Same, device tensors in and out (no host round trip): gt_boxes_list of CUDA tensors [G_i,5],
    anchors CUDA [A,4] -> 4 CUDA tensors.  What RetinaNet.forward wants when rpn_targets is None
    (libs/nets/retinatnet.py:70-75) — the targets are consumed on the device."""
    off = torch.zeros(len(gt_boxes_list) + 1, dtype=torch.int32)
    off[1:] = torch.tensor([int(g.shape[0]) for g in gt_boxes_list]).cumsum(0)
    gt = torch.cat([g.reshape(-1, 5).float() for g in gt_boxes_list], 0) if int(off[-1]) else anchors.new_zeros((0, 5))
    return ops.anchor_target(anchors, gt, off.to(anchors.device), _anchor._target_cfg(), pyramid=pyramid)


def compute_rpn_targets_padded_cuda(gt_padded, gt_counts, anchors, pyramid=None, format="reference"):
    """The collate contract of a device-side target pipeline: gt_padded [B, Gmax, 5] and gt_counts int32
    [B] CUDA tensors — what a collate_fn that pads the per-image GT (libs/datasets/citypersons2.py:1045-1070) uploads
    with the images — anchors CUDA [A,4].  Nothing is concatenated or counted on the host; the targets stay in HBM for
    the loss (libs/nets/model.py:106-128 reads rpn_labels / rpn_bboxes / rpn_bbwghts on the device).
    format="compact": (packed uint8 [B,A], bbox_targets [B,A,4], counts int32 [B,2])."""
    gt, off = ops.pack_padded_gt(gt_padded, gt_counts)
    return ops.anchor_target(anchors, gt, off, _anchor._target_cfg(), pyramid=pyramid, format=format)
