"""libs.nets.model — the hot-path call sites of detection_model (reference libs/nets/model.py) as
free functions with the methods' names, arguments and return tuples.  Tensors stay on the device."""
import torch

import numpy as np

from ... import fused, ops, ops_post
from ..configs import config as cfg


def _objectness(probs, activation=None):
    """model.py:54-61 (eager; the fused path computes it inside objectness_kernel)."""
    activation = cfg.class_activation if activation is None else activation
    if activation == 'softmax':
        return 1. - probs[:, 0]
    if activation == 'sigmoid':
        return probs.max(dim=1)[0]
    raise ValueError('Unknown activation funtion %s' % activation)


def _decode_and_choose_top_n_stage1(rpn_box, rpn_prob, anchors, top_n=1000, activation=None, per_image=False):
    """model.py:308-321: ONE descending selection over batch x levels x anchors (the multi-level
    merge), gather, decode (no clip) -> (boxes [k,4], probs [k,C], img_ids int64 [k], anchors [k,4]).
    Ties between equal scores: lower flat index first (upstream's torch.sort leaves them undefined).
    per_image=True takes top_n inside every image instead (the sharded mode)."""
    activation = cfg.class_activation if activation is None else activation
    out = ops_post.topk_decode(rpn_box, rpn_prob, anchors, top_n, activation, cfg.rpn_box_encoding, per_image,
                               want=("probs", "img_ids", "anchors"))
    return out["boxes"], out["probs"], out["img_ids"], out["anchors"]


def _decoding_and_thresholding_stage1(rpn_box, rpn_prob, anchors, score_threshold=0.3, max_dets=100, activation=None):
    """model.py:323-340: top max_dets*3, then objectness >= threshold (fallback: the single best)."""
    boxes, probs, img_ids, anc = _decode_and_choose_top_n_stage1(rpn_box, rpn_prob, anchors, top_n=max_dets * 3, activation=activation)
    objness = _objectness(probs, activation)
    inds = objness.ge(score_threshold).nonzero().view(-1)
    if inds.numel() == 0:
        inds = objness.sort(dim=0, descending=True, stable=True)[1][:1]
    return boxes[inds], probs[inds], img_ids[inds], anc[inds]


def _thresholding(boxes, probs, batch_ids, score_threshold=0.1, activation=None):
    """model.py:92-104: objectness >= threshold, or the ten best when nothing passes."""
    objness = _objectness(probs, activation)
    inds = objness.ge(score_threshold).nonzero().view(-1)
    if inds.numel() == 0:
        inds = objness.sort(dim=0, descending=True, stable=True)[1][:10]
    return boxes[inds], probs[inds], batch_ids[inds]


def _apply_nms_in_batch(boxes, probs, img_ids, anchors, activation, overlap_threshold=0.5):
    """model.py:343-355: per-image NMS, kept rows image-major and score-descending inside an image.
    Upstream loops over images in Python with a nonzero() sync per image; here the images are
    segments of one launch sequence and the single host read is the total count."""
    objness = _objectness(probs, activation)
    if boxes.shape[0] == 0:
        return boxes, probs, img_ids, anchors
    lo = int(img_ids.min().item())
    n_seg = int(img_ids.max().item()) - lo + 1
    keep, n_keep = ops_post.nms(boxes[:, :4], objness, float(overlap_threshold), ops.NMS_GT_F32,
                                seg_ids=(img_ids - lo).to(torch.int32), n_seg=n_seg)
    all_keeps = keep[:int(n_keep[-1].item())]
    return boxes[all_keeps], probs[all_keeps], img_ids[all_keeps], anchors[all_keeps]


def _stage_one_results(rpn_box, rpn_prob, anchors, top_n=2000, overlap_threshold=0.7, top_n_post_nms=None, activation=None):
    """model.py:79-90."""
    activation = cfg.class_activation if activation is None else activation
    boxes, probs, img_ids, anc = _decode_and_choose_top_n_stage1(rpn_box, rpn_prob, anchors, top_n=top_n, activation=activation)
    boxes, probs, img_ids, anc = _apply_nms_in_batch(boxes, probs, img_ids, anc, activation=activation, overlap_threshold=overlap_threshold)
    if top_n_post_nms is not None:
        return boxes[:top_n_post_nms], probs[:top_n_post_nms], img_ids[:top_n_post_nms]
    return boxes, probs, img_ids


def _to_dets(boxes, probs, img_ids, activation, multi_label, score_threshold, batch_size):
    """Shared body of the four to_Dets* conversions: tensors (or NumPy arrays) in, list of `batch_size` float64
    arrays [m_i,6] out, like upstream's everything2numpy + per-image loop (model.py:357-469)."""
    dev = boxes.device if torch.is_tensor(boxes) and boxes.is_cuda else torch.device("cuda")
    t = lambda x, dt: (x if torch.is_tensor(x) else torch.from_numpy(np.ascontiguousarray(x))).to(dev, dt)
    boxes, probs, img_ids = t(boxes, torch.float32).reshape(-1, 4), t(probs, torch.float32), t(img_ids, torch.int64).reshape(-1)
    n = cfg.batch_size if batch_size is None else int(batch_size)
    if boxes.shape[0] == 0:
        return [np.zeros((0, 6)) for _ in range(n)]
    probs = probs.reshape(boxes.shape[0], -1)
    if boxes.shape[0] > 1 and not bool((img_ids[1:] >= img_ids[:-1]).all()):     # upstream selects rows per image: any order works
        order = torch.sort(img_ids, stable=True)[1]
        boxes, probs, img_ids = boxes[order], probs[order], img_ids[order]
    return fused.split(*ops_post.to_dets(boxes.contiguous(), probs.contiguous(), img_ids, n, activation, multi_label, score_threshold))


def to_Dets(boxes, probs, img_ids, batch_size=None):
    """model.py:358-378: one detection per box, arg-max foreground class of a softmax head."""
    return _to_dets(boxes, probs, img_ids, 'softmax', False, 0.0, batch_size)


def to_Dets_sigmoid(boxes, probs, img_ids, batch_size=None):
    """model.py:381-402."""
    return _to_dets(boxes, probs, img_ids, 'sigmoid', False, 0.0, batch_size)


def to_Dets2(boxes, probs, img_ids, score_threshold=0.1, batch_size=None):
    """model.py:405-436: every (box, class) with prob > threshold; arg-max rows for an image where none passes."""
    return _to_dets(boxes, probs, img_ids, 'softmax', True, score_threshold, batch_size)


def to_Dets2_sigmoid(boxes, probs, img_ids, score_threshold=0.1, batch_size=None):
    """model.py:439-469."""
    return _to_dets(boxes, probs, img_ids, 'sigmoid', True, score_threshold, batch_size)


def get_final_results_stage1(rpn_box, rpn_prob, anchors, score_threshold=0.1, max_dets=100, overlap_threshold=0.5, activation=None,
                             batch_size=None):
    """model.py:474-497: top max_dets*3 -> threshold -> per-image NMS -> to_Dets2 / to_Dets2_sigmoid.  Returns the list
    of per-image float64 arrays [m,6] upstream returns; everything before the final copy runs on the device without a
    host synchronisation (detbox.fused.stage_one)."""
    activation = cfg.class_activation if activation is None else activation
    if activation not in ('softmax', 'sigmoid'):
        raise ValueError('Unknown activation function %s' % activation)
    n = cfg.batch_size if batch_size is None else int(batch_size)
    return fused.split(*fused.stage_one(rpn_box, rpn_prob, anchors, n, score_threshold, max_dets, overlap_threshold, activation,
                                        cfg.rpn_box_encoding))
