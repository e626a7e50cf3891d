"""libs.postprocessings.single_shot (reference libs/postprocessings/single_shot.py): the NumPy-facing
post-processing — compute_detection (:13-60: decode + clip every anchor, per class threshold -> NMS -> top
max_det_num), compute_detection_new (:62-126: class-agnostic NMS on the objectness, then per-class rows),
nms_advanced (:128-152) and nms_stitch (:154-177) over nms_multianchor, dets_cluster (:401-415, box voting).  Decode and every NMS run on the
device; the per-image / per-class Python loops are upstream's."""
import numpy as np
import torch

from ... import ops_post
from ..._lib import DetboxError
from ..boxes.cython_nms import nms_multianchor
from ..boxes.nms_wrapper import nms
from ..configs import config as cfg
from ..layers import anchor


def compute_detection(Outputs, Anchors, ih, iw, num_classes, score_threshold=0.5, overlap_threshold=0.7,
                      only_anchor=False, max_det_num=100, batch_size=None):
    """Outputs = (_, rpn_box [n*A,4], rpn_prob [n*A,C]) NumPy; Anchors [A,4].  Returns a list of n
    arrays [m,6] = (x1,y1,x2,y2,score,class) sorted by score.  `batch_size` defaults to
    rpn_box.shape[0] // Anchors.shape[0] (upstream reads cfg.batch_size, single_shot.py:21)."""
    _, rpn_box, rpn_prob = Outputs
    n = batch_size or rpn_box.shape[0] // Anchors.shape[0]
    assert Anchors.shape[0] * n == rpn_prob.shape[0] == rpn_box.shape[0], 'shape dont match'
    rpn_box_ = rpn_box.copy()
    if only_anchor:
        rpn_box_[...] = 0.0
    out_dim = rpn_box.shape[0] // n
    Dets = []
    for i in range(n):
        s, e = i * out_dim, (i + 1) * out_dim
        probs = rpn_prob[s:e]
        final_boxes, _, _ = anchor.decode(rpn_box_[s:e], probs, Anchors, ih, iw, num_classes)
        dets = []
        for j in range(1, num_classes):
            inds = np.where(probs[:, j] > score_threshold)[0]
            cls_dets = np.hstack((final_boxes[inds, :], probs[inds, j][:, np.newaxis])).astype(np.float32, copy=False)
            keep = nms(cls_dets, overlap_threshold)
            cls_dets = cls_dets[keep, :]
            tmp_det = np.zeros([cls_dets.shape[0], 6])
            tmp_det[:, 0:5] = cls_dets
            tmp_det[:, 5] = j
            dets.append(tmp_det)
        dets = np.vstack(dets) if dets else np.zeros((0, 6))
        order = np.argsort(-dets[:, 4], kind="stable")
        Dets.append(dets[order, :][:max_det_num, :])
    return Dets


def compute_detection_new(batch_rpn_box, batch_rpn_prob, batch_anchors, image_inds, ih, iw, num_classes,
                          score_threshold=0.5, overlap_threshold=0.7, only_anchor=False, max_dets=100, batch_size=None):
    """single_shot.py:62-126: rows already selected (one anchor row per box row, `image_inds` says which image).
    Per image: decode + clip, objectness = 1 - p[:,0], class-agnostic NMS, then every (row, class) with
    prob > score_threshold becomes a detection; sorted by score.  `batch_size` defaults to cfg.batch_size."""
    assert image_inds.shape[0] == batch_rpn_box.shape[0] == batch_rpn_prob.shape[0] == batch_anchors.shape[0], \
        'unmatch output: {} vs {} vs {} vs {}'.format(image_inds.shape[0], batch_rpn_box.shape[0],
                                                      batch_rpn_prob.shape[0], batch_anchors.shape[0])
    n = cfg.batch_size if batch_size is None else batch_size
    rpn_box_ = batch_rpn_box.copy()
    if only_anchor:
        rpn_box_[...] = 0.0
    Dets = []
    for i in range(n):
        inds = np.where(image_inds == i)[0]
        probs = batch_rpn_prob[inds]
        probs_positive = 1 - probs[:, 0]
        final_boxes, _, _ = anchor.decode(rpn_box_[inds], probs, batch_anchors[inds], ih, iw, num_classes)
        object_dets = np.hstack((final_boxes, probs_positive[:, np.newaxis])).astype(np.float32, copy=False)
        keep = np.asarray(nms(object_dets, overlap_threshold), dtype=np.int32)
        object_dets = object_dets[keep, :]
        if keep.size > max_dets:
            keep = keep[:max_dets]                 # upstream truncates `keep` only (object_dets stays longer), :105-106
        probs_objectness = probs[keep][:, 1:]
        d0, d1 = np.where(probs_objectness > score_threshold)
        cls_dets = np.zeros([d0.size, 6])
        cls_dets[:, 0:4] = object_dets[d0, 0:4]
        cls_dets[:, 4] = probs_objectness[d0, d1]
        cls_dets[:, 5] = d1 + 1
        order = np.argsort(-cls_dets[:, 4], kind="stable")
        Dets.append(cls_dets[order, :])
    return Dets


def nms_advanced(Dets):
    """single_shot.py:128-152: per image and class, nms_multianchor with the crowd / non-crowd threshold."""
    Dets_new = []
    for dets in Dets:
        cls_ids = dets[:, 5].astype(np.int32)
        dets_new = []
        for id in np.unique(cls_ids):
            thr = cfg.overlap_threshold_noncrowd if id in cfg.noncrowd_classes else cfg.overlap_threshold
            dt = dets[np.where(cls_ids == id)[0]]
            if dt.shape[0] > 1:
                keep, _ = nms_multianchor(dt[:, 0:5].astype(np.float32), float(thr))
                dt = dt[keep]
            dets_new.append(dt)
        Dets_new.append(np.vstack(dets_new) if dets_new else np.zeros((0, 6), dtype=dets.dtype))
    return Dets_new


def nms_stitch(Dets):
    """single_shot.py:154-177: drop low-score (0.1, 0.5] detections that nms_multianchor at IoU >= 0.95 absorbs."""
    thrsh_scores = [(0.1, 0.5, 0.95)]
    Dets_new = []
    for dets in Dets:
        probs = dets[:, 4].astype(np.float32)
        all_inds = np.arange(probs.shape[0])
        drops_inds = []
        for (score1, score2, th) in thrsh_scores:
            if score2 < cfg.score_threshold:
                continue
            th = max(cfg.overlap_threshold, th)
            inds = np.where(np.logical_and(probs > score1, probs <= score2))[0]
            if inds.size > 0 and probs.shape[0] > 1:
                keep_inds, _ = nms_multianchor(dets[:, 0:5].astype(np.float32), float(th))
                drops_inds.append(np.setdiff1d(inds, keep_inds))
        if len(drops_inds) > 0:
            dets = dets[np.setdiff1d(all_inds, np.unique(np.hstack(drops_inds)))]
        Dets_new.append(dets)
    return Dets_new


def dets_cluster(dets, keeps, clusters):
    """single_shot.py:401-415: score-weighted box voting — one row per kept index, the score-weighted mean box of
    its cluster and the cluster's best score (`keeps`, `clusters` as nms_multianchor returns them).  Runs on the
    device (detbox_dets_cluster); an empty `keeps` returns `dets` like upstream."""
    keeps = np.asarray(keeps, dtype=np.int64).reshape(-1)
    if keeps.size == 0:
        return dets
    if not torch.cuda.is_available():
        raise DetboxError("detbox has no CPU path: a CUDA device is required")
    d = torch.from_numpy(np.ascontiguousarray(dets[:, :5], dtype=np.float32)).cuda()
    out = ops_post.dets_cluster(d, torch.from_numpy(keeps).cuda(), torch.from_numpy(np.asarray(clusters, dtype=np.int64)).cuda())
    return out.cpu().numpy()
