"""libs.postprocessings.single_shot (reference libs/postprocessings/single_shot.py): the NumPy-facing
post-processing — compute_detection (:13-60: decode + clip every anchor, per class threshold -> NMS -> top
max_det_num), compute_detection_new (:62-126: class-agnostic NMS on the objectness, then per-class rows),
nms_advanced (:128-152) and nms_stitch (:154-177) over nms_multianchor, dets_cluster (:401-415, box voting).  Decode and every NMS run on the
device; compute_detection runs all (image, class) pairs as segments of one NMS launch."""
import numpy as np
import torch

from ... import ops_post
from ..._lib import DetboxError
from ..boxes.cython_nms import nms_multianchor
from ..boxes.nms_wrapper import nms, nms_mode
from ..configs import config as cfg
from ..layers import anchor


def compute_detection(Outputs, Anchors, ih, iw, num_classes, score_threshold=0.5, overlap_threshold=0.7,
                      only_anchor=False, max_det_num=100, batch_size=None):
    """Outputs = (_, rpn_box [n*A,4], rpn_prob [n*A,C]) NumPy; Anchors [A,4].  Returns a list of n
    arrays [m,6] = (x1,y1,x2,y2,score,class) sorted by score.  `batch_size` defaults to
    rpn_box.shape[0] // Anchors.shape[0] (upstream reads cfg.batch_size, single_shot.py:21).

    Upstream loops over images and classes with one NMS call each (:30-52) — for COCO that is 80 x batch tiny
    synchronous round trips.  Here every (image, class) pair is one SEGMENT of a single detbox_nms launch (segments
    never suppress each other, tests: segments == per-pair loop), after one decode + clip per image; the rows each pair
    keeps, their order inside the pair, the class-major concatenation and the final score sort are upstream's."""
    _, rpn_box, rpn_prob = Outputs
    n = batch_size or rpn_box.shape[0] // Anchors.shape[0]
    assert Anchors.shape[0] * n == rpn_prob.shape[0] == rpn_box.shape[0], 'shape dont match'
    rpn_box_ = rpn_box.copy()
    if only_anchor:
        rpn_box_[...] = 0.0
    out_dim = rpn_box.shape[0] // n
    n_cls = max(num_classes - 1, 0)
    final_boxes = np.concatenate([anchor.decode(rpn_box_[i * out_dim:(i + 1) * out_dim], rpn_prob[i * out_dim:(i + 1) * out_dim],
                                                Anchors, ih, iw, num_classes)[0] for i in range(n)], 0) if n else np.zeros((0, 4), np.float32)
    # candidates of every (image, class j >= 1) pair: rows with prob > threshold (:32), pair-major like upstream's loops
    fg = rpn_prob[:, 1:num_classes]
    rows, cls0 = np.nonzero(fg > score_threshold)                 # row-major: ascending row, then class
    seg = (rows // out_dim) * n_cls + cls0
    order = np.argsort(seg, kind="stable")                        # group by pair, ascending row inside a pair (np.where order)
    rows, cls0, seg = rows[order], cls0[order], seg[order]
    cand = np.hstack((final_boxes[rows, :], fg[rows, cls0][:, np.newaxis])).astype(np.float32, copy=False)
    n_seg = n * n_cls
    if cand.shape[0] and n_seg:
        dev = lambda a, dt: torch.from_numpy(np.ascontiguousarray(a, dtype=dt)).cuda()
        keep_d, nk_d = ops_post.nms(dev(cand[:, :4], np.float32), dev(cand[:, 4], np.float32), overlap_threshold, nms_mode(),
                                    seg_ids=dev(seg, np.int32) if n_seg > 1 else None, n_seg=n_seg)
        nk = nk_d.cpu().numpy()
        keep = keep_d.cpu().numpy()[:int(nk[-1])]
        counts = nk[:n_seg] if n_seg > 1 else nk[:1]
    else:
        keep, counts = np.zeros((0,), np.int64), np.zeros((max(n_seg, 1),), np.int64)
    starts = np.concatenate([[0], np.cumsum(counts)])
    Dets = []
    for i in range(n):
        k = keep[starts[i * n_cls]:starts[(i + 1) * n_cls]] if n_cls else keep[:0]     # class-major, score-descending inside a class (:47-50)
        dets = np.zeros([k.size, 6])
        dets[:, 0:5] = cand[k]
        dets[:, 5] = cls0[k] + 1
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
