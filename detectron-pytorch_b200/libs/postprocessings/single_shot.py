"""libs.postprocessings.single_shot.compute_detection (reference libs/postprocessings/single_shot.py:
13-60): the NumPy-facing post-processing — decode + clip every anchor, per class threshold -> NMS ->
top max_det_num.  Decode and NMS run on the device; the per-class Python loop is upstream's."""
import numpy as np

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
