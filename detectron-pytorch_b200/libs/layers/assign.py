"""libs.layers.assign (reference libs/layers/assign.py:10-43): FPN level of a box from its area,
k = floor(4 + log2(sqrt(w*h) / base_size)) clamped to [min_k, max_k].  A dozen scalar operations per GT box on the
host in the reference (unused by its nets); here the same arithmetic on the host in the input's dtype."""
import numpy as np


def assign_boxes(gt_boxes, min_k=2, max_k=5, base_size=224.0):
    gt_boxes = np.asarray(gt_boxes)
    if gt_boxes.size == 0:
        return np.asarray([], dtype=np.int32)
    ws = gt_boxes[:, 2] - gt_boxes[:, 0]
    hs = gt_boxes[:, 3] - gt_boxes[:, 1]
    with np.errstate(all='raise'):                               # assign.py:21 np.seterr(all='raise'): a zero area raises
        k = np.floor(4 + np.log2(np.sqrt(ws * hs) / base_size))
    k[k < min_k] = min_k
    k[k > max_k] = max_k
    return k.astype(np.int32)
