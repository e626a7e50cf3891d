"""libs.layers.nms.nms_layer (reference libs/layers/nms/nms_layer.py:11-54)."""
import torch.nn as nn

from .box import apply_nms


class NMSFunction(object):
    """Callable with the old-style Function's shape: NMSFunction(thr)(boxes, scores) -> keep indices.
    (Upstream's forward concatenates boxes and scores on dim 0, nms_layer.py:25 — a bug that makes
    it unusable and it is unused by the nets; this one does what apply_nms does.)"""

    def __init__(self, overlap_threshold):
        self.overlap_threshold = overlap_threshold

    def __call__(self, boxes, scores):
        return apply_nms(boxes, scores, self.overlap_threshold)


class NMSLayer(nn.Module):
    def __init__(self, overlap_threshold):
        super(NMSLayer, self).__init__()
        self.overlap_threshold = overlap_threshold

    def forward(self, boxes, scores):
        return NMSFunction(self.overlap_threshold)(boxes, scores)
