"""libs.layers.anchor_target.modules.anchor_target.AnchorTarget (reference
libs/layers/anchor_target/modules/anchor_target.py:13-59).  Upstream's forward raises
("Implement on cude is buggy, dont use it", :25); this one works and follows the semantics of
libs/layers/anchor.py::encode, returning the module's own triple."""
import numpy as np
import torch
from torch.nn import Module

from ... import ops


class AnchorTarget(Module):
    def __init__(self, bg_threshold, fg_threshold, ignored_threshold=0.2, box_encoding='fastrcnn'):
        super(AnchorTarget, self).__init__()
        self.fg_threshold, self.bg_threshold = fg_threshold, bg_threshold
        self.box_encoding, self.ignored_threshold = box_encoding, ignored_threshold

    def forward(self, anchors, gt_boxes):
        """anchors: CUDA [A,4]; gt_boxes: a NumPy array [G,5] or a list of them (one per image).
        Returns (labels int64 [B*A], deltas [B*A,4], bbwgts [B*A,4]) concatenated over images."""
        gts = gt_boxes if isinstance(gt_boxes, list) else [gt_boxes]
        gts = [np.asarray(g, np.float32).reshape(-1, 5) for g in gts]
        off = np.zeros(len(gts) + 1, np.int32)
        off[1:] = np.cumsum([g.shape[0] for g in gts])
        gt = np.concatenate(gts, 0) if off[-1] else np.zeros((0, 5), np.float32)
        tc = ops.make_target_cfg(self.bg_threshold, self.fg_threshold, True, self.ignored_threshold,
                                 box_encoding=self.box_encoding)
        labels, _, deltas, bbw = ops.anchor_target(anchors.contiguous(), torch.from_numpy(gt).to(anchors.device),
                                                   torch.from_numpy(off).to(anchors.device), tc)
        return labels.reshape(-1).long(), deltas.reshape(-1, 4), bbw.reshape(-1, 4)
