"""detbox.fused.stage_one — get_final_results_stage1 (reference libs/nets/model.py:474-497) as one launch sequence
on one stream with no host synchronisation:

    objectness -> top 9*max_dets of the batch -> gather + decode      _decode_and_choose_top_n_stage1  model.py:308-321
    -> objectness >= score_threshold (or the single best row)         _decoding_and_thresholding_stage1 :323-340
    -> per-image NMS                                                  _apply_nms_in_batch              :343-355
    -> (box, prob, class) rows per image                              to_Dets2 / to_Dets2_sigmoid       :405-469

The reference synchronises at every arrow (nonzero(), a Python loop over images with .nonzero() inside, D2H of
everything for the NumPy epilogue).  Here rows that fail the threshold are parked in a dummy NMS segment instead
of being removed, counts stay on the device, and the caller reads back one [cap,6] array and B+1 offsets — or
nothing, if the detections are consumed on the device.  Every stage is a libdetbox.so kernel; the whole sequence
can be captured in a CUDA graph.
"""
from __future__ import annotations

import torch

from . import ops, ops_post


def stage_one(rpn_box, rpn_prob, anchors, n_images, score_threshold=0.1, max_dets=100, overlap_threshold=0.5,
              activation="softmax", box_encoding="fastrcnn", per_image=False, nms_cmp=ops.NMS_GT_F32):
    """rpn_box [B*A,4], rpn_prob [B*A,C], anchors [A,4] CUDA tensors -> (dets float64 [cap,6], offsets int32 [B+1])
    on the device: image i's detections are dets[offsets[i]:offsets[i+1]] = (x1,y1,x2,y2,prob,class), in
    the reference's order (NMS keep order, then class).  top_n = 9 * max_dets over the whole batch as upstream
    (get_final_results_stage1 passes max_dets*3 to _decoding_and_thresholding_stage1, which triples it again:
    model.py:483, 326); per_image=True takes it inside every image instead — the sharded mode."""
    sel = ops_post.topk_decode(rpn_box, rpn_prob, anchors, int(max_dets) * 9, activation, box_encoding, per_image,
                               want=("probs", "img_ids", "scores"))
    seg32, seg64 = ops_post.threshold_segments(sel["scores"], sel["img_ids"], n_images, score_threshold)
    keep, n_keep = ops_post.nms(sel["boxes"], sel["scores"], float(overlap_threshold), nms_cmp, seg_ids=seg32, n_seg=int(n_images) + 1)
    return ops_post.to_dets(sel["boxes"], sel["probs"], seg64, n_images, activation, True, score_threshold,
                            keep=keep, n_keep=n_keep[-1:])


def split(dets, offsets):
    """(dets, offsets) device tensors -> list of per-image NumPy arrays [m_i,6] (one device->host copy each)."""
    off = offsets.cpu().numpy()
    if int(off[-1]) > dets.shape[0]:
        raise RuntimeError("to_dets output truncated: %d rows needed, capacity %d" % (int(off[-1]), dets.shape[0]))
    d = dets[:int(off[-1])].cpu().numpy()
    return [d[off[i]:off[i + 1]] for i in range(off.size - 1)]
