"""libs.layers.box (reference libs/layers/box.py): torch-tensor box ops on the device."""
import torch

from ... import ops, ops_post
from ..._lib import DetboxError


def decoding_box(deltas, anchors, box_encoding='fastrcnn'):
    """box.py:72-109 ('fastrcnn' | 'simple'; 'linear' also accepted): new [n,4] tensor, no clip."""
    return ops.box_decode(anchors.reshape(-1, 4), deltas.reshape(-1, 4), box_encoding)


def encoding_box(gt_boxes, anchors, box_encoding='fastrcnn'):
    """box.py:112-145."""
    return ops.box_encode(anchors, gt_boxes, box_encoding)


def apply_nms(boxes, scores, overlap_threshold):
    """box.py:148-173: boxes [n,4], scores [n] (CUDA) -> LongTensor [m] of kept ORIGINAL indices in
    descending score order, on the input's device.  The comparator is the reference's CUDA path
    (iou > thr, nms_cuda_kernel.cu:63).  One host read (the count) sizes the result — upstream
    copies the whole n x n/64 mask to the host instead (nms_cuda.c:33-59)."""
    if not boxes.is_cuda:
        raise DetboxError("apply_nms: CPU tensors are not supported (detbox has no CPU path)")
    keep, n_keep = ops_post.nms(boxes[:, :4], scores.reshape(-1), float(overlap_threshold), ops.NMS_GT_F32)
    return keep[:int(n_keep[-1].item())]


def apply_nms_in_batch(boxes, scores, img_ids, n_images, overlap_threshold=0.5, cmp_mode=ops.NMS_GT_F32):
    """The per-image loop of detection_model._apply_nms_in_batch (libs/nets/model.py:343-355) as ONE
    segmented launch sequence: rows of different images never suppress each other.  Returns
    (keep int64 [n] buffer, n_keep int32 [n_images+1]) on the device, without host synchronisation:
    keep[:n_keep[-1]] is image-major, score-descending inside an image."""
    return ops_post.nms(boxes[:, :4], scores.reshape(-1), float(overlap_threshold), cmp_mode,
                        seg_ids=img_ids.to(torch.int32), n_seg=int(n_images))
