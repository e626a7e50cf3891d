"""libs.layers.box (reference libs/layers/box.py): torch-tensor box ops on the device."""
import numpy as np
import torch

from ... import ops, ops_post, ops_target
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
                        seg_ids=img_ids, n_seg=int(n_images))


def _image_major(image_inds):
    """Permutation that groups rows by image, stable (what `boxes_np[image_inds_np == i]` for i = 0..B-1
    produces); None when the rows already are image-major (the case after _apply_nms_in_batch)."""
    if image_inds.numel() < 2 or bool((image_inds[1:] >= image_inds[:-1]).all()):
        return None
    return torch.sort(image_inds, stable=True)[1]


def matching_box(boxes, image_inds, gt_boxes_list, bg_overlap_threshold=0.5, fg_overlap_threshold=0.6):
    """box.py:12-69: (match_labels int64, match_inds int32, match_boxes f32 [n,4]), rows grouped by image in
    image order.  Upstream copies the RoIs to the host, runs the Cython IoU per image and copies three
    arrays back; here it is one launch over the whole batch and nothing leaves the device."""
    if not boxes.is_cuda:
        raise DetboxError("matching_box: CPU tensors are not supported (detbox has no CPU path)")
    ids = image_inds.reshape(-1).long()
    assert boxes.shape[0] == ids.numel()
    order = _image_major(ids)
    if order is not None:
        boxes, ids = boxes[order], ids[order]
    gt, off = ops_target.pack_gt(gt_boxes_list, boxes.device)
    out = ops_target.match_boxes(boxes, ids, gt, off, bg_overlap_threshold, fg_overlap_threshold)
    labels, match, mbox = out["labels"], out["match_inds"], out["match_boxes"]
    inside = labels != -2                                  # rows of images outside gt_boxes_list are dropped upstream
    if not bool(inside.all()):
        labels, match, mbox = labels[inside], match[inside], mbox[inside]
    return labels, match, mbox


def jitter_boxes(boxes, rng=np.random):
    """box.py:295-307 (in place, four rand(num) draws in upstream's order)."""
    num = boxes.shape[0]
    ws = boxes[:, 2] - boxes[:, 0] + 1.0
    hs = boxes[:, 3] - boxes[:, 1] + 1.0
    boxes[:, 0] = boxes[:, 0] + (rng.rand(num) - 0.5) * 0.2 * ws
    boxes[:, 1] = boxes[:, 1] + (rng.rand(num) - 0.5) * 0.2 * hs
    boxes[:, 2] = boxes[:, 2] + (rng.rand(num) - 0.5) * 0.2 * ws
    boxes[:, 3] = boxes[:, 3] + (rng.rand(num) - 0.5) * 0.2 * hs
    return boxes


def sample_rois(boxes, image_inds, gt_boxes_list, fg_overlap_threshold=0.5, rois_per_image=512, fg_fraction=0.25,
                ignore_threshold=0.2, rng=np.random):
    """box.py:175-292: drop small RoIs, label by max IoU, mark RoIs inside ignore regions, append the jittered
    GT boxes, keep label >= 0, then draw at most fg_fraction foregrounds and fill with backgrounds.

    The RoIs never leave the device: labelling, the ordered compaction and the final gather are kernels.
    The random draws stay on the host with upstream's NumPy calls in upstream's order (`rng` defaults to the
    global np.random, so a seeded run reproduces upstream's selection); they need the fg / bg counts, which
    is the one 12-byte device->host read.  GT boxes (tens of rows, host lists) are jittered on the host."""
    if not boxes.is_cuda:
        raise DetboxError("sample_rois: CPU tensors are not supported (detbox has no CPU path)")
    dev = boxes.device
    ids = image_inds.reshape(-1).long()
    assert boxes.shape[0] == ids.numel()
    boxes0, ids0 = boxes, ids
    order = _image_major(ids)
    if order is not None:
        boxes, ids = boxes[order], ids[order]
    batch_size = len(gt_boxes_list)
    gts = [np.asarray(g, np.float32).reshape(-1, 5) for g in gt_boxes_list]
    gt, off = ops_target.pack_gt(gts, dev)
    labels = ops_target.match_boxes(boxes, ids, gt, off, 0.0, fg_overlap_threshold, ignore_threshold, 8.0,
                                    want=("labels",))["labels"]
    extra, extra_lab, eoff = [], [], np.zeros(batch_size + 1, np.int32)
    for i, g in enumerate(gts):
        if g.size > 0:                                       # box.py:223-236
            valid = np.where(g[:, 4] > 0)[0]
            extra.append(jitter_boxes(g[valid][:, :4].astype(np.float32), rng))
            extra_lab.append(g[valid][:, 4].astype(np.int64))
        eoff[i + 1] = eoff[i] + (extra[-1].shape[0] if g.size > 0 else 0)
    eb = np.concatenate(extra, 0).astype(np.float32) if extra else np.zeros((0, 4), np.float32)
    el = np.concatenate(extra_lab, 0) if extra_lab else np.zeros((0,), np.int64)
    sb, sl, si, fg_pos, bg_pos, counts = ops_target.sample_compact(
        boxes[:, :4], ids, labels, torch.from_numpy(eb).to(dev), torch.from_numpy(el).to(dev), torch.from_numpy(eoff).to(dev))
    total, n_fg_all, n_bg_all = (int(v) for v in counts[:3].tolist())
    fg_choice = np.arange(n_fg_all)
    bg_choice = np.arange(n_bg_all)
    num_fg = min(n_fg_all, int(fg_fraction * rois_per_image * batch_size))            # box.py:265-275
    if num_fg > 0:
        fg_choice = rng.choice(fg_choice, num_fg, replace=False)
    num_bg = min(rois_per_image * batch_size - num_fg, n_bg_all)
    if n_bg_all > 0:
        bg_choice = rng.choice(bg_choice, num_bg, replace=False)
    if fg_choice.size + bg_choice.size == 0:                                          # box.py:282-286
        return boxes0[:1, :], torch.full((1,), -1, dtype=torch.int64, device=dev), ids0[:1]
    return ops_target.sample_gather(sb, sl, si, fg_pos, bg_pos, torch.from_numpy(np.asarray(fg_choice, np.int64)).to(dev),
                                    torch.from_numpy(np.asarray(bg_choice, np.int64)).to(dev))
