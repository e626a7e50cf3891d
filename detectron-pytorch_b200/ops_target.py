"""RoI -> ground-truth matching, RoI targets and RoI sampling on the device (roi_target.cu).
torch supplies device memory and the current stream; every computation is a libdetbox.so kernel."""
from __future__ import annotations

import ctypes

import numpy as np
import torch

from . import _lib
from ._lib import check, lib
from .ops import _count, _dev, _ptr, _stream


def pack_gt(gt_boxes_list, device):
    """list of [G_i,5] arrays (NumPy or tensors) -> (gt [sum G,5] f32, offsets int32 [B+1]) on `device`."""
    gts = [np.asarray(g.detach().cpu() if torch.is_tensor(g) else g, np.float32).reshape(-1, 5) for g in gt_boxes_list]
    off = np.zeros(len(gts) + 1, np.int32)
    off[1:] = np.cumsum([g.shape[0] for g in gts])
    gt = np.concatenate(gts, 0) if off[-1] else np.zeros((0, 5), np.float32)
    return torch.from_numpy(gt).to(device), torch.from_numpy(off).to(device)


def match_boxes(boxes, image_inds, gt, gt_offsets, bg_thresh=0.5, fg_thresh=0.6, ignore_thresh=None, min_size=0.0,
                want=("labels", "match_inds", "match_boxes")):
    """detbox_match_boxes: per-row (labels int64, match_inds int32, match_boxes [R,4]) in input order."""
    b = _dev(boxes)
    if b.dim() != 2 or b.shape[1] < 4:
        raise _lib.DetboxError("boxes must be [R, >=4]")
    ids = _dev(image_inds, torch.int64, "image_inds").view(-1)
    g = _dev(gt).view(-1, 5)
    off = _dev(gt_offsets, torch.int32, "gt_offsets").view(-1)
    R, B = b.shape[0], off.numel() - 1
    if ids.numel() != R:
        raise _lib.DetboxError("image_inds must have one entry per box")
    cfg = _lib.MatchCfg(float(bg_thresh), float(fg_thresh), -1.0 if ignore_thresh is None else float(ignore_thresh), float(min_size))
    out = {
        "labels": torch.empty((R,), dtype=torch.int64, device=b.device) if "labels" in want else None,
        "match_inds": torch.empty((R,), dtype=torch.int32, device=b.device) if "match_inds" in want else None,
        "match_boxes": torch.empty((R, 4), dtype=torch.float32, device=b.device) if "match_boxes" in want else None,
    }
    check(lib().detbox_match_boxes(b.data_ptr(), b.stride(0), ids.data_ptr(), R, _ptr(g), off.data_ptr(), B, ctypes.byref(cfg),
                                   _ptr(out["labels"]), _ptr(out["match_inds"]), _ptr(out["match_boxes"]), _stream()),
          "match_boxes")
    _count(1 if R else 0)
    return out


def roi_target(rois, roi_batch_inds, gt_boxes, gt_batch_inds, fg_overlap, bg_overlap=0.0):
    """detbox_roi_target == roi_target_forward_cuda: (labels int64 [R], deltas [R,4], bbwgts [R,4])."""
    r = _dev(rois)
    if r.dim() != 2 or r.shape[1] != 4:
        raise _lib.DetboxError("rois should have Nx4 dims (x1, y1, x2, y2)")         # roi_target_cuda.c:42-47
    ri = _dev(roi_batch_inds, torch.int64, "roi_batch_inds").view(-1)
    g = _dev(gt_boxes)
    if g.numel() and (g.dim() != 2 or g.shape[1] != 5):
        raise _lib.DetboxError("gt_boxes should have Gx5 dims (x1, y1, x2, y2, cls)")  # roi_target_cuda.c:56-61
    g = g.view(-1, 5)
    gi = _dev(gt_batch_inds, torch.int64, "gt_batch_inds").view(-1)
    if ri.numel() != r.shape[0] or gi.numel() != g.shape[0]:
        raise _lib.DetboxError("batch indices and boxes should have the same number of rows")
    R = r.shape[0]
    labels = torch.empty((R,), dtype=torch.int64, device=r.device)
    deltas = torch.empty((R, 4), dtype=torch.float32, device=r.device)
    bbw = torch.empty((R, 4), dtype=torch.float32, device=r.device)
    check(lib().detbox_roi_target(r.data_ptr(), 4, ri.data_ptr(), R, _ptr(g), _ptr(gi), g.shape[0], float(bg_overlap), float(fg_overlap),
                                  labels.data_ptr(), deltas.data_ptr(), bbw.data_ptr(), _stream()), "roi_target")
    _count(1 if R else 0)
    return labels, deltas, bbw


def sample_compact(boxes, image_inds, labels, extra_boxes, extra_labels, extra_offsets):
    """detbox_sample_compact: (boxes, labels, image_inds, fg_pos, bg_pos, counts) device buffers sized R + E."""
    b = _dev(boxes)
    ids = _dev(image_inds, torch.int64, "image_inds").view(-1)
    lab = _dev(labels, torch.int64, "labels").view(-1)
    eb = _dev(extra_boxes).view(-1, 4)
    el = _dev(extra_labels, torch.int64, "extra_labels").view(-1)
    eo = _dev(extra_offsets, torch.int32, "extra_offsets").view(-1)
    R, E, B = b.shape[0], eb.shape[0], eo.numel() - 1
    n = max(R + E, 1)
    dev = eo.device
    ob = torch.empty((n, 4), dtype=torch.float32, device=dev)
    ol = torch.empty((n,), dtype=torch.int64, device=dev)
    oi = torch.empty((n,), dtype=torch.int64, device=dev)
    fg = torch.empty((n,), dtype=torch.int32, device=dev)
    bg = torch.empty((n,), dtype=torch.int32, device=dev)
    counts = torch.zeros((4,), dtype=torch.int32, device=dev)
    check(lib().detbox_sample_compact(_ptr(b), b.stride(0) if R else 4, _ptr(ids), _ptr(lab), R, _ptr(eb), _ptr(el), eo.data_ptr() if E else 0, B,
                                      ob.data_ptr(), ol.data_ptr(), oi.data_ptr(), fg.data_ptr(), bg.data_ptr(), counts.data_ptr(),
                                      _stream()), "sample_compact")
    _count(1)
    return ob, ol, oi, fg, bg, counts


def sample_gather(boxes, labels, image_inds, fg_pos, bg_pos, fg_choice, bg_choice):
    """detbox_sample_gather: rows fg_pos[fg_choice] then bg_pos[bg_choice] of the compacted arrays."""
    dev = boxes.device
    fc = _dev(fg_choice, torch.int64, "fg_choice").view(-1)
    bc = _dev(bg_choice, torch.int64, "bg_choice").view(-1)
    n = fc.numel() + bc.numel()
    ob = torch.empty((n, 4), dtype=torch.float32, device=dev)
    ol = torch.empty((n,), dtype=torch.int64, device=dev)
    oi = torch.empty((n,), dtype=torch.int64, device=dev)
    check(lib().detbox_sample_gather(boxes.data_ptr(), labels.data_ptr(), image_inds.data_ptr(), fg_pos.data_ptr(), bg_pos.data_ptr(),
                                     _ptr(fc), fc.numel(), _ptr(bc), bc.numel(), ob.data_ptr(), ol.data_ptr(), oi.data_ptr(), _stream()),
          "sample_gather")
    _count(1 if n else 0)
    return ob, ol, oi
