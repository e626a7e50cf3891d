"""Post-processing operators: stable sort, objectness top-k + decode, NMS (device tensors in/out).
See ops.py for the conventions."""
from __future__ import annotations

import ctypes
import os

import torch

from . import _lib
from ._lib import check, lib
from .ops import ACT, ENC, NMS_GE_F32, NMS_GE_F64, NMS_GT_F32, _count, _dev, _ptr, _stream, _ws  # noqa: F401


def sort_desc(scores):
    """Stable descending argsort of float32 scores (ties: lower index first) -> int32 [n].
    The explicit sorts of libs/layers/box.py:160 and libs/nms/gpu_nms.pyx:25."""
    s = _dev(scores).view(-1)
    n = s.numel()
    order = torch.empty((n,), dtype=torch.int32, device=s.device)
    if n == 0:
        return order
    ws = _ws.get(lib().detbox_sort_desc_workspace_bytes(n), s.device)
    check(lib().detbox_sort_desc(s.data_ptr(), n, order.data_ptr(), ws.data_ptr(), ws.numel(), _stream()), "sort_desc")
    _count(1 if n <= 4096 else 13)           # one-block bitonic sort, else keys + 4 x (count, scan, scatter)
    return order


def topk_decode(rpn_box, rpn_prob, anchors, top_n, activation="softmax", box_encoding="fastrcnn",
                per_image=False, want=("probs", "img_ids", "anchors")):
    """detection_model._decode_and_choose_top_n_stage1 (libs/nets/model.py:308-321).
    rpn_box [B*A,4], rpn_prob [B*A,C], anchors [A,4].  Returns a dict with 'boxes' [K,4] and, as
    requested in `want`, 'probs' [K,C], 'img_ids' int64 [K], 'anchors' [K,4], 'flat_idx' int64 [K],
    'scores' [K]; K = min(top_n, B*A) (per_image: B*min(top_n, A), image-major).
    rpn_box may be a PINNED HOST float32 tensor: only the selected rows are gathered, straight over PCIe through the
    buffer's device-side address, instead of copying every row to the device first."""
    prob, anc = _dev(rpn_prob), _dev(anchors).view(-1, 4)
    if isinstance(rpn_box, torch.Tensor) and not rpn_box.is_cuda:
        if not (rpn_box.is_pinned() and rpn_box.dtype == torch.float32 and rpn_box.is_contiguous()):
            raise _lib.DetboxError("topk_decode: a host rpn_box must be a pinned contiguous float32 tensor (there is no CPU path)")
        box = rpn_box.view(-1, 4)
        mapped = ctypes.c_void_p()
        check(lib().detbox_host_device_pointer(box.data_ptr(), ctypes.byref(mapped)), "host_device_pointer")
        box_ptr = mapped.value
    else:
        box = _dev(rpn_box).view(-1, 4)
        box_ptr = box.data_ptr()
    if prob.dim() == 1:
        prob = prob.view(-1, 1)
    N, C, A = box.shape[0], prob.shape[1], anc.shape[0]
    if A == 0 or N % A or prob.shape[0] != N:
        raise _lib.DetboxError("topk_decode: rows of rpn_box/rpn_prob (%d/%d) must be a multiple of the anchors (%d)"
                               % (N, prob.shape[0], A))
    B = N // A
    K = B * min(int(top_n), A) if per_image else min(int(top_n), N)
    dev = prob.device
    out = {"boxes": torch.empty((K, 4), dtype=torch.float32, device=dev)}
    if "probs" in want:
        out["probs"] = torch.empty((K, C), dtype=torch.float32, device=dev)
    if "img_ids" in want:
        out["img_ids"] = torch.empty((K,), dtype=torch.int64, device=dev)
    if "anchors" in want:
        out["anchors"] = torch.empty((K, 4), dtype=torch.float32, device=dev)
    if "flat_idx" in want:
        out["flat_idx"] = torch.empty((K,), dtype=torch.int64, device=dev)
    if "scores" in want:
        out["scores"] = torch.empty((K,), dtype=torch.float32, device=dev)
    if K == 0:
        return out
    ws = _ws.get(lib().detbox_topk_decode_workspace_bytes(B, A, int(top_n), int(per_image)), dev)
    check(lib().detbox_topk_decode(prob.data_ptr(), C, box_ptr, anc.data_ptr(), B, A, ACT[activation], int(top_n),
                                   int(per_image), ENC[box_encoding], out["boxes"].data_ptr(), _ptr(out.get("probs")),
                                   _ptr(out.get("img_ids")), _ptr(out.get("anchors")), _ptr(out.get("flat_idx")),
                                   _ptr(out.get("scores")), ws.data_ptr(), ws.numel(), _stream()), "topk_decode")
    # kernels launched (memsets not counted): init, objectness, select, then
    #   k <= 2048: partition, then ONE tail kernel (candidate select + per-segment sort + gather/decode) -> 5
    #   k <= the one-block sort's limit (4096 by default): partition, candidate select, per-segment sort -> 7 with the gather
    #   else     : 2 x (refine, select), count / scan / write, 13-launch radix sort (+3 per image-id pass)
    k_seg = min(int(top_n), A if per_image else B * A)
    limit = int(os.environ.get('DETBOX_SEG_SORT_LIMIT', 4096))
    fused_tail = k_seg <= min(2048, limit) and 'DETBOX_TOPK_UNFUSED_TAIL' not in os.environ
    _count(5 if fused_tail else 7 if k_seg <= limit else 24 + (3 if per_image and B > 1 else 0))
    return out


NMS_VARIANTS = {None: 0, "plain": 0, "new": 1, "multianchor": 2}


def nms(boxes, scores, thresh, cmp_mode=NMS_GT_F32, seg_ids=None, n_seg=1, return_counts=False, presorted=False,
        variant=None, want_clusters=False):
    """Greedy NMS on the device.  boxes [n,>=4], scores [n] -> (keep int64 [n] buffer, n_keep int32
    [n_seg+1] on the device); no host synchronisation happens here.  keep[:n_keep[n_seg]] are the
    kept ORIGINAL row indices, segment-major, score-descending inside a segment.
    presorted=True: the rows already are in that processing order (segment ids non-decreasing, scores
    non-increasing inside a segment, boxes [n,4]) — e.g. topk_decode(per_image=True) output — and the
    sort + gather are skipped.
    variant='new' | 'multianchor': the pair rules of libs/boxes/nms.pyx:70-195; want_clusters=True adds
    nms_multianchor's cluster ids (int64 [n], original indices) as a third result."""
    b, s = _dev(boxes), _dev(scores).view(-1)
    n = b.shape[0]
    if s.numel() != n:
        raise _lib.DetboxError("nms: %d boxes but %d scores" % (n, s.numel()))
    seg, seg_i64 = None, False
    if seg_ids is not None and n_seg > 1:
        seg_i64 = torch.is_tensor(seg_ids) and seg_ids.dtype == torch.int64       # narrowed inside the library
        seg = _dev(seg_ids, torch.int64 if seg_i64 else torch.int32, "seg_ids").view(-1)
        if seg.numel() != n:
            raise _lib.DetboxError("nms: %d boxes but %d segment ids" % (n, seg.numel()))
    else:
        n_seg = 1
    keep = torch.empty((n,), dtype=torch.int64, device=b.device)
    n_keep = torch.empty((n_seg + 1,), dtype=torch.int32, device=b.device)          # zeroed by the library
    clusters = torch.empty((n,), dtype=torch.int64, device=b.device) if want_clusters else None
    if n == 0:
        n_keep.zero_()
        return (keep, n_keep, clusters) if want_clusters else (keep, n_keep)
    ws = _ws.get(lib().detbox_nms_workspace_bytes(n, n_seg), b.device)
    if presorted and b.shape[1] != 4:
        raise _lib.DetboxError("nms(presorted=True) needs boxes of shape [n,4]")
    mode = int(cmp_mode) | (0x100 if presorted else 0) | (0x200 if seg_i64 else 0)
    if NMS_VARIANTS[variant] or want_clusters:
        check(lib().detbox_nms_variant(b.data_ptr(), b.shape[1], s.data_ptr(), n, _ptr(seg), n_seg, float(thresh), mode,
                                       NMS_VARIANTS[variant], keep.data_ptr(), n_keep.data_ptr(), _ptr(clusters),
                                       ws.data_ptr(), ws.numel(), _stream()), "nms_variant")
    else:
        check(lib().detbox_nms(b.data_ptr(), b.shape[1], s.data_ptr(), n, _ptr(seg), n_seg, float(thresh), mode,
                               keep.data_ptr(), n_keep.data_ptr(), ws.data_ptr(), ws.numel(), _stream()), "nms")
    # launches: sort (1 for n <= 4096, else 13) + gather unless presorted; segment count/scan (+ id passes); mask, resolve, compact
    _count((0 if presorted else (2 if n <= 4096 and n_seg == 1 else 14)) + 3 + ((2 if presorted else 5) if n_seg > 1 else 0)
           + (1 if want_clusters else 0))
    return (keep, n_keep, clusters) if want_clusters else (keep, n_keep)


def threshold_segments(scores, img_ids, n_images, score_threshold):
    """detbox_threshold_segments: (seg int32 [n], seg int64 [n]) — the image id of rows with score >= threshold (or of
    the single best row when none passes, model.py:331-333), `n_images` for every other row.  No host sync."""
    s = _dev(scores).view(-1)
    ids = _dev(img_ids, torch.int64, "img_ids").view(-1)
    n = s.numel()
    if ids.numel() != n:
        raise _lib.DetboxError("threshold_segments: %d scores but %d image ids" % (n, ids.numel()))
    seg32 = torch.empty((n,), dtype=torch.int32, device=s.device)
    seg64 = torch.empty((n,), dtype=torch.int64, device=s.device)
    check(lib().detbox_threshold_segments(_ptr(s), _ptr(ids), n, int(n_images), float(score_threshold),
                                          _ptr(seg32), _ptr(seg64), _stream()), "threshold_segments")
    _count(1 if n else 0)
    return seg32, seg64


def to_dets(boxes, probs, img_ids, n_images, activation="softmax", multi_label=True, score_threshold=0.1,
            keep=None, n_keep=None, cap=None):
    """detbox_to_dets (model.py:357-469 on the device): rows keep[:n_keep] (or all rows) -> (dets float64 [cap,6],
    offsets int32 [n_images+1]) device tensors; image i owns dets[offsets[i]:offsets[i+1]].  `n_keep` is a device
    scalar (int32 tensor, e.g. nms()'s n_keep[-1:]) so nothing is read back here."""
    b = _dev(boxes)
    if b.dim() != 2 or b.shape[1] != 4:
        raise _lib.DetboxError("to_dets: boxes must be [n,4]")
    p = _dev(probs)
    p = p.view(b.shape[0], -1)
    ids = _dev(img_ids, torch.int64, "img_ids").view(-1)
    C = p.shape[1]
    k = _dev(keep, torch.int64, "keep").view(-1) if keep is not None else None
    m_max = k.numel() if k is not None else b.shape[0]
    nk = _dev(n_keep, torch.int32, "n_keep").view(-1) if n_keep is not None else None
    first = 1 if activation == "softmax" else 0
    per_row = max(C - first, 1) if multi_label else 1
    cap = int(cap) if cap is not None else max(m_max * per_row, 1)
    dets = torch.empty((cap, 6), dtype=torch.float64, device=b.device)
    off = torch.zeros((int(n_images) + 1,), dtype=torch.int32, device=b.device)
    check(lib().detbox_to_dets(_ptr(b), _ptr(p), C, _ptr(ids), _ptr(k), _ptr(nk), m_max, int(n_images), ACT[activation],
                               int(bool(multi_label)), float(score_threshold), dets.data_ptr(), cap, off.data_ptr(), _stream()),
          "to_dets")
    _count(1)
    return dets, off


def dets_cluster(dets, keep, clusters, n_keep=None):
    """detbox_dets_cluster (single_shot.py:401-415 on the device): score-weighted box voting.  dets float32 [n,5],
    keep int64 [k] / clusters int64 [n] as nms(..., variant="multianchor", want_clusters=True) returns them, `n_keep`
    an optional device int32 scalar (nms()'s n_keep[-1:]).  Returns float32 [k,5]; rows past n_keep are zero."""
    d = _dev(dets)
    if d.dim() != 2 or d.shape[1] != 5:
        raise _lib.DetboxError("dets_cluster: dets must be [n,5]")
    k = _dev(keep, torch.int64, "keep").view(-1)
    c = _dev(clusters, torch.int64, "clusters").view(-1)
    if c.numel() != d.shape[0]:
        raise _lib.DetboxError("dets_cluster: clusters must have one entry per row of dets")
    nk = _dev(n_keep, torch.int32, "n_keep").view(-1) if n_keep is not None else None
    out = torch.zeros((k.numel(), 5), dtype=torch.float32, device=d.device)
    check(lib().detbox_dets_cluster(_ptr(d), d.shape[0], _ptr(k), _ptr(nk), k.numel(), _ptr(c), out.data_ptr(), _stream()),
          "dets_cluster")
    _count(1)
    return out
