"""Device-tensor operators over the C ABI (include/detbox.h).

Every function takes CUDA torch tensors (float32 unless said otherwise), launches hand-written
sm_100a kernels from libdetbox.so on torch's CURRENT stream and returns CUDA tensors.  PyTorch is
used for device memory and streams only — there is no eager/torch fallback: a CPU tensor raises.
The reference-shaped wrappers (module paths, NumPy in/out, autograd Functions) live in
``libs/`` next to this file and call these.
"""
from __future__ import annotations

import ctypes

import numpy as np
import torch

from . import _lib
from ._lib import check, lib

ENC = {"fastrcnn": 0, "linear": 1, "simple": 2}
ACT = {"softmax": 0, "sigmoid": 1}
NMS_GE_F64, NMS_GE_F32, NMS_GT_F32 = 0, 1, 2
ROI_BIN_CENTRE, ROI_GRID_POINT, ROI_CROP_NORM = 0, 1, 2
POOL = {None: 0, "none": 0, "avg": 1, "max": 2}

_launches = 0          # kernels launched through this module (bench.py reports it)


def launches() -> int:
    return _launches


def _count(n: int):
    global _launches
    _launches += n


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _dev(t: torch.Tensor, dtype=torch.float32, name="tensor") -> torch.Tensor:
    if not isinstance(t, torch.Tensor) or not t.is_cuda:
        raise _lib.DetboxError("%s must be a CUDA tensor: detbox has no CPU path" % name)
    if t.dtype != dtype:
        t = t.to(dtype)
    return t.contiguous()


def _ptr(t) -> int:
    return 0 if t is None else t.data_ptr()


class _Workspace:
    """Grow-only per-(device, stream) scratch buffer handed to the *_workspace_bytes() APIs.

    A CUDA graph bakes the buffer's address into its kernel nodes, so a buffer that was handed out while a stream
    was capturing is never freed: when a later, larger request on the same (device, stream handle) replaces it, the
    old tensor is retired (kept alive here) and the graph keeps writing into memory nothing else owns."""

    def __init__(self):
        self.bufs = {}
        self.captured = set()
        self.retired = []

    def get(self, nbytes: int, device) -> torch.Tensor:
        key = (device.index, torch.cuda.current_stream(device).cuda_stream)
        buf = self.bufs.get(key)
        if buf is None or buf.numel() < nbytes:
            if buf is not None and key in self.captured:
                self.retired.append(buf)
                self.captured.discard(key)
            buf = torch.empty(int(nbytes * 1.25) + 1024, dtype=torch.uint8, device=device)
            self.bufs[key] = buf
        if torch.cuda.is_current_stream_capturing():
            self.captured.add(key)
        return buf


_ws = _Workspace()


# ---------------------------------------------------------------------------------- anchors ----
_cell_cache = {}       # anchor configuration -> (cell anchors on the device, per-level sizes)


def base_anchors(scales, ratios, base) -> np.ndarray:
    """libs/boxes/anchor.py:101-103 (host, float64 [n_ratios*n_scales, 4])."""
    s = np.ascontiguousarray(scales, np.float64).ravel()
    r = np.ascontiguousarray(ratios, np.float64).ravel()
    out = np.empty((r.size * s.size, 4), np.float64)
    check(lib().detbox_base_anchors_host(s.ctypes.data_as(_lib.P_double), s.size, r.ctypes.data_as(_lib.P_double),
                                         r.size, int(base), out.ctypes.data_as(_lib.P_double)), "base_anchors")
    return out


def make_pyramid_desc(n_cell, heights, widths):
    """detbox_pyramid_desc: the grid geometry anchor rows were expanded over (level -> y -> x -> cell anchor)."""
    d = _lib.PyramidDesc()
    d.n_levels = len(n_cell)
    if d.n_levels > _lib.MAX_LEVELS:
        raise _lib.DetboxError("at most %d pyramid levels" % _lib.MAX_LEVELS)
    for l, (n, h, w) in enumerate(zip(n_cell, heights, widths)):
        d.n_cell[l], d.height[l], d.width[l] = int(n), int(h), int(w)
    return d


def pyramid_desc(strides, ih, iw, anchor_scales, anchor_ratios, anchor_base, anchor_shift=((0.0, 0.0),), device="cuda"):
    """The descriptor of the rows anchor_pyramid(...) returns for the same arguments."""
    return anchor_pyramid(strides, ih, iw, anchor_scales, anchor_ratios, anchor_base, anchor_shift, device, desc_only=True)


def anchor_pyramid(strides, ih, iw, anchor_scales, anchor_ratios, anchor_base, anchor_shift=((0.0, 0.0),),
                   device="cuda", dtype=torch.float32, desc_only=False) -> torch.Tensor:
    """libs/boxes/anchor.py:30-44 on the device: [sum_l H_l*W_l*A, 4], level -> y -> x -> (shift, ratio, scale)."""
    key = (repr(list(strides)), int(ih), int(iw), repr(anchor_scales), repr(anchor_ratios), int(anchor_base),
           repr(np.asarray(anchor_shift, np.float64).tolist()), str(device))
    hit = _cell_cache.get(key)
    if hit is None:
        sh = np.asarray(anchor_shift, np.float64).reshape(-1, 2)
        cells, n_cell, hs, ws = [], [], [], []
        for i, stride in enumerate(strides):
            sc = anchor_scales[i] if isinstance(anchor_scales[i], (list, tuple, np.ndarray)) else anchor_scales
            base = base_anchors(sc, anchor_ratios, anchor_base)
            shift4 = np.hstack((sh, sh)) * stride                      # anchor.py:54-55
            cell = np.vstack([base + s for s in shift4])
            cells.append(cell); n_cell.append(cell.shape[0])
            hs.append(int(ih / stride)); ws.append(int(iw / stride))
        if len(_cell_cache) > 64:
            _cell_cache.clear()
        hit = _cell_cache[key] = (torch.from_numpy(np.vstack(cells)).to(device), n_cell, hs, ws)   # <= 1 KB, uploaded once
    cell_dev, n_cell, hs, ws = hit
    if desc_only:
        return make_pyramid_desc(n_cell, hs, ws)
    L = len(strides)
    total = sum(h * w * n for h, w, n in zip(hs, ws, n_cell))
    out = torch.empty((total, 4), dtype=dtype, device=cell_dev.device)
    arr = lambda v: (ctypes.c_int * L)(*[int(x) for x in v])
    check(lib().detbox_anchor_pyramid(cell_dev.data_ptr(), L, arr(n_cell), arr(hs), arr(ws), arr(strides),
                                      out.data_ptr(), 1 if dtype == torch.float64 else 0, _stream()), "anchor_pyramid")
    _count(1)
    return out


# -------------------------------------------------------------------------------------- IoU ----
def _pairwise(fn, boxes, query):
    b, q = _dev(boxes, torch.float64, "boxes")[:, :4].contiguous(), _dev(query, torch.float64, "query_boxes")[:, :4].contiguous()
    out = torch.empty((b.shape[0], q.shape[0]), dtype=torch.float64, device=b.device)
    check(fn(b.data_ptr(), b.shape[0], q.data_ptr(), q.shape[0], out.data_ptr(), _stream()), "pairwise")
    _count(1)
    return out


def bbox_overlaps(boxes, query_boxes):
    """libs/boxes/bbox.pyx:15-55: float64 [N,K] IoU, +1 convention, bit-exact."""
    return _pairwise(lib().detbox_bbox_overlaps_f64, boxes, query_boxes)


def bbox_intersections(boxes, query_boxes):
    """libs/boxes/bbox.pyx:57-94: inter / area(query)."""
    return _pairwise(lib().detbox_bbox_intersections_f64, boxes, query_boxes)


def bbox_exclude_ignored_areas(boxes, ignored, threshold):
    """libs/boxes/bbox.pyx:96-135 -> int64 [N]."""
    b, q = _dev(boxes, torch.float64)[:, :4].contiguous(), _dev(ignored, torch.float64)[:, :4].contiguous()
    out = torch.empty((b.shape[0],), dtype=torch.int64, device=b.device)
    check(lib().detbox_bbox_exclude_ignored_f64(b.data_ptr(), b.shape[0], q.data_ptr(), q.shape[0], float(threshold),
                                                out.data_ptr(), _stream()), "bbox_exclude_ignored_areas")
    _count(1)
    return out


# ---------------------------------------------------------------------------- encode/decode ----
def box_encode(anchors, gt_boxes, box_encoding="fastrcnn"):
    """libs/boxes/bbox_transform.py:11-41 / libs/layers/box.py:112-145."""
    a, g = _dev(anchors).view(-1, 4), _dev(gt_boxes)[:, :4].contiguous()
    out = torch.empty_like(a)
    check(lib().detbox_box_encode(a.data_ptr(), g.data_ptr(), a.shape[0], ENC[box_encoding], out.data_ptr(), _stream()), "box_encode")
    _count(1)
    return out


def box_decode(anchors, deltas, box_encoding="fastrcnn", im_shape=None):
    """libs/boxes/bbox_transform.py:43-76 (+ clip :78-91 when im_shape=(ih, iw)) / libs/layers/box.py:72-109."""
    a, d = _dev(anchors).view(-1, 4), _dev(deltas).view(-1, 4)
    out = torch.empty_like(d)
    ih, iw = (0, 0) if im_shape is None else (int(im_shape[0]), int(im_shape[1]))
    check(lib().detbox_box_decode(a.data_ptr(), d.data_ptr(), d.shape[0], ENC[box_encoding], ih, iw, out.data_ptr(), _stream()), "box_decode")
    _count(1)
    return out


def clip_boxes_(boxes, im_shape):
    """libs/boxes/bbox_transform.py:78-91, in place on a contiguous [n, 4k] float32 CUDA tensor."""
    if not (boxes.is_cuda and boxes.dtype == torch.float32 and boxes.is_contiguous()):
        raise _lib.DetboxError("clip_boxes_ needs a contiguous float32 CUDA tensor")
    n, k4 = boxes.shape
    check(lib().detbox_clip_boxes(boxes.data_ptr(), n, k4 // 4, int(im_shape[0]), int(im_shape[1]), _stream()), "clip_boxes")
    _count(1)
    return boxes


# ---------------------------------------------------------------------------- anchor target ----
def make_target_cfg(bg, fg, clobber=True, ignore_fraction=0.3, bbweights=(1., 1., 1., 1.), box_encoding="fastrcnn"):
    c = _lib.TargetCfg()
    c.bg_threshold, c.fg_threshold, c.ignore_fraction = float(bg), float(fg), float(ignore_fraction)
    c.clobber_positives, c.box_encoding = int(bool(clobber)), ENC[box_encoding]
    for i in range(4):
        c.bbweights[i] = float(bbweights[i])
    return c


def anchor_target(anchors, gt, gt_offsets, tcfg, return_counts=False, out=None, pyramid=None, format="reference"):
    """Batched libs/layers/anchor.py:17-145.  anchors [A,4]; gt [sumG,5] (images concatenated);
    gt_offsets int32 [B+1] (CUDA).  Returns labels [B,A], label_weights [B,A], bbox_targets [B,A,4],
    bbox_inside_weights [B,A,4] (+ counts int32 [B,2] = (#fg, #bg)).  `out`: the four tensors to write
    into (contiguous float32 CUDA tensors of those shapes) instead of fresh allocations.

    pyramid: the detbox_pyramid_desc of the anchors (pyramid_desc(...)): GT-0 targets then come from separable
    per-image tables (verified on the device, bit-identical results).
    format="compact": returns (packed uint8 [B,A], bbox_targets [B,A,4], counts int32 [B,2]) — 17 bytes per
    anchor instead of 40; expand_targets() rebuilds the reference's arrays on the host.  `out` = those three."""
    a = _dev(anchors).view(-1, 4)
    g = _dev(gt).view(-1, 5)
    off = _dev(gt_offsets, torch.int32, "gt_offsets")
    B, A, G = off.numel() - 1, a.shape[0], g.shape[0]
    dev = a.device
    if format not in ("reference", "compact"):
        raise _lib.DetboxError("format must be 'reference' or 'compact'")
    compact = format == "compact"
    L = lib()
    pd = ctypes.byref(pyramid) if pyramid is not None else None
    nbytes = L.detbox_anchor_target_ex_workspace_bytes(A, B, G, pd)
    ws = _ws.get(nbytes, dev)
    if compact:
        if out is not None:
            packed, bt, counts = out
            for t, shape, dt in ((packed, (B, A), torch.uint8), (bt, (B, A, 4), torch.float32), (counts, (B, 2), torch.int32)):
                if not (t.is_cuda and t.dtype == dt and t.is_contiguous() and tuple(t.shape) == shape):
                    raise _lib.DetboxError("anchor_target(format='compact', out=...): expected contiguous CUDA tensors uint8 %s, "
                                           "float32 %s, int32 %s" % ((B, A), (B, A, 4), (B, 2)))
        else:
            packed = torch.empty((B, A), dtype=torch.uint8, device=dev)
            bt = torch.empty((B, A, 4), dtype=torch.float32, device=dev)
            counts = torch.empty((B, 2), dtype=torch.int32, device=dev)
        check(L.detbox_anchor_target_ex(a.data_ptr(), A, pd, g.data_ptr() if G else 0, off.data_ptr(), B, G, ctypes.byref(tcfg), 1,
                                        0, 0, bt.data_ptr(), 0, packed.data_ptr(), counts.data_ptr(),
                                        ws.data_ptr(), ws.numel(), _stream()), "anchor_target")
        _count((3 if G else 1) if A and B else 0)
        return packed, bt, counts
    if out is not None:
        labels, lw, bt, bw = out
        for t, shape in ((labels, (B, A)), (lw, (B, A)), (bt, (B, A, 4)), (bw, (B, A, 4))):
            if not (t.is_cuda and t.dtype == torch.float32 and t.is_contiguous() and tuple(t.shape) == shape):
                raise _lib.DetboxError("anchor_target(out=...): expected contiguous float32 CUDA tensors of shape %s" % (shape,))
    else:
        labels = torch.empty((B, A), dtype=torch.float32, device=dev)
        lw = torch.empty((B, A), dtype=torch.float32, device=dev)
        bt = torch.empty((B, A, 4), dtype=torch.float32, device=dev)
        bw = torch.empty((B, A, 4), dtype=torch.float32, device=dev)
    counts = torch.empty((B, 2), dtype=torch.int32, device=dev) if return_counts else None
    check(L.detbox_anchor_target_ex(a.data_ptr(), A, pd, g.data_ptr() if G else 0, off.data_ptr(), B, G, ctypes.byref(tcfg), 0,
                                    labels.data_ptr(), lw.data_ptr(), bt.data_ptr(), bw.data_ptr(), 0, _ptr(counts),
                                    ws.data_ptr(), ws.numel(), _stream()), "anchor_target")
    _count((4 if G else 2) if A and B else 0)
    if return_counts:
        return labels, lw, bt, bw, counts
    return labels, lw, bt, bw


def pack_padded_gt(gt_padded, gt_counts):
    """[B, Gmax, 5] padded GT + int32 [B] counts (CUDA) -> (gt [B*Gmax, 5] buffer whose first offsets[B] rows are the
    concatenated GT, offsets int32 [B+1]) for anchor_target, on the device (no host synchronisation: the row count
    stays on the device, so the full-size buffer is handed on and total_gt = B*Gmax bounds the work)."""
    g = _dev(gt_padded)
    if g.dim() != 3 or g.shape[2] != 5:
        raise _lib.DetboxError("pack_padded_gt: gt_padded must be [B, Gmax, 5]")
    c = _dev(gt_counts, torch.int32, "gt_counts").view(-1)
    B, gmax = g.shape[0], g.shape[1]
    if c.numel() != B:
        raise _lib.DetboxError("pack_padded_gt: %d images but %d counts" % (B, c.numel()))
    out = torch.empty((B * gmax, 5), dtype=torch.float32, device=g.device)
    off = torch.empty((B + 1,), dtype=torch.int32, device=g.device)
    check(lib().detbox_pack_padded_gt(g.data_ptr(), c.data_ptr(), B, gmax, out.data_ptr(), off.data_ptr(), _stream()), "pack_padded_gt")
    _count(2 if gmax else 1)
    return out, off


def expand_targets(packed, counts, bbweights=(1., 1., 1., 1.), n_threads=None, out=None):
    """HOST: the compact anchor-target format -> the reference's labels [B,A], label_weights [B,A],
    bbox_inside_weights [B,A,4] (libs/layers/anchor.py:111-118,136-138), bit for bit.  packed uint8 [B,A] and
    counts int32 [B,2] are CPU tensors / arrays (e.g. pinned buffers a D2H copy filled); bbox_targets needs no work."""
    pk = packed if isinstance(packed, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(packed))
    ct = counts if isinstance(counts, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(counts, np.int32))
    if pk.is_cuda or ct.is_cuda or pk.dtype != torch.uint8 or ct.dtype != torch.int32:
        raise _lib.DetboxError("expand_targets takes HOST uint8 / int32 arrays (copy the device results first)")
    pk, ct = pk.contiguous(), ct.contiguous()
    B, A = pk.shape
    if out is None:
        out = (torch.empty((B, A), dtype=torch.float32), torch.empty((B, A), dtype=torch.float32),
               torch.empty((B, A, 4), dtype=torch.float32))
    labels, lw, bw = out
    w = (ctypes.c_float * 4)(*[float(x) for x in bbweights])
    if n_threads is None:
        import os
        n_threads = min(16, len(os.sched_getaffinity(0)))
    check(lib().detbox_expand_targets_host(pk.data_ptr(), ct.data_ptr(), A, B, w, labels.data_ptr(), lw.data_ptr(),
                                           bw.data_ptr(), int(n_threads)), "expand_targets")
    return labels, lw, bw
