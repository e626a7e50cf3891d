"""CPU oracle for the box-and-RoI hot path: ctypes bindings over ``liboracle.so``
(``detbox_oracle.c``, the plain-C restatement) plus the NumPy glue that restates the reference's
Python-level composition (top-k, batch NMS, pyramid wrappers).

TEST INFRASTRUCTURE ONLY.  Imported by ``tests/``, ``__graft_entry__.smoke()`` and the
``cpu_baseline`` / ``--impl reference`` legs of ``bench.py`` — never by the product package.
Every function cites the reference file:line it follows (paths relative to the reference root).

Tie rule (the reference's ``argsort()[::-1]`` / ``torch.sort`` are unstable, i.e. ties have no
defined order upstream): equal scores are ordered by ascending original index.
"""
from __future__ import annotations

import ctypes
import os
import subprocess
from dataclasses import dataclass, field

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

c_f = ctypes.POINTER(ctypes.c_float)
c_d = ctypes.POINTER(ctypes.c_double)
c_i32 = ctypes.POINTER(ctypes.c_int32)
c_i64 = ctypes.POINTER(ctypes.c_int64)


def build(force: bool = False) -> str:
    so = os.path.join(HERE, "liboracle.so")
    src = os.path.join(HERE, "detbox_oracle.c")
    if force or not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", HERE, "liboracle.so"], stdout=subprocess.DEVNULL)
    return so


def lib():
    global _LIB
    if _LIB is None:
        _LIB = ctypes.CDLL(build())
        _LIB.orc_nms.restype = ctypes.c_long
    return _LIB


def _p(a, t):
    return a.ctypes.data_as(t)


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


# ------------------------------------------------------------------------------- config ----
@dataclass
class Cfg:
    """The hot-path keys of libs/configs/config.py (defaults of config.py:95-172)."""
    strides: tuple = (8, 16, 32, 64, 128)
    anchor_scales: list = field(default_factory=lambda: [[2, 2.52, 3.17], [4, 5.04, 6.35],
                                                         [8, 10.08, 12.70], [16, 20.16, 25.40],
                                                         [32, 40.32, 50.80]])
    anchor_ratios: list = field(default_factory=lambda: [0.5, 1, 2.0])
    anchor_base: int = 16
    anchor_shift: list = field(default_factory=lambda: [[0.0, 0.0]])
    rpn_bg_threshold: float = 0.4
    rpn_fg_threshold: float = 0.6
    rpn_clobber_positives: bool = True
    rpn_sample_strategy: str = "simple"
    rpn_box_encoding: str = "fastrcnn"
    bbweights: list = field(default_factory=lambda: [1., 1., 1., 1.])
    ignored_area_intersection_fraction: float = 0.3
    num_classes: int = 81


def citypersons_retinanet_cfg() -> Cfg:
    """exper/citypersons/retinanet/train_test_resnet50_softmax.yaml:19-46 (tag CP-R)."""
    return Cfg(strides=(8, 16, 32, 64, 128),
               anchor_scales=[[2, 2.520, 3.175], [4, 5.040, 6.349], [8, 10.079, 12.700],
                              [16, 20.159, 25.398], [32, 40.317, 50.797]],
               anchor_ratios=[1, 2.5], anchor_base=16,
               rpn_bg_threshold=0.5, rpn_fg_threshold=0.6, num_classes=2)


ENCODINGS = {"fastrcnn": 0, "linear": 1, "simple": 2}


# ------------------------------------------------------------------------------ anchors ----
def anchors(scales, ratios, base):
    """libs/boxes/anchor.py:101-103 -> float64 [len(ratios)*len(scales), 4]."""
    s, r = _f64(scales), _f64(ratios)
    out = np.empty((r.size * s.size, 4), np.float64)
    lib().orc_base_anchors(_p(s, c_d), int(s.size), _p(r, c_d), int(r.size),
                           ctypes.c_long(int(base)), _p(out, c_d))
    return out


def anchor_one_plane(height, width, stride, anchor_scales, anchor_ratios, anchor_base, anchor_shifts):
    """libs/boxes/anchor.py:47-70."""
    base = anchors(anchor_scales, anchor_ratios, anchor_base)
    sh = _f64(anchor_shifts).reshape(-1, 2)
    out = np.empty((height * width * sh.shape[0] * base.shape[0], 4), np.float64)
    lib().orc_anchor_plane(_p(base, c_d), base.shape[0], _p(sh, c_d), sh.shape[0],
                           int(height), int(width), ctypes.c_long(int(stride)), _p(out, c_d))
    return out


def anchor_pyramid(strides, ih, iw, anchor_scales, anchor_ratios, anchor_base, anchor_shift=((0.0, 0.0),)):
    """libs/boxes/anchor.py:30-44: levels vstacked fine -> coarse, float64."""
    planes = []
    for i, stride in enumerate(strides):
        h, w = int(ih / stride), int(iw / stride)
        sc = anchor_scales[i] if isinstance(anchor_scales[i], (list, tuple)) else anchor_scales
        planes.append(anchor_one_plane(h, w, stride, sc, anchor_ratios, anchor_base, anchor_shift))
    return np.vstack(planes)


# ---------------------------------------------------------------------------------- IoU ----
def bbox_overlaps(boxes, query_boxes):
    """libs/boxes/bbox.pyx:15-55 -> float64 [N,K]."""
    b, q = _f64(boxes), _f64(query_boxes)
    out = np.zeros((b.shape[0], q.shape[0]), np.float64)
    lib().orc_bbox_overlaps(_p(b, c_d), ctypes.c_long(b.shape[0]), _p(q, c_d), ctypes.c_long(q.shape[0]), _p(out, c_d))
    return out


def bbox_intersections(boxes, query_boxes):
    """libs/boxes/bbox.pyx:57-94 -> float64 [N,K] = inter / area(query_k)."""
    b, q = _f64(boxes), _f64(query_boxes)
    out = np.zeros((b.shape[0], q.shape[0]), np.float64)
    lib().orc_bbox_intersections(_p(b, c_d), ctypes.c_long(b.shape[0]), _p(q, c_d), ctypes.c_long(q.shape[0]), _p(out, c_d))
    return out


def bbox_exclude_ignored_areas(boxes, ignored_areas, threshold):
    """libs/boxes/bbox.pyx:96-135 -> int64 [N]."""
    b, q = _f64(boxes), _f64(ignored_areas)
    out = np.zeros((b.shape[0],), np.int64)
    lib().orc_bbox_exclude_ignored_areas(_p(b, c_d), ctypes.c_long(b.shape[0]), _p(q, c_d),
                                         ctypes.c_long(q.shape[0]), ctypes.c_double(threshold), _p(out, c_i64))
    return out


# ------------------------------------------------------------------------ encode/decode ----
def bbox_transform(ex_rois, gt_rois, box_encoding="fastrcnn"):
    """libs/boxes/bbox_transform.py:11-41 / :93-116; torch twin libs/layers/box.py:112-145. f32."""
    a, g = _f32(ex_rois), _f32(gt_rois)[:, :4].copy()
    out = np.empty((a.shape[0], 4), np.float32)
    lib().orc_box_encode(_p(a, c_f), _p(g, c_f), ctypes.c_long(a.shape[0]), ENCODINGS[box_encoding], _p(out, c_f))
    return out


def bbox_transform_inv(boxes, deltas, box_encoding="fastrcnn", im_shape=None):
    """libs/boxes/bbox_transform.py:43-76 (+ clip_boxes :78-91 when im_shape=(ih,iw));
    torch twin decoding_box libs/layers/box.py:72-109 (no clip).  f32, [n,4] deltas."""
    a, d = _f32(boxes), _f32(deltas)
    out = np.empty((a.shape[0], 4), np.float32)
    ih, iw = (0, 0) if im_shape is None else (int(im_shape[0]), int(im_shape[1]))
    lib().orc_box_decode(_p(a, c_f), _p(d, c_f), ctypes.c_long(a.shape[0]), ENCODINGS[box_encoding], ih, iw, _p(out, c_f))
    return out


# ------------------------------------------------------------------------ anchor target ----
def anchor_encode(gt_boxes, all_anchors, cfg: Cfg, diagnostics: bool = False):
    """libs/layers/anchor.py:17-145 ('simple' strategy) -> (labels, label_weights,
    bbox_targets, bbox_inside_weights) float32; with diagnostics=True also
    (max_overlaps f64, gt_assignment i32, ignore_sum f64)."""
    if cfg.rpn_sample_strategy != "simple":
        raise ValueError("oracle restates the deterministic 'simple' strategy only")
    anc = _f32(all_anchors).reshape(-1, 4)
    gt = _f32(gt_boxes).reshape(-1, 5)
    A, G = anc.shape[0], gt.shape[0]
    labels = np.empty((A,), np.float32)
    lw = np.empty((A,), np.float32)
    bt = np.empty((A, 4), np.float32)
    bw = np.empty((A, 4), np.float32)
    mo = np.empty((A,), np.float64)
    asg = np.empty((A,), np.int32)
    isum = np.empty((A,), np.float64)
    bbw = _f32(cfg.bbweights)
    lib().orc_anchor_encode(_p(gt, c_f), G, _p(anc, c_f), ctypes.c_long(A),
                            ctypes.c_double(cfg.rpn_bg_threshold), ctypes.c_double(cfg.rpn_fg_threshold),
                            int(bool(cfg.rpn_clobber_positives)),
                            ctypes.c_double(cfg.ignored_area_intersection_fraction),
                            _p(bbw, c_f), ENCODINGS[cfg.rpn_box_encoding],
                            _p(labels, c_f), _p(lw, c_f), _p(bt, c_f), _p(bw, c_f),
                            _p(mo, c_d), _p(asg, c_i32), _p(isum, c_d))
    if diagnostics:
        return labels, lw, bt, bw, mo, asg, isum
    return labels, lw, bt, bw


def anchor_decode(boxes, scores, all_anchors, ih, iw, num_classes, box_encoding="fastrcnn"):
    """libs/layers/anchor.py:148-176 -> (final_boxes f32 [R,4], classes i32 [R], scores [R,K])."""
    scores = np.asarray(scores).reshape(-1, num_classes)
    fb = bbox_transform_inv(np.asarray(all_anchors).reshape(-1, 4), np.asarray(boxes).reshape(-1, 4),
                            box_encoding, (ih, iw))
    return fb, np.argmax(scores, axis=1).astype(np.int32), scores


# ---------------------------------------------------------------------------------- NMS ----
NMS_GE_F64, NMS_GE_F32, NMS_GT_F32 = 0, 1, 2


def stable_desc_order(scores):
    """Descending score, ties by ascending index (our documented tie rule)."""
    s = np.asarray(scores)
    return np.argsort(-s.astype(np.float64), kind="stable").astype(np.int64)


def nms(boxes, scores, thresh, cmp_mode=NMS_GE_F64):
    """Greedy NMS; cmp_mode selects cpu_nms.pyx:65 (>=, f64 thr), nms_cpu.c:76 (>=, f32 thr) or
    nms_kernel.cu:83 (>, f32 thr).  Returns kept ORIGINAL indices (int64) in score order."""
    b = _f32(boxes)
    n = b.shape[0]
    if n == 0:
        return np.zeros((0,), np.int64)
    order = stable_desc_order(scores)
    keep = np.empty((n,), np.int64)
    nk = lib().orc_nms(_p(b, c_f), ctypes.c_long(b.shape[1]), _p(order, c_i64), ctypes.c_long(n),
                       ctypes.c_double(float(thresh)), int(cmp_mode), _p(keep, c_i64))
    return keep[:nk].copy()


def cpu_nms(dets, thresh):
    """libs/nms/cpu_nms.pyx:17-68: dets [n,5] f32 -> list of kept indices."""
    d = _f32(dets)
    return [int(i) for i in nms(d[:, :4], d[:, 4], thresh, NMS_GE_F64)]


def gpu_nms(dets, thresh):
    """libs/nms/gpu_nms.pyx:16-31 + nms_kernel.cu:46-155 ('>' comparator, f32 threshold)."""
    d = _f32(dets)
    return [int(i) for i in nms(d[:, :4], d[:, 4], thresh, NMS_GT_F32)]


# ------------------------------------------------------------------- top-k / stage one ----
def objectness(probs, activation):
    """libs/nets/model.py:54-61."""
    p = _f32(probs)
    out = np.empty((p.shape[0],), np.float32)
    lib().orc_objectness(_p(p, c_f), ctypes.c_long(p.shape[0]), p.shape[1], 0 if activation == "softmax" else 1, _p(out, c_f))
    return out


def decode_and_choose_top_n(rpn_box, rpn_prob, anchors_, top_n, activation="softmax",
                            box_encoding="fastrcnn", per_image=False):
    """libs/nets/model.py:308-321: one global descending sort over batch x levels x anchors,
    top_n, gather, decode (no clip).  per_image=True takes top_n inside every image instead
    (the sharded mode of SURVEY §8e; identical to the reference when the batch is 1).
    Returns (boxes [k,4] f32, probs [k,C] f32, img_ids [k] i64, anchors [k,4] f32, flat_idx i64)."""
    rpn_box, rpn_prob, anc = _f32(rpn_box), _f32(rpn_prob), _f32(anchors_)
    A = anc.shape[0]
    obj = objectness(rpn_prob, activation)
    if per_image:
        B = obj.shape[0] // A
        parts = [stable_desc_order(obj[b * A:(b + 1) * A])[:top_n] + b * A for b in range(B)]
        inds = np.concatenate(parts) if parts else np.zeros((0,), np.int64)
    else:
        inds = stable_desc_order(obj)[:top_n]
    anchor_ids = inds % A
    sel_anchors = anc[anchor_ids]
    boxes = bbox_transform_inv(sel_anchors, rpn_box[inds], box_encoding)
    return boxes, rpn_prob[inds], inds // A, sel_anchors, inds


def apply_nms_in_batch(boxes, probs, img_ids, activation, overlap_threshold=0.5, cmp_mode=NMS_GT_F32):
    """libs/nets/model.py:343-355: per image i in [min,max]: NMS over the image's rows; keep
    indices concatenated image-major, score-descending inside an image.  The reference's CUDA
    path (libs/layers/box.py:170-173 -> nms_cuda_kernel.cu:63) uses '>'; its CPU path '>='
    (nms_cpu.c:76)."""
    obj = objectness(probs, activation)
    img_ids = np.asarray(img_ids)
    keeps = []
    if img_ids.size:
        for i in range(int(img_ids.min()), int(img_ids.max()) + 1):
            inds = np.nonzero(img_ids == i)[0]
            if inds.size:
                k = nms(boxes[inds][:, :4], obj[inds], overlap_threshold, cmp_mode)
                keeps.append(inds[k])
    return np.concatenate(keeps) if keeps else np.zeros((0,), np.int64)


# ----------------------------------------------------------------------------- RoIAlign ----
def tf_boxes(rois, spatial_scale, ph, pw, H, W):
    """roi_align_tf.py:24-40: (b,x1,y1,x2,y2) image coords -> normalised (y1,x1,y2,x2) + int32 index."""
    r = _f32(rois).reshape(-1, 5)
    boxes = np.empty((r.shape[0], 4), np.float32)
    ind = np.empty((r.shape[0],), np.int32)
    lib().orc_tf_boxes(_p(r, c_f), ctypes.c_long(r.shape[0]), ctypes.c_float(spatial_scale), ph, pw,
                       int(H), int(W), _p(boxes, c_f), _p(ind, c_i32))
    return boxes, ind


def crop_and_resize_fwd(image, boxes, box_ind, extrapolation_value, crop_h, crop_w):
    """crop_and_resize.c:115-161: image NCHW f32, boxes normalised (y1,x1,y2,x2), box_ind int32."""
    f, b, ind = _f32(image), _f32(boxes), np.ascontiguousarray(box_ind, np.int32)
    B, C, H, W = f.shape
    out = np.empty((b.shape[0], C, crop_h, crop_w), np.float32)
    lib().orc_crop_and_resize_fwd(_p(f, c_f), B, C, H, W, _p(b, c_f), _p(ind, c_i32), ctypes.c_long(b.shape[0]),
                                  ctypes.c_float(extrapolation_value), crop_h, crop_w, _p(out, c_f))
    return out


def crop_and_resize_bwd(grads, boxes, box_ind, image_shape):
    """crop_and_resize.c:164-268 (serial scatter; gradient w.r.t. the image only)."""
    g, b, ind = _f32(grads), _f32(boxes), np.ascontiguousarray(box_ind, np.int32)
    B, C, H, W = image_shape
    gi = np.zeros((B, C, H, W), np.float32)
    lib().orc_crop_and_resize_bwd(_p(g, c_f), B, C, H, W, _p(b, c_f), _p(ind, c_i32), ctypes.c_long(b.shape[0]),
                                  g.shape[2], g.shape[3], _p(gi, c_f))
    return gi


def roi_align_tf_fwd(feat, rois, spatial_scale, ph, pw):
    """roi_align_tf.py:16-42 + crop_and_resize.c:6-161 (bin-centre, 1 sample per bin)."""
    f = _f32(feat)
    boxes, ind = tf_boxes(rois, spatial_scale, ph, pw, f.shape[2], f.shape[3])
    return crop_and_resize_fwd(f, boxes, ind, 0.0, ph, pw)


def roi_align_tf_bwd(grad_out, rois, spatial_scale, feat_shape):
    """CropAndResizeFunction.backward crop_and_resize.py:39-55 -> crop_and_resize.c:164-268."""
    g = _f32(grad_out)
    boxes, ind = tf_boxes(rois, spatial_scale, g.shape[2], g.shape[3], feat_shape[2], feat_shape[3])
    return crop_and_resize_bwd(g, boxes, ind, feat_shape)


def roi_align_grid_fwd(feat, rois, spatial_scale, ph, pw):
    """roi_align/src/cuda/roi_align_kernel.cu:15-68 (grid-point, both ends inclusive)."""
    f, r = _f32(feat), _f32(rois).reshape(-1, 5)
    B, C, H, W = f.shape
    out = np.empty((r.shape[0], C, ph, pw), np.float32)
    lib().orc_roi_align_grid_fwd(_p(f, c_f), B, C, H, W, _p(r, c_f), ctypes.c_long(r.shape[0]),
                                 ctypes.c_float(spatial_scale), ph, pw, _p(out, c_f))
    return out


def roi_align_grid_bwd(grad_out, rois, spatial_scale, feat_shape):
    """roi_align_kernel.cu:97-153."""
    g, r = _f32(grad_out), _f32(rois).reshape(-1, 5)
    B, C, H, W = feat_shape
    gi = np.zeros((B, C, H, W), np.float32)
    lib().orc_roi_align_grid_bwd(_p(g, c_f), B, C, H, W, _p(r, c_f), ctypes.c_long(r.shape[0]),
                                 ctypes.c_float(spatial_scale), g.shape[2], g.shape[3], _p(gi, c_f))
    return gi


def pool2x2(x, kind):
    """avg_pool2d / max_pool2d(kernel 2, stride 1) of roi_align/modules/roi_align.py:23-46."""
    a, b, c, d = x[:, :, :-1, :-1], x[:, :, :-1, 1:], x[:, :, 1:, :-1], x[:, :, 1:, 1:]
    if kind == "avg":
        return ((a + b + c + d) * np.float32(0.25)).astype(np.float32)
    return np.maximum(np.maximum(a, b), np.maximum(c, d))


def pyramid_ref_areas(cfg: Cfg):
    """pyramid_roi_align.py:18: (mean(anchor_scales[l]) * anchor_base) ** 2 in float32."""
    return (np.asarray(cfg.anchor_scales, dtype=np.float32).mean(axis=1) * cfg.anchor_base) ** 2


def assign_levels(bboxes, ref_areas):
    """pyramid_roi_align.py:33-37 -> int32 level per RoI."""
    b, ra = _f32(bboxes)[:, :4].copy(), _f32(ref_areas)
    out = np.empty((b.shape[0],), np.int32)
    lib().orc_assign_levels(_p(b, c_f), ctypes.c_long(b.shape[0]), _p(ra, c_f), int(ra.size), _p(out, c_i32))
    return out


def pyramid_roi_align(pyramids, bboxes, batch_inds, strides, ref_areas, ph, pw, mode="tf", all_levels=False):
    """PyramidRoIAlign (pyramid_roi_align.py:30-60: level-assigned, output in the caller's RoI
    order) or, all_levels=True, PyramidRoIAlign2 (:82-91: every RoI on every level, concatenated
    on channels; the 1x1 conv that follows is out of scope)."""
    fwd = roi_align_tf_fwd if mode == "tf" else roi_align_grid_fwd
    bboxes = _f32(bboxes)
    rois = np.concatenate([np.asarray(batch_inds, np.float32).reshape(-1, 1), bboxes], axis=1)
    if all_levels:
        return np.concatenate([fwd(p, rois, np.float32(1.0 / s), ph, pw) for p, s in zip(pyramids, strides)], axis=1)
    lv = assign_levels(bboxes, ref_areas)
    C = pyramids[0].shape[1]
    out = np.zeros((rois.shape[0], C, ph, pw), np.float32)
    for l, (p, s) in enumerate(zip(pyramids, strides)):
        sel = np.nonzero(lv == l)[0]
        if sel.size:
            out[sel] = fwd(p, rois[sel], np.float32(1.0 / s), ph, pw)
    return out


# ------------------------------------------------------------------------------------------------
# RoI -> GT matching, RoI sampling, RoI targets
# ------------------------------------------------------------------------------------------------
def matching_box(boxes, image_inds, gt_boxes_list, bg_overlap_threshold=0.5, fg_overlap_threshold=0.6):
    """libs/layers/box.py:12-69 on NumPy arrays: per image (rows `image_inds == i`, in image order) the
    float64 IoU argmax -> (labels int64, match int32, match_box f32), concatenated."""
    boxes = np.asarray(boxes, np.float32); image_inds = np.asarray(image_inds)
    labels_all, match_all, box_all = [], [], []
    for i, gt in enumerate(gt_boxes_list):
        gt = np.asarray(gt, np.float32).reshape(-1, 5)
        b = boxes[image_inds == i]
        match = np.zeros((b.shape[0],), np.int32) - 1                       # :33
        labels = np.zeros((b.shape[0],), np.int64)
        mbox = np.zeros((b.shape[0], 4), np.float32)
        if gt.size > 0 and b.size > 0:                                      # :36
            ov = bbox_overlaps(b[:, :4].astype(np.float64), gt[:, :4].astype(np.float64))
            arg = ov.argmax(axis=1)                                         # :42 (first maximum)
            mx = ov[np.arange(b.shape[0]), arg]
            match[:] = arg
            match[mx < bg_overlap_threshold] = -1                           # :45
            mbox[:, :4] = gt[arg, :4]
            labels[:] = gt[arg, 4]                                          # :48 (float -> int64 truncation)
            labels[mx < fg_overlap_threshold] = 0                           # :52
        labels_all.append(labels); match_all.append(match); box_all.append(mbox)
    return np.concatenate(labels_all), np.concatenate(match_all), np.concatenate(box_all)


def jitter_boxes(boxes, rng):
    """libs/layers/box.py:295-307 (in place; rand(num) drawn for x1, y1, then x2, y2)."""
    num = boxes.shape[0]
    ws = boxes[:, 2] - boxes[:, 0] + 1.0
    hs = boxes[:, 3] - boxes[:, 1] + 1.0
    dws = (rng.rand(num) - 0.5) * 0.2 * ws
    dhs = (rng.rand(num) - 0.5) * 0.2 * hs
    boxes[:, 0] = boxes[:, 0] + dws
    boxes[:, 1] = boxes[:, 1] + dhs
    dws = (rng.rand(num) - 0.5) * 0.2 * ws
    dhs = (rng.rand(num) - 0.5) * 0.2 * hs
    boxes[:, 2] = boxes[:, 2] + dws
    boxes[:, 3] = boxes[:, 3] + dhs
    return boxes


def sample_rois(boxes, image_inds, gt_boxes_list, fg_overlap_threshold=0.5, rois_per_image=512, fg_fraction=0.25,
                ignore_threshold=0.2, rng=np.random):
    """libs/layers/box.py:175-292 on NumPy arrays with an injectable RandomState (upstream uses the global
    np.random): filter small boxes (:310-315), label by float64 IoU argmax, ignore regions (class <= 0,
    inter/area(box) > ignore_threshold), append jittered valid GT, keep label >= 0, then the
    'faster rcnn sampling' branch (:264-276)."""
    boxes = np.asarray(boxes, np.float32); image_inds = np.asarray(image_inds)
    batch_size = len(gt_boxes_list)
    s_boxes, s_labels, s_inds = [], [], []
    for i, gt in enumerate(gt_boxes_list):
        gt = np.asarray(gt, np.float32).reshape(-1, 5)
        b = boxes[image_inds == i]
        ii = image_inds[image_inds == i]
        ws = b[:, 2] - b[:, 0] + 1.0; hs = b[:, 3] - b[:, 1] + 1.0           # filter_boxes, float32
        keep = np.where(np.logical_and(ws >= 8, hs >= 8))[0]
        b = b[keep]; ii = ii[keep]
        labels = np.zeros((b.shape[0],), np.int64)
        if gt.size > 0:
            ov = bbox_overlaps(b[:, :4].astype(np.float64), gt[:, :4].astype(np.float64))
            arg = ov.argmax(axis=1)
            mx = ov[np.arange(b.shape[0]), arg]
            labels[:] = gt[arg, 4]
            labels[mx < fg_overlap_threshold] = 0
            ign = gt[:, 4] <= 0                                               # :213
            if np.any(ign):
                flag = bbox_exclude_ignored_areas(b[:, :4].astype(np.float64), gt[ign][:, :4].astype(np.float64), ignore_threshold)
                labels[flag == 1] = -1
            valid = np.where(gt[:, 4] > 0)[0]                                 # :224-236
            gb = jitter_boxes(gt[valid][:, :4].astype(np.float32), rng)
            b = np.concatenate((b[:, :4], gb), axis=0)
            labels = np.concatenate((labels, gt[valid][:, 4].astype(np.int64)), axis=0)
            ii = np.concatenate((ii, np.zeros((gb.shape[0],), ii.dtype) + i), axis=0)
        s_boxes.append(b[labels >= 0][:, :4]); s_labels.append(labels[labels >= 0]); s_inds.append(ii[labels >= 0])
    s_boxes = np.concatenate(s_boxes, 0); s_labels = np.concatenate(s_labels, 0).astype(np.int64)
    s_inds = np.concatenate(s_inds, 0).astype(np.int64)
    bg_inds = np.where(s_labels == 0)[0]
    fg_inds = np.where(s_labels > 0)[0]
    num_fg = min(fg_inds.size, int(fg_fraction * rois_per_image * batch_size))
    if num_fg > 0:
        fg_inds = rng.choice(fg_inds, num_fg, replace=False)
    num_bg = min(rois_per_image * batch_size - num_fg, bg_inds.size)
    if bg_inds.size > 0:
        bg_inds = rng.choice(bg_inds, num_bg, replace=False)
    keep_inds = np.append(fg_inds, bg_inds).astype(np.int64)
    s_labels, s_boxes, s_inds = s_labels[keep_inds], s_boxes[keep_inds], s_inds[keep_inds]
    if s_labels.size == 0:                                                    # :282-286
        return boxes[:1, :4], np.array([-1], np.int64), image_inds[:1].astype(np.int64)
    return s_boxes, s_labels, s_inds


def roi_target(rois, roi_batch_inds, gt_boxes, gt_batch_inds, fg_overlap):
    """libs/layers/roi_target/src/cuda/roi_target_kernel.cu:16-90 restated (CUDA-only upstream; parity
    pinned on the GPU box against the kernel compiled unmodified, tests/test_gpu_ref_cuda.py).  float32
    arithmetic with the kernel's double-literal promotions; `<=` keeps the LAST maximum."""
    f = np.float32
    rois = np.asarray(rois, f).reshape(-1, 4); gt = np.asarray(gt_boxes, f).reshape(-1, 5)
    R = rois.shape[0]
    labels = np.zeros(R, np.int64); deltas = np.zeros((R, 4), f); bbw = np.zeros((R, 4), f)
    one = f(1.0)
    for r in range(R):
        a = rois[r]
        best, best_ind, match_cls = f(0.0), -1, -1
        for j in range(gt.shape[0]):
            if int(gt_batch_inds[j]) != int(roi_batch_inds[r]):
                continue
            b = gt[j]; cls = int(b[4])
            left, right = max(a[0], b[0]), min(a[2], b[2])
            top, bottom = max(a[1], b[1]), min(a[3], b[3])
            width = max(f(f(right - left) + one), f(0)); height = max(f(f(bottom - top) + one), f(0))
            inter = f(width * height)
            sa = f(f(f(a[2] - a[0]) + one) * f(f(a[3] - a[1]) + one))
            sb = f(f(f(b[2] - b[0]) + one) * f(f(b[3] - b[1]) + one))
            with np.errstate(all="ignore"):
                iou = f(inter / sa) if cls <= 0 else f(inter / f(f(sa + sb) - inter))
            if best <= iou:
                best, best_ind, match_cls = iou, j, cls
        labels[r] = (match_cls if match_cls > 0 else -1) if best >= f(fg_overlap) else 0
        if best >= min(f(0.5), f(fg_overlap)) and match_cls > 0 and best_ind >= 0:
            b = gt[best_ind]
            bw = f(np.float64(f(b[2] - b[0])) + 1.0); bh = f(np.float64(f(b[3] - b[1])) + 1.0)
            bx = f(np.float64(b[0]) + np.float64(bw) * 0.5); by = f(np.float64(b[1]) + np.float64(bh) * 0.5)
            aw = f(np.float64(f(a[2] - a[0])) + 1.0); ah = f(np.float64(f(a[3] - a[1])) + 1.0)
            ax = f(np.float64(a[0]) + np.float64(aw) * 0.5); ay = f(np.float64(a[1]) + np.float64(ah) * 0.5)
            with np.errstate(all="ignore"):
                deltas[r, 0] = f(np.float64(f(f(bx - ax) / aw)) / 0.1)
                deltas[r, 1] = f(np.float64(f(f(by - ay) / ah)) / 0.1)
                deltas[r, 2] = f(np.float64(np.log(f(bw / aw))) / 0.2)     # logf: float32 in, float32 out
                deltas[r, 3] = f(np.float64(np.log(f(bh / ah))) / 0.2)
            bbw[r] = 1.0
    return labels, deltas, bbw


# ------------------------------------------------------------------------------------------------
# NMS variants
# ------------------------------------------------------------------------------------------------
def _nms_variant(dets, thresh, variant):
    """libs/boxes/nms.pyx:70-123 (nms_new) / :125-195 (nms_multianchor) restated with the typed arithmetic the
    generated C has: float32 for every `cdef np.float32_t`, double where an untyped Python float or a double
    literal enters.  Ties: ascending index (upstream's argsort()[::-1] is unstable)."""
    f = np.float32
    dets = np.asarray(dets, f)
    x1, y1, x2, y2, sc = (dets[:, k] for k in range(5))
    areas = (x2 - x1 + f(1)) * (y2 - y1 + f(1))
    order = stable_desc_order(sc)
    n = dets.shape[0]
    suppressed = np.zeros(n, np.int64)
    clusters = np.arange(n)
    keep = []
    thresh = float(thresh)
    with np.errstate(all="ignore"):
        for _i in range(n):
            i = order[_i]
            if suppressed[i]:
                continue
            keep.append(int(i))
            icx = 0.5 * float(f(x1[i] + x2[i])); icy = 0.5 * float(f(y1[i] + y2[i]))
            for _j in range(_i + 1, n):
                j = order[_j]
                if suppressed[j]:
                    continue
                xx1 = max(x1[i], x1[j]); yy1 = max(y1[i], y1[j]); xx2 = min(x2[i], x2[j]); yy2 = min(y2[i], y2[j])
                w = f(max(0.0, float(f(f(xx2 - xx1) + f(1))))); h = f(max(0.0, float(f(f(yy2 - yy1) + f(1)))))
                inter = f(w * h)
                if variant == "new":
                    ovr = f(inter / f(f(areas[i] + areas[j]) - inter))
                    ovr1 = f(inter / areas[i]); ovr2 = f(inter / areas[j])
                    if float(ovr) >= thresh or float(ovr1) > 0.95 or float(ovr2) > 0.95:
                        suppressed[j] = 1
                else:
                    ovr = f((float(inter) + 0.1) / (float(f(f(areas[i] + areas[j]) - inter)) + 0.1))
                    if float(ovr) >= thresh:
                        suppressed[j] = 1; clusters[j] = i
                        continue
                    jcx = 0.5 * float(f(x1[j] + x2[j])); jcy = 0.5 * float(f(y1[j] + y2[j]))
                    dist = (jcx - icx) * (jcx - icx) + (jcy - icy) * (jcy - icy)
                    if np.sqrt(dist) <= 0.2 * float(min(w, h)):
                        suppressed[j] = 1; clusters[j] = i
    return keep, clusters


def nms_new(dets, thresh):
    return _nms_variant(dets, thresh, "new")[0]


def nms_multianchor(dets, thresh):
    return _nms_variant(dets, thresh, "multianchor")


# ------------------------------------------------------------------------------------------------
# to_Dets* epilogues and the stage-one composite
# ------------------------------------------------------------------------------------------------
def to_dets(boxes, probs, img_ids, batch_size, activation="softmax", multi_label=True, score_threshold=0.1):
    """libs/nets/model.py:357-469 restated (to_Dets / to_Dets_sigmoid for multi_label=False, to_Dets2 /
    to_Dets2_sigmoid for True): list of `batch_size` arrays [m,6] = (x1,y1,x2,y2,prob,class), float64."""
    boxes = np.asarray(boxes, np.float32).reshape(-1, 4); img_ids = np.asarray(img_ids).reshape(-1)
    probs = np.asarray(probs, np.float32).reshape(boxes.shape[0], -1)
    first = 1 if activation == "softmax" else 0
    out = []
    for i in range(batch_size):
        inds = np.where(img_ids == i)[0]
        p, b = probs[inds], boxes[inds]
        fg = p[:, first:]
        if fg.shape[1] == 1:                                               # :365-367 / :387-389
            cls_ids = np.ones((p.shape[0],), np.int32); cls_probs = fg[:, 0]
        else:
            d0, d1 = np.where(fg > score_threshold) if multi_label else (np.zeros(0, np.int64), np.zeros(0, np.int64))
            if d0.size > 0:                                                # :421-427
                cls_ids = d1 + 1; cls_probs = fg[d0, d1]; b = b[d0, :]
            else:                                                          # :369-370 / :428-431
                am = fg.argmax(axis=1) if fg.shape[0] else np.zeros(0, np.int64)
                cls_ids = am + 1; cls_probs = fg[np.arange(fg.shape[0]), am]
        out.append(np.concatenate((b.reshape(-1, 4), cls_probs[:, None], cls_ids[:, None]), axis=1).astype(np.float64))
    return out


def final_results_stage1(rpn_box, rpn_prob, anchors_, batch_size, score_threshold=0.1, max_dets=100, overlap_threshold=0.5,
                         activation="softmax", box_encoding="fastrcnn"):
    """libs/nets/model.py:474-497 composed from the pieces above: top 9*max_dets (:483, :326) -> objectness >=
    threshold with the single-best fallback (:329-333) -> per-image NMS ('>' comparator of the CUDA path) ->
    to_Dets2 / to_Dets2_sigmoid."""
    boxes, probs, img_ids, _, _ = decode_and_choose_top_n(rpn_box, rpn_prob, anchors_, max_dets * 9, activation, box_encoding)
    obj = objectness(probs, activation)
    inds = np.nonzero(obj >= np.float32(score_threshold))[0]
    if inds.size == 0:
        inds = stable_desc_order(obj)[:1]
    boxes, probs, img_ids = boxes[inds], probs[inds], img_ids[inds]
    keep = apply_nms_in_batch(boxes, probs, img_ids, activation, overlap_threshold, NMS_GT_F32)
    return to_dets(boxes[keep], probs[keep], img_ids[keep], batch_size, activation, True, score_threshold)


# ------------------------------------------------------------------------------------------------
# dets_cluster: score-weighted box voting over nms_multianchor's clusters
# ------------------------------------------------------------------------------------------------
def dets_cluster(dets, keeps, clusters):
    """libs/postprocessings/single_shot.py:401-415 restated: per kept index, the rows of its cluster
    (`clusters == ind`, ascending) are averaged with their scores as weights (float32, NumPy's own summation
    order), the score becomes the cluster's maximum; a cluster of one keeps its row; no keeps -> `dets`."""
    dets = np.asarray(dets, np.float32); clusters = np.asarray(clusters).reshape(-1)
    rows = []
    for ind in np.asarray(keeps).reshape(-1):
        members = dets[np.where(clusters == ind)[0], :]
        if members.shape[0] > 1:
            score = members[:, 4]
            voted = np.sum(members[:, :4] * score[:, None] / np.sum(score), axis=0)
            members = np.concatenate([voted, [np.max(score)]]).astype(np.float32)
        rows.append(members)
    return np.vstack(rows) if rows else dets
