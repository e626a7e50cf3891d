#!/usr/bin/env python
"""Generate tests/golden/*.npz by running the UNMODIFIED reference (compiled into oracle/_ref by
oracle/build_ref.py) on seeded inputs.  TEST INFRASTRUCTURE ONLY.

Run in the build container (where /root/reference exists and oracle/_ref was built):
    python oracle/build_ref.py && python oracle/make_golden.py
The vectors are committed; the GPU box only reads them.
Each file stores inputs AND the reference's outputs, so the tests need no reference at run time.
"""
from __future__ import annotations

import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import ref_loader  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")

CPR_SCALES = [[2, 2.520, 3.175], [4, 5.040, 6.349], [8, 10.079, 12.700],
              [16, 20.159, 25.398], [32, 40.317, 50.797]]
CPR_STRIDES = [8, 16, 32, 64, 128]
CPR_RATIOS = [1, 2.5]

# libs/layers/anchor_target/modules/anchor_target_test.py:15-31 — the reference's own fixture
FIX_ANCHORS = np.asarray([[120, 120, 140, 150], [95, 100, 120, 130], [120, 120, 125, 125],
                          [40, 50, 50, 60], [95, 100, 120, 130], [0, 0, 9, 9], [1, 1, 8, 8],
                          [10, 10, 20, 20], [16, 10, 26, 20], [70, 60, 80, 90]], np.float32)
FIX_GT = np.asarray([[121, 122, 144, 155, 2], [1, 1, 3, 3, 1], [0, 0, 10, 10, -1]], np.float32)


def synth_gt(rng, G, ih, iw, ignore_frac=0.15, ncls=1):
    """pedestrian-shaped boxes, some of them ignore regions (cls -1)."""
    h = np.exp(rng.uniform(np.log(12), np.log(264), G))
    w = h / rng.uniform(2.0, 2.8, G)
    cx, cy = rng.uniform(0, iw, G), rng.uniform(0, ih, G)
    b = np.stack([cx - w / 2, cy - h / 2, cx + w / 2, cy + h / 2], 1)
    b[:, 0::2] = np.clip(b[:, 0::2], 0, iw - 1)
    b[:, 1::2] = np.clip(b[:, 1::2], 0, ih - 1)
    cls = rng.randint(1, ncls + 1, G).astype(np.float64)
    cls[rng.rand(G) < ignore_frac] = -1
    return np.concatenate([b, cls[:, None]], 1).astype(np.float32)


def set_cfg(cfg, bg, fg, clobber=True, enc="fastrcnn"):
    cfg.rpn_bg_threshold, cfg.rpn_fg_threshold = bg, fg
    cfg.rpn_clobber_positives = clobber
    cfg.rpn_sample_strategy = "simple"
    cfg.rpn_box_encoding = enc
    cfg.bbweights = [1., 1., 1., 1.]
    cfg.ignored_area_intersection_fraction = 0.3


def synth_rois(seed, n=None, B=None, sorted_ids=True):
    """Proposals + per-image GT lists for matching_box / sample_rois: log-uniform sizes (some below
    filter_boxes' 8 px), GT classes in {-1, 0, 1, 2, 3} (<= 0 are ignore regions), a few proposals planted
    on top of GT boxes so that foregrounds exist, one image possibly without GT."""
    rng = np.random.RandomState(seed)
    n = rng.randint(1, 400) if n is None else n
    B = rng.randint(1, 5) if B is None else B
    wh = np.exp(rng.uniform(np.log(3), np.log(200), (n, 2)))
    c = rng.uniform(0, 300, (n, 2))
    b = np.concatenate([c - wh / 2, c + wh / 2], 1).astype(np.float32)
    ids = rng.randint(0, B, n)
    if sorted_ids:
        ids = np.sort(ids)
    gts = []
    for i in range(B):
        G = rng.randint(0, 6)
        gwh = np.exp(rng.uniform(np.log(10), np.log(200), (G, 2)))
        gc = rng.uniform(0, 300, (G, 2))
        cls = rng.choice([-1, 0, 1, 2, 3], G).astype(np.float32)
        gts.append(np.concatenate([gc - gwh / 2, gc + gwh / 2, cls[:, None]], 1).astype(np.float32))
    for i in range(B):
        for g in gts[i]:
            for _ in range(3):
                k = rng.randint(0, n)
                b[k] = g[:4] + rng.uniform(-3, 3, 4).astype(np.float32)
                ids[k] = i
    if sorted_ids:
        o = np.argsort(ids, kind="stable"); b, ids = b[o], ids[o]
    return b, ids.astype(np.int64), gts


def roi_sampling_golden(r):
    """matching_box / sample_rois of the reference's libs/layers/box.py (CPU torch tensors in and out;
    np.random seeded per case, which fixes the jitter and both np.random.choice draws)."""
    import torch
    out = {}
    for case, (seed, sorted_ids) in enumerate([(0, True), (1, True), (2, False), (3, True), (4, False), (5, True)]):
        b, ids, gts = synth_rois(seed, sorted_ids=sorted_ids)
        out["c%d_boxes" % case] = b; out["c%d_ids" % case] = ids
        out["c%d_ngt" % case] = np.asarray([g.shape[0] for g in gts], np.int32)
        out["c%d_gt" % case] = np.concatenate(gts, 0) if sum(g.shape[0] for g in gts) else np.zeros((0, 5), np.float32)
        lab, match, mbox = r.layers_box.matching_box(torch.from_numpy(b), torch.from_numpy(ids), gts, 0.4, 0.6)
        out["c%d_match_labels" % case] = lab.numpy(); out["c%d_match_inds" % case] = match.numpy()
        out["c%d_match_boxes" % case] = mbox.numpy()
        np.random.seed(100 + case)
        sb, sl, si = r.layers_box.sample_rois(torch.from_numpy(b), torch.from_numpy(ids), gts, 0.5, 16, 0.25, 0.2)
        out["c%d_sample_boxes" % case] = sb.numpy(); out["c%d_sample_labels" % case] = sl.numpy()
        out["c%d_sample_ids" % case] = si.numpy()
    np.savez_compressed(os.path.join(GOLD, "roi_sampling.npz"), **out)


def variant_dets(seed):
    rng = np.random.RandomState(seed)
    n = [200, 500, 40, 700][seed % 4]
    wh = np.exp(rng.uniform(np.log(4), np.log(80), (n, 2)))
    c = rng.uniform(0, 250, (n, 2))
    b = np.concatenate([c - wh / 2, c + wh / 2], 1)
    if seed % 2:
        b = np.round(b)                     # integer boxes: quotients exactly on thresholds, coincident centres
    s = rng.permutation(n) / float(n)
    return np.concatenate([b, s[:, None]], 1).astype(np.float32)


def nms_variants_golden(r):
    """nms_new / nms_multianchor of the reference's libs/boxes/nms.pyx (compiled as cython_nms)."""
    out = {}
    for seed in range(4):
        d = variant_dets(seed)
        out["d%d" % seed] = d
        for thr in (0.3, 0.5):
            out["d%d_new_%g" % (seed, thr)] = np.asarray(r.cython_nms.nms_new(d, thr), np.int64)
            keep, clusters = r.cython_nms.nms_multianchor(d, thr)
            out["d%d_multi_keep_%g" % (seed, thr)] = np.asarray(keep, np.int64)
            out["d%d_multi_clusters_%g" % (seed, thr)] = np.asarray(clusters, np.int64)
    np.savez_compressed(os.path.join(GOLD, "nms_variants.npz"), **out)


def to_dets_cases():
    """(name, activation, multi_label, C, threshold, batch, sorted ids) for the to_Dets* conversions."""
    return [("softmax2", "softmax", True, 2, 0.1, 3, True), ("softmax5", "softmax", True, 5, 0.3, 4, True),
            ("softmax5_unsorted", "softmax", True, 5, 0.45, 4, False), ("softmax5_single", "softmax", False, 5, 0.0, 3, True),
            ("sigmoid4", "sigmoid", True, 4, 0.5, 4, True), ("sigmoid4_single", "sigmoid", False, 4, 0.0, 3, False)]


def to_dets_inputs(case):
    name, act, multi, C, thr, B, sorted_ids = case
    import zlib
    rng = np.random.RandomState(zlib.crc32(name.encode()) % 1000)
    n = 120
    xy = rng.uniform(0, 300, (n, 2)); wh = rng.uniform(5, 80, (n, 2))
    boxes = np.concatenate([xy, xy + wh], 1).astype(np.float32)
    probs = rng.dirichlet(np.ones(C) * 0.7, n).astype(np.float32) if act == "softmax" else rng.rand(n, C).astype(np.float32)
    ids = rng.randint(0, B - 1, n)                       # the last image has no rows
    if sorted_ids:
        ids = np.sort(ids)
    if C > 2:
        probs[ids == 0] *= np.float32(0.2)               # image 0: nothing above the threshold -> arg-max rule
    return boxes, probs, ids.astype(np.int64)


def to_dets_golden(ref_root="/root/reference"):
    """to_Dets / to_Dets_sigmoid / to_Dets2 / to_Dets2_sigmoid: the four static methods are taken out of the
    reference's libs/nets/model.py:357-469 AT RUN TIME (ast) and executed on NumPy inputs — model.py itself imports
    the whole network stack and cannot be imported here.  Nothing of the reference's text is stored."""
    import ast, textwrap, types
    src = open(os.path.join(ref_root, "libs/nets/model.py")).read()
    tree = ast.parse(src)
    ns = {"np": np, "everything2numpy": lambda xs: [np.asarray(x) for x in xs], "cfg": types.SimpleNamespace(batch_size=0)}
    for node in ast.walk(tree):
        if isinstance(node, ast.FunctionDef) and node.name in ("to_Dets", "to_Dets_sigmoid", "to_Dets2", "to_Dets2_sigmoid"):
            node.decorator_list = []
            exec(compile(ast.Module(body=[node], type_ignores=[]), "<reference model.py>", "exec"), ns)
    out = {}
    for case in to_dets_cases():
        name, act, multi, C, thr, B, _ = case
        boxes, probs, ids = to_dets_inputs(case)
        ns["cfg"].batch_size = B
        fn = ns[{("softmax", False): "to_Dets", ("sigmoid", False): "to_Dets_sigmoid",
                 ("softmax", True): "to_Dets2", ("sigmoid", True): "to_Dets2_sigmoid"}[(act, multi)]]
        dets = fn(boxes, probs, ids, thr) if multi else fn(boxes, probs, ids)
        out[name + "_boxes"], out[name + "_probs"], out[name + "_ids"] = boxes, probs, ids
        out[name + "_dets"] = np.concatenate([np.asarray(d, np.float64).reshape(-1, 6) for d in dets], 0)
        out[name + "_offsets"] = np.cumsum([0] + [np.asarray(d).reshape(-1, 6).shape[0] for d in dets]).astype(np.int32)
    np.savez_compressed(os.path.join(GOLD, "to_dets.npz"), **out)


def real_gt_sets(ref_root="/root/reference", picks=6):
    """GT sets of real CityPersons validation images: data/citypersons/annotations/anno_val.mat rows
    (class, x, y, w, h, instance, x_vis, y_vis, w_vis, h_vis) -> [x1, y1, x2, y2, class] with the conversion of
    libs/datasets/citypersons2.py:85-89 (x2 = x + w) and the class / ignore rules of :801-812 (0, 4, 5 -> -1; 2, 3 -> 1;
    visible fraction < 0.3 or height < 32 -> -1).  Picks image 0, the images with the most GT / most ignore boxes /
    most pedestrians, and two more."""
    import scipy.io as sio
    ann = sio.loadmat(os.path.join(ref_root, "data/citypersons/annotations/anno_val.mat"))["anno_val_aligned"].reshape(-1)
    sets = []
    for a in ann:
        bbs = a.reshape(-1)[0][2].astype(np.float32)
        if bbs.size == 0:
            sets.append(np.zeros((0, 5), np.float32)); continue
        cls, box, vis = bbs[:, 0].astype(np.int32), bbs[:, 1:5].copy(), bbs[:, 6:10].copy()
        box[:, 2:4] += box[:, 0:2]; vis[:, 2:4] += vis[:, 0:2]
        cls[cls == 0] = -1; cls[cls == 2] = 1; cls[cls == 3] = 1; cls[cls == 4] = -1; cls[cls == 5] = -1
        frac = ((vis[:, 3] - vis[:, 1] + 1) * (vis[:, 2] - vis[:, 0] + 1)) / ((box[:, 3] - box[:, 1] + 1) * (box[:, 2] - box[:, 0] + 1))
        cls[frac < 0.3] = -1; cls[(box[:, 3] - box[:, 1] + 1) < 32] = -1
        sets.append(np.concatenate([box, cls[:, None].astype(np.float32)], 1).astype(np.float32))
    if picks is None:                                        # every validation image, in file order
        return list(enumerate(sets))
    n = np.asarray([g.shape[0] for g in sets]); ign = np.asarray([(g[:, 4] < 0).sum() for g in sets]); ped = n - ign
    chosen = []
    for i in [0, int(n.argmax()), int(ign.argmax()), int(ped.argmax()), 123, 377]:
        if i not in chosen:
            chosen.append(i)
    return [(i, sets[i]) for i in chosen[:picks]]


def real_gt_golden(r):
    """anchor.encode of the compiled reference on real CityPersons GT x the CP-R pyramid (rpn_bg 0.5 / rpn_fg 0.6, the
    yaml's thresholds): labels stored sparsely, label-weight values, every 1009th target row plus the rows that carry
    inside weights."""
    A, cfg = r.boxes_anchor, r.cfg
    pyr = A.anchor_pyramid(CPR_STRIDES, 1024, 2048, CPR_SCALES, CPR_RATIOS, 16).astype(np.float32)
    set_cfg(cfg, 0.5, 0.6)
    out = {"image_ids": np.asarray([i for i, _ in real_gt_sets()], np.int64)}
    for i, gt in real_gt_sets():
        lab, lw, bt, bw = r.layers_anchor.encode(gt.copy(), pyr.copy())
        k = "img%d_" % i
        nz = np.nonzero(lab != 0)[0]
        out[k + "gt"] = gt
        out[k + "labels_nz_idx"], out[k + "labels_nz_val"] = nz, lab[nz]
        out[k + "counts"] = np.asarray([(lab >= 1).sum(), (lab == 0).sum(), (lab < 0).sum()])
        out[k + "bg_weight"] = lw[lab == 0][:1]
        wrows = np.nonzero(bw.any(1))[0]
        rows = np.union1d(np.arange(0, pyr.shape[0], 1009), wrows)
        out[k + "inside_rows"] = wrows
        out[k + "rows_idx"], out[k + "bbox_targets_rows"] = rows, bt[rows]
    np.savez_compressed(os.path.join(GOLD, "anchor_target_real.npz"), **out)


def dets_cluster_cases():
    """Synthetic (dets, keep, clusters) with clusters of 1, 2, 7, 8, 9, 130 and 400 members (NumPy's pairwise
    summation changes shape at 8 and at 128 elements); members interleaved, keepers not in index order."""
    rng = np.random.RandomState(11)
    sizes = [1, 2, 7, 8, 9, 130, 400, 1, 33]
    n = sum(sizes)
    owner = rng.permutation(np.repeat(np.arange(len(sizes)), sizes))
    keep = np.asarray([np.where(owner == c)[0][0] for c in range(len(sizes))], np.int64)
    clusters = keep[owner]
    xy = rng.uniform(0, 1800, (n, 2)); wh = rng.uniform(5, 200, (n, 2))
    dets = np.concatenate([xy, xy + wh, rng.uniform(0.01, 1, (n, 1))], 1).astype(np.float32)
    return [("synth", dets, keep[rng.permutation(len(sizes))], clusters)]


def dets_cluster_golden(ref_root="/root/reference"):
    """dets_cluster of libs/postprocessings/single_shot.py:401-415, lifted out of the module with `ast` at run time
    (single_shot.py imports cv2-based code) and run on nms_multianchor's golden (keep, clusters) plus a synthetic
    case with large clusters.  Nothing of the reference's text is stored."""
    import ast
    tree = ast.parse(open(os.path.join(ref_root, "libs/postprocessings/single_shot.py")).read())
    ns = {"np": np}
    for node in ast.walk(tree):
        if isinstance(node, ast.FunctionDef) and node.name == "dets_cluster":
            exec(compile(ast.Module(body=[node], type_ignores=[]), "<reference single_shot.py>", "exec"), ns)
    g = np.load(os.path.join(GOLD, "nms_variants.npz"))
    out = {}
    for seed in range(4):
        for thr in (0.3, 0.5):
            d = g["d%d" % seed]
            out["d%d_voted_%g" % (seed, thr)] = np.asarray(
                ns["dets_cluster"](d.copy(), g["d%d_multi_keep_%g" % (seed, thr)], g["d%d_multi_clusters_%g" % (seed, thr)]), np.float32)
    for name, d, keep, clusters in dets_cluster_cases():
        out[name + "_dets"], out[name + "_keep"], out[name + "_clusters"] = d, keep, clusters
        out[name + "_voted"] = np.asarray(ns["dets_cluster"](d.copy(), keep, clusters), np.float32)
    out["numpy_version"] = np.asarray(np.__version__)
    np.savez_compressed(os.path.join(GOLD, "dets_cluster.npz"), **out)


def main():
    r = ref_loader.load()
    os.makedirs(GOLD, exist_ok=True)
    if "--only-to-dets" in sys.argv:
        to_dets_golden()
        return
    if "--only-dets-cluster" in sys.argv:
        dets_cluster_golden()
        return
    if "--only-real-gt" in sys.argv:
        real_gt_golden(r)
        return
    if "--only-roi-sampling" in sys.argv:
        roi_sampling_golden(r)
        return
    if "--only-nms-variants" in sys.argv:
        nms_variants_golden(r)
        return
    A = r.boxes_anchor

    # ---- anchors -------------------------------------------------------------------------
    out = {}
    for lvl in range(5):
        out["base_cpr_l%d" % lvl] = A.anchors(CPR_SCALES[lvl], CPR_RATIOS, 16)
    out["base_default"] = A.anchors([2., 4., 8., 16., 32.], [0.5, 1, 2.0], 16)
    out["base_odd15"] = A.anchors([1.5, 3], [0.5, 1, 2.0, 3.0], 15)
    pyr = A.anchor_pyramid(CPR_STRIDES, 1024, 2048, CPR_SCALES, CPR_RATIOS, 16, [[0.0, 0.0]])
    out["cpr_pyramid_shape"] = np.asarray(pyr.shape)
    out["cpr_pyramid_rows_idx"] = np.arange(0, pyr.shape[0], 997)
    out["cpr_pyramid_rows"] = pyr[::997]
    out["cpr_pyramid_colsum"] = pyr.sum(0)
    out["small_pyramid"] = A.anchor_pyramid([4, 8, 16], 64, 96, [[2, 3], [4, 6], [8, 12]], [0.5, 1, 2.0], 16,
                                           [[0.0, 0.0], [0.5, 0.25]])
    out["plane_shifted"] = A.anchor_one_plane(5, 7, 8, [2., 4.], [0.5, 2.0], 16, [[0, 0], [0.5, 0]])
    np.savez_compressed(os.path.join(GOLD, "anchors.npz"), **out)

    # ---- IoU -----------------------------------------------------------------------------
    rng = np.random.RandomState(10)
    b = rng.uniform(0, 200, (300, 4)); b[:, 2:] = b[:, :2] + rng.uniform(0, 80, (300, 2))
    q = rng.uniform(0, 200, (17, 4)); q[:, 2:] = q[:, :2] + rng.uniform(0, 80, (17, 2))
    b32, q32 = b.astype(np.float32).astype(np.float64), q.astype(np.float32).astype(np.float64)
    bi = np.floor(b / 8); qi = np.floor(q / 8)         # integer boxes: many exact ties
    out = dict(fix_anchors=FIX_ANCHORS, fix_gt=FIX_GT,
               fix_ov=r.cython_bbox.bbox_overlaps(FIX_ANCHORS.astype(np.float64), FIX_GT[:, :4].astype(np.float64)),
               boxes=b32, query=q32, ov=r.cython_bbox.bbox_overlaps(b32, q32),
               inter=r.cython_bbox.bbox_intersections(b32, q32),
               excl=r.cython_bbox.bbox_exclude_ignored_areas(b32, q32, 0.2),
               boxes_int=bi, query_int=qi, ov_int=r.cython_bbox.bbox_overlaps(bi, qi))
    np.savez_compressed(os.path.join(GOLD, "iou.npz"), **out)

    # ---- encode / decode -----------------------------------------------------------------
    rng = np.random.RandomState(11)
    anc = pyr[rng.choice(pyr.shape[0], 500, replace=False)].astype(np.float32)
    gt = synth_gt(rng, 500, 1024, 2048, 0.0)[:, :4]
    deltas = rng.normal(0, 0.5, (500, 4)).astype(np.float32)
    T = r.bbox_transform
    dec = T.bbox_transform_inv(anc, deltas)
    out = dict(anchors=anc, gt=gt, deltas=deltas,
               enc=T.bbox_transform(anc, gt), dec=dec,
               dec_clip=T.clip_boxes(dec.copy(), (1024, 2048)),
               enc_linear=T.bbox_transform_linear(anc, gt), dec_linear=T.bbox_transform_inv_linear(anc, deltas))
    np.savez_compressed(os.path.join(GOLD, "transform.npz"), **out)

    # ---- anchor target -------------------------------------------------------------------
    cfg = r.cfg
    out = {}
    cases = []
    set_cfg(cfg, 0.3, 0.6)
    cases.append(("fixture", FIX_ANCHORS, FIX_GT, 0.3, 0.6, True))
    small = A.anchor_pyramid([8, 16, 32], 128, 256, CPR_SCALES[:3], CPR_RATIOS, 16).astype(np.float32)
    rng = np.random.RandomState(12)
    g1 = synth_gt(rng, 9, 128, 256, 0.3); g1[:, :4] *= 1.0
    g1[:, 2] = np.minimum(g1[:, 0] + np.maximum(g1[:, 2] - g1[:, 0], 10), 255)
    cases.append(("small_a", small, g1, 0.5, 0.6, True))
    cases.append(("small_noclobber", small, g1, 0.4, 0.6, False))
    g2 = np.asarray([[10, 10, 60, 90, 1], [400, 300, 420, 330, 1], [30, 20, 80, 100, -1]], np.float32)  # gt#1 off-canvas: zero column
    cases.append(("small_zero_col", small, g2, 0.5, 0.6, True))
    cases.append(("small_zero_col_noclobber", small, g2, 0.5, 0.6, False))
    cases.append(("small_nogt", small, np.zeros((0, 5), np.float32), 0.5, 0.6, True))
    g3 = np.asarray([[0, 0, 32, 32, 3], [0, 0, 32, 32, 2], [8, 8, 40, 40, 5]], np.float32)             # duplicate gt: argmax ties
    cases.append(("small_dup_gt", small, g3, 0.3, 0.7, True))
    cpr = pyr.astype(np.float32)
    g4 = synth_gt(np.random.RandomState(0), 12, 1024, 2048)
    for name, anc_, gt_, bg, fg, clob in cases:
        set_cfg(cfg, bg, fg, clob)
        lab, lw, bt, bw = r.layers_anchor.encode(gt_.copy(), anc_.copy())
        out[name + "_anchors"], out[name + "_gt"] = anc_, gt_
        out[name + "_cfg"] = np.asarray([bg, fg, float(clob)])
        out[name + "_labels"], out[name + "_label_weights"] = lab, lw
        out[name + "_bbox_targets"], out[name + "_bbox_inside_weights"] = bt, bw
    # full CP-R image: store labels compactly (+ sparse target rows)
    set_cfg(cfg, 0.5, 0.6)
    lab, lw, bt, bw = r.layers_anchor.encode(g4.copy(), cpr.copy())
    nz = np.nonzero(lab != 0)[0]
    out["cpr_gt"] = g4
    out["cpr_labels_nz_idx"], out["cpr_labels_nz_val"] = nz, lab[nz]
    out["cpr_counts"] = np.asarray([(lab >= 1).sum(), (lab == 0).sum(), (lab < 0).sum()])
    out["cpr_bg_weight"] = lw[lab == 0][:1]
    rows = np.arange(0, cpr.shape[0], 1009)
    out["cpr_rows_idx"], out["cpr_bbox_targets_rows"] = rows, bt[rows]
    out["cpr_bbox_inside_weights_sum"] = bw.sum(0)
    np.savez_compressed(os.path.join(GOLD, "anchor_target.npz"), **out)

    # ---- NMS -----------------------------------------------------------------------------
    out = {}
    rng = np.random.RandomState(2)
    for name, n in (("rand1k", 1000), ("rand3k", 3000)):
        h = np.exp(rng.uniform(np.log(12), np.log(264), n)); w = h / 2.44
        cx, cy = rng.uniform(0, 2048, n), rng.uniform(0, 1024, n)
        d = np.stack([cx - w / 2, cy - h / 2, cx + w / 2, cy + h / 2, rng.permutation(np.linspace(0, 1, n))], 1)
        d[:, 0:4:2] = np.clip(d[:, 0:4:2], 0, 2047); d[:, 1:4:2] = np.clip(d[:, 1:4:2], 0, 1023)
        out[name] = d.astype(np.float32)
    # clustered: few survive
    n = 2000
    gtc = synth_gt(rng, 40, 1024, 2048, 0.0)[:, :4]
    base = gtc[rng.randint(0, 40, n)]
    wh = np.stack([base[:, 2] - base[:, 0], base[:, 3] - base[:, 1]] * 2, 1)
    d = np.concatenate([base + rng.normal(0, 0.1, (n, 4)) * wh, rng.permutation(np.linspace(0, 1, n))[:, None]], 1)
    out["clustered2k"] = d.astype(np.float32)
    # integer grid boxes -> many IoUs exactly at 0.5 / 1/3: separates '>=' from '>'
    n = 600
    xy = rng.randint(0, 40, (n, 2)).astype(np.float64)
    wh = rng.choice([1, 2, 3, 5], (n, 2)).astype(np.float64)
    d = np.concatenate([xy, xy + wh, rng.permutation(np.linspace(0, 1, n))[:, None]], 1)
    out["intgrid600"] = d.astype(np.float32)
    for name in ("rand1k", "rand3k", "clustered2k", "intgrid600"):
        d = out[name]
        for thr in (0.5, 0.3, 0.7):
            out["%s_keep_ge_%g" % (name, thr)] = np.asarray(r.cpu_nms.cpu_nms(d, thr), np.int64)
            out["%s_keep_gt_%g" % (name, thr)] = np.asarray(r.py_cpu_nms.py_cpu_nms(d, thr), np.int64)
            out["%s_keep_cynms_%g" % (name, thr)] = np.asarray(r.cython_nms.nms(d, thr), np.int64)
    np.savez_compressed(os.path.join(GOLD, "nms.npz"), **out)

    # ---- RoIAlign (crop_and_resize.c) ------------------------------------------------------
    rng = np.random.RandomState(3)
    img = rng.normal(0, 1, (2, 5, 12, 17)).astype(np.float32)
    nb = 23
    y1 = rng.uniform(-0.1, 0.8, nb); x1 = rng.uniform(-0.1, 0.8, nb)
    boxes = np.stack([y1, x1, y1 + rng.uniform(0.05, 0.5, nb), x1 + rng.uniform(0.05, 0.5, nb)], 1).astype(np.float32)
    ind = rng.randint(0, 2, nb).astype(np.int32)
    g = rng.normal(0, 1, (nb, 5, 7, 7)).astype(np.float32)
    out = dict(image=img, boxes=boxes, box_ind=ind, grad=g,
               crops7=r.crop_and_resize_forward(img, boxes, ind, 0.0, 7, 7),
               crops3x5=r.crop_and_resize_forward(img, boxes, ind, 0.0, 3, 5),
               crops1=r.crop_and_resize_forward(img, boxes, ind, 0.0, 1, 1),
               grad_image7=r.crop_and_resize_backward(g, boxes, ind, img.shape))
    ramp = np.arange(16, dtype=np.float32).reshape(1, 1, 4, 4)
    # SURVEY KAT-5: RoI (0,0,3,3) at scale 1, 2x2 bins -> normalised box through the wrapper arithmetic
    sw = np.float32(3.0) / np.float32(2.0)
    n0 = (np.float32(0.0) + sw / np.float32(2) - np.float32(0.5)) / np.float32(3.0)
    nw = sw * np.float32(1.0) / np.float32(3.0)
    kb = np.asarray([[n0, n0, n0 + nw, n0 + nw]], np.float32)
    out["kat5_box"] = kb
    out["kat5_fwd"] = r.crop_and_resize_forward(ramp, kb, np.zeros(1, np.int32), 0.0, 2, 2)
    out["kat5_bwd"] = r.crop_and_resize_backward(np.ones((1, 1, 2, 2), np.float32), kb, np.zeros(1, np.int32), ramp.shape)
    np.savez_compressed(os.path.join(GOLD, "roialign.npz"), **out)

    roi_sampling_golden(r)
    nms_variants_golden(r)
    to_dets_golden()
    dets_cluster_golden()
    real_gt_golden(r)

    sz = sum(os.path.getsize(os.path.join(GOLD, f)) for f in os.listdir(GOLD))
    print("golden vectors written to", GOLD, "(%d KiB)" % (sz // 1024))


if __name__ == "__main__":
    main()
