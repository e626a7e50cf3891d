"""The drop-in boundary: the reference's module paths, names and call shapes, served by
detbox.libs after install(), checked against the golden vectors the compiled reference produced
(tests/golden/*.npz) and against the CPU oracle.  These read like the reference's own scripts:
`import libs.boxes.cython_bbox as cython_bbox`, `from libs.layers.box import apply_nms`, ..."""
import numpy as np
import pytest
import torch

import detbox
from oracle import oracle as O
from util import assert_boxes_close

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module", autouse=True)
def libs():
    yield detbox.libs.install()
    detbox.libs.uninstall()          # tests/test_oracle_vs_ref.py imports the compiled reference as `libs`


def cu(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


# ------------------------------------------------------------------------------- libs.boxes ----
def test_cython_bbox_matches_golden(golden):
    import libs.boxes.cython_bbox as cython_bbox
    g = golden("iou")
    assert np.array_equal(cython_bbox.bbox_overlaps(g["boxes"], g["query"]), g["ov"])
    assert np.array_equal(cython_bbox.bbox_intersections(g["boxes"], g["query"]), g["inter"])
    assert np.array_equal(cython_bbox.bbox_overlaps(g["boxes_int"], g["query_int"]), g["ov_int"])
    assert np.array_equal(cython_bbox.bbox_exclude_ignored_areas(g["boxes"], g["query"], 0.2), g["excl"])
    # the reference's own fixture (anchor_target_test.py:15-31)
    ov = cython_bbox.bbox_overlaps(g["fix_anchors"].astype(np.float64), np.ascontiguousarray(g["fix_gt"][:, :4], np.float64))
    assert np.array_equal(ov, g["fix_ov"])
    with pytest.raises(ValueError):
        cython_bbox.bbox_overlaps(g["boxes"].astype(np.float32), g["query"])


def test_anchor_module_matches_golden(golden):
    from libs.boxes.anchor import anchor_one_plane, anchor_pyramid, anchors, anchors_plane
    g = golden("anchors")
    scales = [[2, 2.520, 3.175], [4, 5.040, 6.349], [8, 10.079, 12.700], [16, 20.159, 25.398], [32, 40.317, 50.797]]
    for l in range(5):
        assert np.array_equal(anchors(scales[l], [1, 2.5], 16), g["base_cpr_l%d" % l])
    pyr = anchor_pyramid([8, 16, 32, 64, 128], 1024, 2048, scales, [1, 2.5], 16, [[0.0, 0.0]])
    assert pyr.dtype == np.float64 and tuple(pyr.shape) == tuple(g["cpr_pyramid_shape"])
    assert np.array_equal(pyr[g["cpr_pyramid_rows_idx"]], g["cpr_pyramid_rows"])
    assert np.array_equal(anchor_pyramid([4, 8, 16], 64, 96, [[2, 3], [4, 6], [8, 12]], [0.5, 1, 2.0], 16,
                                         [[0.0, 0.0], [0.5, 0.25]]), g["small_pyramid"])
    assert np.array_equal(anchor_one_plane(5, 7, 8, [2., 4.], [0.5, 2.0], 16, [[0, 0], [0.5, 0]]), g["plane_shifted"])
    pl = anchors_plane(5, 7, 8, [2., 4.], [0.5, 2.0], 16, [[0, 0]])
    assert pl.shape == (5, 7, 4, 4)
    assert np.array_equal(pl.reshape(-1, 4), O.anchor_one_plane(5, 7, 8, [2., 4.], [0.5, 2.0], 16, [[0, 0]]))


def test_bbox_transform_matches_golden(golden):
    from libs.boxes import bbox_transform as bt
    g = golden("transform")
    assert_boxes_close(bt.bbox_transform(g["anchors"], g["gt"]), g["enc"])
    assert_boxes_close(bt.bbox_transform_inv(g["anchors"], g["deltas"]), g["dec"])
    assert_boxes_close(bt.bbox_transform_linear(g["anchors"], g["gt"]), g["enc_linear"])
    assert_boxes_close(bt.bbox_transform_inv_linear(g["anchors"], g["deltas"]), g["dec_linear"])
    b = g["dec"].copy()
    r = bt.clip_boxes(b, (600, 800))
    assert r is b                                                  # in place, same array returned
    assert_boxes_close(b, np.stack([np.clip(g["dec"][:, 0], 0, 799), np.clip(g["dec"][:, 1], 0, 599),
                                    np.clip(g["dec"][:, 2], 0, 799), np.clip(g["dec"][:, 3], 0, 599)], 1))
    assert bt.bbox_transform_inv(np.zeros((0, 4), np.float32), np.zeros((0, 8), np.float32)).shape == (0, 8)
    two = bt.bbox_transform_inv(g["anchors"], np.hstack([g["deltas"], g["deltas"]]))      # k-class layout
    assert_boxes_close(two[:, :4], g["dec"]); assert_boxes_close(two[:, 4:], g["dec"])


@pytest.mark.parametrize("name", ["rand1k", "rand3k", "clustered2k", "intgrid600"])
@pytest.mark.parametrize("thr", [0.3, 0.5, 0.7])
def test_nms_modules_match_golden(golden, name, thr):
    from libs.boxes.nms_wrapper import nms
    from libs.nms.cpu_nms import cpu_nms
    from libs.nms.gpu_nms import gpu_nms
    g = golden("nms")
    dets = g[name]
    assert cpu_nms(dets, thr) == list(g["%s_keep_ge_%s" % (name, thr)])          # the reference's cpu_nms
    assert gpu_nms(dets, thr) == list(g["%s_keep_gt_%s" % (name, thr)])          # py_cpu_nms ('>')
    assert nms(dets, thr, force_cpu=True) == list(g["%s_keep_ge_%s" % (name, thr)])
    assert nms(dets[:0], thr) == []
    with pytest.raises(TypeError):
        cpu_nms(dets, 1)                                                          # thresh must be a float


# ------------------------------------------------------------------------------ libs.layers ----
@pytest.mark.parametrize("case", ["fixture", "small_a", "small_noclobber", "small_zero_col", "small_nogt", "small_dup_gt"])
def test_layers_anchor_encode_matches_golden(golden, case):
    import libs.configs.config as cfg
    from libs.layers import anchor
    from libs.layers.data_layer import compute_rpn_targets_in_batch
    g = golden("anchor_target")
    bg, fg, clobber = g[case + "_cfg"]
    cfg.rpn_bg_threshold, cfg.rpn_fg_threshold, cfg.rpn_clobber_positives = float(bg), float(fg), bool(clobber)
    try:
        labels, lw, bt, bw = anchor.encode(g[case + "_gt"], g[case + "_anchors"])
        assert np.array_equal(labels, g[case + "_labels"])
        assert np.array_equal(lw, g[case + "_label_weights"])
        assert_boxes_close(bt, g[case + "_bbox_targets"], extra_scale=1.0)
        assert np.array_equal(bw, g[case + "_bbox_inside_weights"])
        batch = compute_rpn_targets_in_batch([g[case + "_gt"], np.zeros((0, 5), np.float32)], g[case + "_anchors"])
        assert batch[0].shape == (2, labels.shape[0]) and np.array_equal(batch[0][0], labels) and (batch[0][1] == 0).all()
        assert np.array_equal(batch[3][0], bw)
    finally:
        cfg.rpn_bg_threshold, cfg.rpn_fg_threshold, cfg.rpn_clobber_positives = 0.4, 0.6, True


def test_layers_anchor_decode_and_single_shot(golden):
    import libs.configs.config as cfg
    from libs.layers import anchor
    from libs.postprocessings.single_shot import compute_detection
    g = golden("transform")
    rng = np.random.RandomState(3)
    scores = rng.rand(500, 3).astype(np.float32)
    fb, classes, sc = anchor.decode(g["deltas"], scores, g["anchors"], 600, 800, 3)
    want = O.anchor_decode(g["deltas"], scores, g["anchors"], 600, 800, 3)
    assert_boxes_close(fb, want[0]); assert np.array_equal(classes, want[1]) and classes.dtype == np.int32
    # compute_detection == decode + per-class threshold + NMS + top-k, composed from the oracle's pieces
    dets = compute_detection((None, g["deltas"], scores), g["anchors"], 600, 800, 3, score_threshold=0.5,
                             overlap_threshold=0.5, max_det_num=50)
    assert len(dets) == 1
    rows = []
    for j in (1, 2):
        inds = np.where(scores[:, j] > 0.5)[0]
        d = np.hstack([want[0][inds], scores[inds, j][:, None]]).astype(np.float32)
        k = O.nms(d[:, :4], d[:, 4], 0.5, O.NMS_GT_F32)
        rows.append(np.hstack([d[k], np.full((len(k), 1), j)]))
    w = np.vstack(rows); w = w[np.argsort(-w[:, 4], kind="stable")][:50]
    assert np.array_equal(dets[0][:, 4:], w[:, 4:])
    assert_boxes_close(dets[0][:, :4], w[:, :4])


def test_layers_box_torch_functions(golden):
    from libs.layers.box import apply_nms, decoding_box, encoding_box
    g, n = golden("transform"), golden("nms")
    assert_boxes_close(decoding_box(cu(g["deltas"]), cu(g["anchors"])).cpu().numpy(), g["dec"])
    assert_boxes_close(encoding_box(cu(g["gt"]), cu(g["anchors"])).cpu().numpy(), g["enc"])
    dets = n["rand3k"]
    keep = apply_nms(cu(dets[:, :4]), cu(dets[:, 4]), 0.5)
    assert keep.dtype == torch.int64 and keep.is_cuda
    assert keep.cpu().tolist() == list(n["rand3k_keep_gt_0.5"])
    with pytest.raises(detbox.DetboxError):
        apply_nms(torch.from_numpy(dets[:, :4]), torch.from_numpy(dets[:, 4]), 0.5)


def test_anchor_target_module(golden):
    from libs.layers.anchor_target.modules.anchor_target import AnchorTarget
    g = golden("anchor_target")
    labels, deltas, bbw = AnchorTarget(0.3, 0.6, ignored_threshold=0.3)(cu(g["fixture_anchors"]), g["fixture_gt"])
    assert labels.dtype == torch.int64
    assert np.array_equal(labels.cpu().numpy(), g["fixture_labels"].astype(np.int64))
    assert np.array_equal(bbw.cpu().numpy(), g["fixture_bbox_inside_weights"])


# -------------------------------------------------------------------------------- libs.nets ----
def test_nets_model_stage_one():
    import libs.configs.config as cfg
    from libs.nets import model
    rng = np.random.RandomState(5)
    anc = O.anchor_pyramid([8, 16], 128, 256, [[2, 3], [4, 6]], [0.5, 1, 2.0], 16).astype(np.float32)
    A, B = anc.shape[0], 3
    deltas = rng.normal(0, 0.5, (B * A, 4)).astype(np.float32)
    p1 = rng.permutation(np.linspace(0, 1, B * A)).astype(np.float32)
    prob = np.stack([1 - p1, p1], 1).astype(np.float32)
    b, p, ids, a = model._decode_and_choose_top_n_stage1(cu(deltas), cu(prob), cu(anc), top_n=500, activation="softmax")
    wb, wp, wi, wa, _ = O.decode_and_choose_top_n(deltas, prob, anc, 500, "softmax")
    assert_boxes_close(b.cpu().numpy(), wb)
    assert np.array_equal(p.cpu().numpy(), wp) and np.array_equal(ids.cpu().numpy(), wi) and np.array_equal(a.cpu().numpy(), wa)
    kb, kp, kid, ka = model._apply_nms_in_batch(b, p, ids, a, "softmax", overlap_threshold=0.5)
    wk = O.apply_nms_in_batch(b.cpu().numpy(), wp, wi, "softmax", 0.5, O.NMS_GT_F32)
    assert np.array_equal(kp.cpu().numpy(), wp[wk]) and np.array_equal(kid.cpu().numpy(), wi[wk])
    sb, sp, sid = model._stage_one_results(cu(deltas), cu(prob), cu(anc), top_n=500, overlap_threshold=0.5,
                                           top_n_post_nms=100, activation="softmax")
    assert np.array_equal(sp.cpu().numpy(), wp[wk][:100])


# ----------------------------------------------------------------- libs.layers.roi_align(_tf) ----
def test_crop_and_resize_function_matches_golden(golden):
    from libs.layers.roi_align_tf.crop_and_resize import CropAndResizeFunction
    g = golden("roialign")
    img = cu(g["image"]).requires_grad_(True)
    out = CropAndResizeFunction(7, 7, 0)(img, cu(g["boxes"]), cu(g["box_ind"]))
    assert np.array_equal(out.detach().cpu().numpy(), g["crops7"])               # bit-exact vs crop_and_resize.c
    out.backward(cu(g["grad"]))
    np.testing.assert_allclose(img.grad.cpu().numpy(), g["grad_image7"], rtol=1e-5, atol=1e-5)
    assert np.array_equal(CropAndResizeFunction(3, 5, 0)(cu(g["image"]), cu(g["boxes"]), cu(g["box_ind"])).cpu().numpy(), g["crops3x5"])
    assert np.array_equal(CropAndResizeFunction(1, 1, 0)(cu(g["image"]), cu(g["boxes"]), cu(g["box_ind"])).cpu().numpy(), g["crops1"])
    ext = CropAndResizeFunction(7, 7, -3.5)(cu(g["image"]), cu(g["boxes"]), cu(g["box_ind"])).cpu().numpy()
    assert np.array_equal(ext, O.crop_and_resize_fwd(g["image"], g["boxes"], g["box_ind"], -3.5, 7, 7))


def test_crop_and_resize_at_the_reference_test_shape():
    """The reference's own comparison recipe (roi_align_tf/test.py:25-56): image 16 x 256 x 200 x 300, 100 random boxes in
    normalised (y1, x1, y2, x2) form, 7 x 7 crops, forward and backward — here against the restated crop_and_resize.c
    (pinned bit-exactly to the compiled reference): forward bit-exact, backward within the atomic-order tolerance."""
    from libs.layers.roi_align_tf.crop_and_resize import CropAndResizeFunction
    rng = np.random.RandomState(1234)
    B, C, H, W, n = 16, 256, 200, 300, 100
    xs = np.sort(rng.uniform(0, W, (n, 2)) / (W - 1), axis=1); ys = np.sort(rng.uniform(0, H, (n, 2)) / (H - 1), axis=1)
    boxes = np.stack((ys[:, 0], xs[:, 0], ys[:, 1], xs[:, 1]), axis=-1).astype(np.float32)      # some reach past 1: extrapolation
    ind = rng.randint(0, B, n).astype(np.int32)
    image = rng.randn(B, C, H, W).astype(np.float32)
    img = cu(image).requires_grad_(True)
    out = CropAndResizeFunction(7, 7, 0)(img, cu(boxes), cu(ind))
    assert np.array_equal(out.detach().cpu().numpy(), O.crop_and_resize_fwd(image, boxes, ind, 0.0, 7, 7))
    g = rng.randn(n, C, 7, 7).astype(np.float32)
    out.backward(cu(g))
    np.testing.assert_allclose(img.grad.cpu().numpy(), O.crop_and_resize_bwd(g, boxes, ind, image.shape), rtol=1e-4, atol=1e-5)


def make_pyramid(rng, B=2, C=6, ih=128, iw=256, strides=(4, 8, 16, 32)):
    return [rng.normal(0, 1, (B, C, ih // s, iw // s)).astype(np.float32) for s in strides]


def make_rois(rng, R, B, ih=128, iw=256):
    wh = np.exp(rng.uniform(np.log(6), np.log(120), (R, 2)))
    c = rng.uniform(0, 1, (R, 2)) * [iw, ih]
    boxes = np.concatenate([c - wh / 2, c + wh / 2], 1).astype(np.float32)
    return boxes, rng.randint(0, B, R).astype(np.int64)


def test_roi_align_tf_modules_vs_oracle():
    import libs.configs.config as cfg
    from libs.layers.roi_align_tf.pyramid_roi_align import PyramidRoIAlign, PyramidRoIAlign2
    from libs.layers.roi_align_tf.roi_align_tf import RoIAlign
    rng = np.random.RandomState(8)
    strides = (4, 8, 16, 32)
    old = (cfg.strides, cfg.anchor_scales)
    cfg.strides, cfg.anchor_scales = strides, [[2, 2.52, 3.17], [4, 5.04, 6.35], [8, 10.08, 12.70], [16, 20.16, 25.40]]
    try:
        feats = make_pyramid(rng)
        boxes, bidx = make_rois(rng, 40, 2)
        rois = np.concatenate([bidx[:, None].astype(np.float32), boxes], 1)
        f0 = cu(feats[1]).requires_grad_(True)
        y = RoIAlign(7, 7)(f0, cu(rois), 1.0 / 8)
        assert np.array_equal(y.detach().cpu().numpy(), O.roi_align_tf_fwd(feats[1], rois, np.float32(1.0 / 8), 7, 7))
        gy = rng.normal(0, 1, y.shape).astype(np.float32)
        y.backward(cu(gy))
        np.testing.assert_allclose(f0.grad.cpu().numpy(), O.roi_align_tf_bwd(gy, rois, np.float32(1.0 / 8), feats[1].shape), rtol=1e-4, atol=1e-5)
        ra = (np.asarray(cfg.anchor_scales, np.float32).mean(axis=1) * cfg.anchor_base) ** 2
        pyr = [cu(f).requires_grad_(True) for f in feats]
        out = PyramidRoIAlign(7, 7)(pyr, cu(boxes), cu(bidx))
        want = O.pyramid_roi_align(feats, boxes, bidx, strides, ra, 7, 7, "tf")
        assert np.array_equal(out.detach().cpu().numpy(), want)
        g = rng.normal(0, 1, want.shape).astype(np.float32)
        out.backward(cu(g))
        lv = O.assign_levels(boxes, ra)
        for l, s in enumerate(strides):
            sel = lv == l
            wg = O.roi_align_tf_bwd(g[sel], rois[sel], np.float32(1.0 / s), feats[l].shape) if sel.any() else np.zeros_like(feats[l])
            np.testing.assert_allclose(pyr[l].grad.cpu().numpy(), wg, rtol=1e-4, atol=1e-5)
        m2 = PyramidRoIAlign2(7, 7, num_channels=6).cuda()
        crops = m2.crop([cu(f) for f in feats], cu(boxes), cu(bidx))
        assert np.array_equal(crops.cpu().numpy(), O.pyramid_roi_align(feats, boxes, bidx, strides, ra, 7, 7, "tf", all_levels=True))
        assert m2([cu(f) for f in feats], cu(boxes), cu(bidx)).shape == (40, 6, 7, 7)
    finally:
        cfg.strides, cfg.anchor_scales = old


@pytest.mark.parametrize("pool", [None, "avg", "max"])
def test_roi_align_grid_modules_vs_oracle(pool):
    import torch.nn.functional as F
    from libs.layers.roi_align.modules.roi_align import RoIAlign, RoIAlignAvg, RoIAlignMax
    rng = np.random.RandomState(9)
    feat = rng.normal(0, 1, (2, 5, 24, 40)).astype(np.float32)
    boxes, bidx = make_rois(rng, 30, 2, 96, 160)
    rois = np.concatenate([bidx[:, None].astype(np.float32), boxes], 1)
    mod = {None: RoIAlign, "avg": RoIAlignAvg, "max": RoIAlignMax}[pool](7, 7)
    f = cu(feat).requires_grad_(True)
    y = mod(f, cu(rois), 0.25)
    P = 7 if pool is None else 8
    grid = O.roi_align_grid_fwd(feat, rois, np.float32(0.25), P, P)
    want = grid if pool is None else O.pool2x2(grid, pool)
    np.testing.assert_allclose(y.detach().cpu().numpy(), want, rtol=1e-5, atol=1e-6)
    gy = rng.normal(0, 1, y.shape).astype(np.float32)
    y.backward(cu(gy))
    # the reference composite: pool backward (torch autograd on the oracle's grid) then the grid-point scatter
    if pool is None:
        ggrid = gy
    else:
        t = torch.from_numpy(grid).requires_grad_(True)
        (F.avg_pool2d(t, 2, 1) if pool == "avg" else F.max_pool2d(t, 2, 1)).backward(torch.from_numpy(gy))
        ggrid = t.grad.numpy()
    wg = O.roi_align_grid_bwd(ggrid, rois, np.float32(0.25), feat.shape)
    np.testing.assert_allclose(f.grad.cpu().numpy(), wg, rtol=1e-4, atol=1e-4)


def test_single_shot_new_advanced_stitch(golden):
    """compute_detection_new / nms_advanced / nms_stitch (single_shot.py:62-177) against the same compositions
    over the oracle's pieces."""
    import libs.configs.config as cfg
    from libs.postprocessings.single_shot import compute_detection_new, nms_advanced, nms_stitch
    g = golden("transform")
    rng = np.random.RandomState(5)
    n = g["deltas"].shape[0]
    probs = rng.dirichlet([1.0, 0.6, 0.6], n).astype(np.float32)
    img = np.sort(rng.randint(0, 2, n))
    dets = compute_detection_new(g["deltas"], probs, g["anchors"], img, 600, 800, 3, score_threshold=0.3,
                                 overlap_threshold=0.5, max_dets=40, batch_size=2)
    assert len(dets) == 2
    for i in range(2):
        inds = np.where(img == i)[0]
        fb = O.anchor_decode(g["deltas"][inds], probs[inds], g["anchors"][inds], 600, 800, 3)[0]
        obj = 1 - probs[inds][:, 0]
        od = np.hstack([fb, obj[:, None]]).astype(np.float32)
        keep = np.asarray(O.nms(od[:, :4], od[:, 4], 0.5, O.NMS_GT_F32))
        od = od[keep]
        keep = keep[:40]
        po = probs[inds][keep][:, 1:]
        d0, d1 = np.where(po > 0.3)
        w = np.zeros((d0.size, 6)); w[:, :4] = od[d0, :4]; w[:, 4] = po[d0, d1]; w[:, 5] = d1 + 1
        w = w[np.argsort(-w[:, 4], kind="stable")]
        assert np.array_equal(dets[i][:, 4:], w[:, 4:])
        assert_boxes_close(dets[i][:, :4], w[:, :4])
    # nms_advanced: per class nms_multianchor with the crowd / non-crowd thresholds
    D = []
    for s in (0, 1):
        b = np.round(rng.uniform(0, 120, (300, 2))); wh = np.round(np.exp(rng.uniform(np.log(6), np.log(60), (300, 2))))
        D.append(np.hstack([b, b + wh, rng.rand(300, 1), rng.choice([1, 2, 5], 300)[:, None]]))
    got = nms_advanced(D)
    for dets_i, got_i in zip(D, got):
        rows = []
        cls = dets_i[:, 5].astype(np.int32)
        for c in np.unique(cls):
            thr = cfg.overlap_threshold_noncrowd if c in cfg.noncrowd_classes else cfg.overlap_threshold
            dt = dets_i[cls == c]
            rows.append(dt[O.nms_multianchor(dt[:, :5].astype(np.float32), thr)[0]])
        assert np.array_equal(got_i, np.vstack(rows))
    got = nms_stitch(D)
    for dets_i, got_i in zip(D, got):
        p = dets_i[:, 4].astype(np.float32)
        inds = np.where((p > 0.1) & (p <= 0.5))[0]
        keep = O.nms_multianchor(dets_i[:, :5].astype(np.float32), 0.95)[0]
        assert np.array_equal(got_i, dets_i[np.setdiff1d(np.arange(p.size), np.setdiff1d(inds, keep))])


def test_pyramid_roi_align_reference_fixture():
    """The reference's own script (roi_align_tf/pyramid_roi_align.py:96-135): five boxes that land one per level of the
    default 5-level pyramid, level i's plane holding the plateau value i under its box; then the same call with the boxes
    permuted.  Box i must come back from level i (its 5x5 crop reads the plateau), in the caller's order either way."""
    import libs.configs.config as cfg
    from libs.layers.roi_align_tf.pyramid_roi_align import PyramidRoIAlign
    assert tuple(cfg.strides) == (8, 16, 32, 64, 128)
    bboxes = np.asarray([[100, 100, 120, 130], [10, 10, 70, 70], [200, 200, 350, 350], [0, 0, 250, 250], [0, 0, 500, 500]], np.float32)
    rng = np.random.RandomState(0)
    pyramid, batch_inds = [], []
    for i, s in enumerate(cfg.strides):
        t = np.zeros((2, 1, 512 // s, 512 // s), np.float32)
        ind = int(rng.randint(0, 2))
        box = bboxes[i] // s
        t[ind, :, int(box[1]):int(box[3] + 1), int(box[0]):int(box[2] + 1)] = i
        pyramid.append(t); batch_inds.append(ind)
    batch_inds = np.asarray(batch_inds, np.float32)
    ra = (np.asarray(cfg.anchor_scales, np.float32).mean(axis=1) * cfg.anchor_base) ** 2
    np.testing.assert_array_equal(O.assign_levels(bboxes, ra), np.arange(5))            # one box per level
    crop = PyramidRoIAlign(5, 5)
    out = crop([cu(f) for f in pyramid], cu(bboxes), cu(batch_inds)).cpu().numpy()
    want = O.pyramid_roi_align(pyramid, bboxes, batch_inds, cfg.strides, ra, 5, 5, "tf")
    assert out.shape == (5, 1, 5, 5) and np.array_equal(out, want)
    for i in range(5):
        assert out[i].max() == i and (i == 0 or out[i, 0, 1:4, 1:4].min() > 0), "box %d did not read level %d" % (i, i)
    perm = rng.permutation(5)
    out_p = crop([cu(f) for f in pyramid], cu(bboxes[perm]), cu(batch_inds[perm])).cpu().numpy()
    assert np.array_equal(out_p, out[perm])


def test_compute_rpn_targets_device_entries(golden):
    """the device-side hooks — compute_rpn_targets_in_batch_cuda (list of CUDA GT tensors) and
    compute_rpn_targets_padded_cuda ([B, Gmax, 5] + counts, no host concatenation) — against the host-list entry and the
    oracle: labels / weights bit-exact, targets identical bits across the three entries."""
    import libs.configs.config as cfg
    from libs.layers.data_layer import (compute_rpn_targets_in_batch, compute_rpn_targets_in_batch_cuda,
                                        compute_rpn_targets_padded_cuda)
    from oracle.make_golden import synth_gt
    cfg.rpn_bg_threshold, cfg.rpn_fg_threshold, cfg.rpn_clobber_positives = 0.5, 0.6, True
    try:
        args = ([8, 16, 32], 192, 320, [[2, 2.52, 3.175], [4, 5.04, 6.349], [8, 10.079, 12.7]], [1, 2.5], 16)
        anchors = detbox.anchor_pyramid(*args)
        pyr = detbox.pyramid_desc(*args)
        anc = anchors.cpu().numpy()
        rng = np.random.RandomState(8)
        counts = [5, 0, 17, 1, 40]
        gts = [synth_gt(rng, G, 192, 320, 0.2) if G else np.zeros((0, 5), np.float32) for G in counts]
        host = compute_rpn_targets_in_batch(gts, anc)
        dev = compute_rpn_targets_in_batch_cuda([cu(g) for g in gts], anchors, pyramid=pyr)
        Gmax = 48
        padded = np.full((len(gts), Gmax, 5), np.nan, np.float32)          # the padding must never be read
        for b, g in enumerate(gts):
            padded[b, :g.shape[0]] = g
        pad = compute_rpn_targets_padded_cuda(cu(padded), cu(np.asarray(counts, np.int32)), anchors, pyramid=pyr)
        for h, d, q in zip(host, dev, pad):
            assert d.is_cuda and q.is_cuda
            assert np.array_equal(h.view(np.uint32), d.cpu().numpy().view(np.uint32))
            assert np.array_equal(h.view(np.uint32), q.cpu().numpy().view(np.uint32))
        ocfg = O.Cfg(rpn_bg_threshold=0.5, rpn_fg_threshold=0.6)
        for b, g in enumerate(gts):
            want = O.anchor_encode(g, anc, ocfg)
            assert np.array_equal(host[0][b], want[0]) and np.array_equal(host[1][b], want[1]) and np.array_equal(host[3][b], want[3])
            assert_boxes_close(host[2][b], want[2], extra_scale=1.0)
        packed, bt, cnt = compute_rpn_targets_padded_cuda(cu(padded), cu(np.asarray(counts, np.int32)), anchors, pyramid=pyr, format="compact")
        labels, lw, bw = detbox.expand_targets(packed.cpu(), cnt.cpu())
        assert np.array_equal(labels.numpy(), host[0]) and np.array_equal(lw.numpy(), host[1]) and np.array_equal(bw.numpy(), host[3])
        assert np.array_equal(bt.cpu().numpy().view(np.uint32), host[2].view(np.uint32))
    finally:
        cfg.rpn_bg_threshold, cfg.rpn_fg_threshold, cfg.rpn_clobber_positives = 0.4, 0.6, True


# ---- round 2: the remaining rows of SURVEY 8a / 8f ------------------------------------------------------------------
def test_layers_roi_encode_decode_match_reference(golden):
    """libs/layers/roi.py:16-125 (8f-1): encode with np.random seeded like the golden run of the COMPILED reference —
    labels identical (same draws in the same order), inside weights identical, targets within the float32 bar; the
    no-GT branch; decode (arg-max class box, clipped)."""
    import libs.configs.config as cfg
    from libs.layers import roi
    g = golden("r2_rows")
    cfg.fg_threshold, cfg.roi_bg_threshold, cfg.rois_per_image, cfg.roi_fg_fraction = 0.5, 0.3, 128, 0.25
    for name in ("roi_a", "roi_b", "roi_nogt"):
        np.random.seed(int(g[name + "_seed"]))
        lab, bt, bw = roi.encode(g[name + "_gt"].copy(), g[name + "_rois"].copy(), 5)
        assert lab.dtype == np.float32 and np.array_equal(lab, g[name + "_labels"]), name
        assert np.array_equal(bw, g[name + "_bbox_inside_weights"]), name
        assert bt.shape == g[name + "_bbox_targets"].shape
        np.testing.assert_allclose(bt, g[name + "_bbox_targets"], rtol=1e-5, atol=1e-5)
    fb, cls, sc = roi.decode(g["dec_deltas"].copy(), g["dec_scores"].copy(), g["dec_rois"].copy(), 512, 768)
    assert np.array_equal(cls, g["dec_classes"]) and cls.dtype == np.int32 and np.array_equal(sc, g["dec_max"])
    assert_boxes_close(fb, g["dec_boxes"])


@pytest.mark.parametrize("name", ["trad_a", "trad_b", "trad_c"])
def test_anchor_encode_traditional_sampling(golden, name):
    """libs/layers/anchor.py:95-109 'traditional': the device assignment + the reference's two np.random.choice draws in
    its order — labels identical to the compiled reference's seeded run, label weights all zero (only 'simple' fills
    them), inside weights unaffected by the sampling; also through the batched entry."""
    import libs.configs.config as cfg
    from libs.layers import anchor
    from libs.layers.data_layer import compute_rpn_targets_in_batch
    g = golden("r2_rows")
    cfg.rpn_bg_threshold, cfg.rpn_fg_threshold, cfg.rpn_clobber_positives = 0.4, 0.6, True
    cfg.rpn_sample_strategy, cfg.rpn_batch_size, cfg.rpn_fg_fraction = 'traditional', int(g[name + "_batch_size"]), 0.25
    try:
        np.random.seed(int(g[name + "_seed"]))
        lab, lw, bt, bw = anchor.encode(g[name + "_gt"].copy(), g["trad_anchors"].copy())
        assert np.array_equal(lab, g[name + "_labels"]) and not lw.any() and np.array_equal(lw, g[name + "_label_weights"])
        assert np.array_equal(bw, g[name + "_bbox_inside_weights"])
        assert_boxes_close(bt, g[name + "_bbox_targets"], extra_scale=1.0)
        np.random.seed(int(g[name + "_seed"]))
        batch = compute_rpn_targets_in_batch([g[name + "_gt"].copy()], g["trad_anchors"].copy())
        assert np.array_equal(batch[0][0], g[name + "_labels"]) and not batch[1].any()
    finally:
        cfg.rpn_sample_strategy, cfg.rpn_batch_size = 'simple', 384


def test_assign_boxes_and_linear_alpha(golden):
    """libs/layers/assign.py:10-43 (a21) incl. areas whose sqrt / base is an exact power of two, and
    libs/boxes/bbox_transform.py:93-144 with alpha = 4 (encode bit-exact, decode within the 1e-5 bar) and alpha = 10."""
    from libs.boxes.bbox_transform import bbox_transform_inv_linear, bbox_transform_linear
    from libs.layers.assign import assign_boxes
    g = golden("r2_rows")
    assert np.array_equal(assign_boxes(g["assign_boxes"].copy()), g["assign_default"])
    k = assign_boxes(g["assign_boxes"].copy(), 3, 6, 112.0)
    assert k.dtype == np.int32 and np.array_equal(k, g["assign_k36"])
    assert assign_boxes(np.zeros((0, 5), np.float32)).shape == (0,)
    for al in (4.0, 10.0):
        np.testing.assert_array_equal(bbox_transform_linear(g["lin_anchors"].copy(), g["lin_gt"].copy(), al), g["lin_enc_%g" % al])
        assert_boxes_close(bbox_transform_inv_linear(g["lin_anchors"].copy(), g["lin_deltas"].copy(), al), g["lin_dec_%g" % al])


def test_compute_detection_batched_equals_per_pair_loop():
    """single_shot.compute_detection (:13-60) runs every (image, class) pair as ONE segmented NMS launch: the result must
    equal upstream's per-image x per-class loop composed from the oracle's pieces (decode + clip, per-class threshold,
    '>' NMS, class-major concatenation, score sort, truncation) — 3 images, 5 classes, one class without candidates."""
    from libs.postprocessings.single_shot import compute_detection
    rng = np.random.RandomState(12)
    n, A, C = 3, 700, 5
    anchors = synth_boxes(rng, A)
    deltas = rng.normal(0, 0.3, (n * A, 4)).astype(np.float32)
    scores = rng.rand(n * A, C).astype(np.float32)
    scores[:, 3] *= 0.4                                                   # class 3 never passes the 0.5 threshold
    scores[A:2 * A, 1] *= 0.45                                            # image 1 has no class-1 candidate either
    dets = compute_detection((None, deltas, scores), anchors, 480, 640, C, score_threshold=0.5, overlap_threshold=0.5, max_det_num=60)
    assert len(dets) == n
    for i in range(n):
        want = O.anchor_decode(deltas[i * A:(i + 1) * A], scores[i * A:(i + 1) * A], anchors, 480, 640, C)[0]
        rows = []
        for j in range(1, C):
            inds = np.where(scores[i * A:(i + 1) * A, j] > 0.5)[0]
            d = np.hstack([want[inds], scores[i * A:(i + 1) * A][inds, j][:, None]]).astype(np.float32)
            k = O.nms(d[:, :4], d[:, 4], 0.5, O.NMS_GT_F32) if d.shape[0] else np.zeros((0,), np.int64)
            rows.append(np.hstack([d[k], np.full((len(k), 1), j)]))
        w = np.vstack(rows); w = w[np.argsort(-w[:, 4], kind="stable")][:60]
        assert dets[i].shape == w.shape and np.array_equal(dets[i][:, 4:], w[:, 4:]), i
        assert_boxes_close(dets[i][:, :4], w[:, :4])


def synth_boxes(rng, n):
    h = np.exp(rng.uniform(np.log(16), np.log(200), n)); w = h / rng.uniform(0.5, 2.5, n)
    cx, cy = rng.uniform(0, 640, n), rng.uniform(0, 480, n)
    return np.stack([cx - w / 2, cy - h / 2, cx + w / 2, cy + h / 2], 1).astype(np.float32)
