"""The CPU restatement (oracle/detbox_oracle.c + oracle/oracle.py) against the golden vectors
that oracle/make_golden.py produced from the UNMODIFIED reference.  CPU only.

Bars: bit-exact for IoU (f64), labels, keep indices, anchors; float32 box math and RoIAlign
within rtol 1e-5 (+ atol 1e-6), the tolerance BASELINE.json's north_star states.
"""
import numpy as np
import pytest

from oracle import oracle as O
from util import assert_boxes_close

RTOL, ATOL = 1e-5, 1e-6
CPR = O.citypersons_retinanet_cfg()


def test_base_anchors(golden):
    g = golden("anchors")
    for lvl in range(5):
        np.testing.assert_array_equal(O.anchors(CPR.anchor_scales[lvl], CPR.anchor_ratios, 16), g["base_cpr_l%d" % lvl])
    np.testing.assert_array_equal(O.anchors([2., 4., 8., 16., 32.], [0.5, 1, 2.0], 16), g["base_default"])
    np.testing.assert_array_equal(O.anchors([1.5, 3], [0.5, 1, 2.0, 3.0], 15), g["base_odd15"])
    # SURVEY KAT-1
    np.testing.assert_allclose(g["base_cpr_l0"][1], [-20.91999984, -20.91999984, 20.91999984, 20.91999984], atol=1e-8)


def test_anchor_pyramid(golden):
    g = golden("anchors")
    pyr = O.anchor_pyramid(CPR.strides, 1024, 2048, CPR.anchor_scales, CPR.anchor_ratios, 16)
    assert tuple(g["cpr_pyramid_shape"]) == pyr.shape == (261888, 4)
    np.testing.assert_array_equal(pyr[g["cpr_pyramid_rows_idx"]], g["cpr_pyramid_rows"])
    np.testing.assert_array_equal(pyr.sum(0), g["cpr_pyramid_colsum"])
    np.testing.assert_array_equal(pyr[6], [-8.5, -16.5, 24.5, 16.5])
    small = O.anchor_pyramid([4, 8, 16], 64, 96, [[2, 3], [4, 6], [8, 12]], [0.5, 1, 2.0], 16, [[0.0, 0.0], [0.5, 0.25]])
    np.testing.assert_array_equal(small, g["small_pyramid"])
    np.testing.assert_array_equal(O.anchor_one_plane(5, 7, 8, [2., 4.], [0.5, 2.0], 16, [[0, 0], [0.5, 0]]), g["plane_shifted"])


def test_iou(golden):
    g = golden("iou")
    np.testing.assert_array_equal(O.bbox_overlaps(g["fix_anchors"], g["fix_gt"][:, :4]), g["fix_ov"])
    np.testing.assert_array_equal(O.bbox_overlaps(g["boxes"], g["query"]), g["ov"])
    np.testing.assert_array_equal(O.bbox_intersections(g["boxes"], g["query"]), g["inter"])
    np.testing.assert_array_equal(O.bbox_exclude_ignored_areas(g["boxes"], g["query"], 0.2), g["excl"])
    np.testing.assert_array_equal(O.bbox_overlaps(g["boxes_int"], g["query_int"]), g["ov_int"])
    # SURVEY KAT-2
    assert abs(g["fix_ov"][0, 0] - 0.65389) < 1e-5 and abs(g["fix_ov"][5, 2] - 0.826446) < 1e-6


def test_transform(golden):
    g = golden("transform")
    np.testing.assert_allclose(O.bbox_transform(g["anchors"], g["gt"]), g["enc"], rtol=RTOL, atol=ATOL)
    assert_boxes_close(O.bbox_transform_inv(g["anchors"], g["deltas"]), g["dec"])
    assert_boxes_close(O.bbox_transform_inv(g["anchors"], g["deltas"], im_shape=(1024, 2048)), g["dec_clip"])
    np.testing.assert_allclose(O.bbox_transform(g["anchors"], g["gt"], "linear"), g["enc_linear"], rtol=RTOL, atol=ATOL)
    assert_boxes_close(O.bbox_transform_inv(g["anchors"], g["deltas"], "linear"), g["dec_linear"])


CASES = ["fixture", "small_a", "small_noclobber", "small_zero_col", "small_zero_col_noclobber", "small_nogt", "small_dup_gt"]


@pytest.mark.parametrize("name", CASES)
def test_anchor_target(golden, name):
    g = golden("anchor_target")
    bg, fg, clob = g[name + "_cfg"]
    cfg = O.Cfg(rpn_bg_threshold=float(bg), rpn_fg_threshold=float(fg), rpn_clobber_positives=bool(clob))
    lab, lw, bt, bw = O.anchor_encode(g[name + "_gt"], g[name + "_anchors"], cfg)
    np.testing.assert_array_equal(lab, g[name + "_labels"])
    np.testing.assert_array_equal(lw, g[name + "_label_weights"])
    np.testing.assert_array_equal(bw, g[name + "_bbox_inside_weights"])
    np.testing.assert_allclose(bt, g[name + "_bbox_targets"], rtol=RTOL, atol=ATOL)


def test_anchor_target_kat3(golden):
    g = golden("anchor_target")
    np.testing.assert_array_equal(g["fixture_labels"], [2, 0, 0, 0, 0, -1, -1, 0, 0, 0])
    np.testing.assert_allclose(g["fixture_bbox_targets"][0], [1.190476, 1.129032, 0.667657, 0.461867], atol=1e-5)


def test_anchor_target_cpr_image(golden):
    g = golden("anchor_target")
    pyr = O.anchor_pyramid(CPR.strides, 1024, 2048, CPR.anchor_scales, CPR.anchor_ratios, 16).astype(np.float32)
    lab, lw, bt, bw = O.anchor_encode(g["cpr_gt"], pyr, CPR)
    ref = np.zeros_like(lab)
    ref[g["cpr_labels_nz_idx"]] = g["cpr_labels_nz_val"]
    np.testing.assert_array_equal(lab, ref)
    np.testing.assert_array_equal([(lab >= 1).sum(), (lab == 0).sum(), (lab < 0).sum()], g["cpr_counts"])
    np.testing.assert_array_equal(lw[lab == 0][:1], g["cpr_bg_weight"])
    np.testing.assert_allclose(bt[g["cpr_rows_idx"]], g["cpr_bbox_targets_rows"], rtol=RTOL, atol=ATOL)
    np.testing.assert_array_equal(bw.sum(0), g["cpr_bbox_inside_weights_sum"])


def test_anchor_target_real_citypersons_gt(golden):
    """GT sets of real CityPersons validation images (6..60 boxes, up to 31 ignore regions) against the
    CP-R pyramid — outputs recorded from the compiled reference (oracle/make_golden.py::real_gt_golden).  Image 0 has
    35 positive / 259,248 negative / 2,605 ignored anchors."""
    g = golden("anchor_target_real")
    pyr = O.anchor_pyramid(CPR.strides, 1024, 2048, CPR.anchor_scales, CPR.anchor_ratios, 16).astype(np.float32)
    np.testing.assert_array_equal(g["img0_counts"], [35, 259248, 2605])
    for i in g["image_ids"][:3]:                       # (three of the six keep the CPU suite short; the GPU test runs all)
        k = "img%d_" % i
        lab, lw, bt, bw = O.anchor_encode(g[k + "gt"], pyr, CPR)
        ref = np.zeros_like(lab)
        ref[g[k + "labels_nz_idx"]] = g[k + "labels_nz_val"]
        np.testing.assert_array_equal(lab, ref)
        np.testing.assert_array_equal([(lab >= 1).sum(), (lab == 0).sum(), (lab < 0).sum()], g[k + "counts"])
        np.testing.assert_array_equal(lw[lab == 0][:1], g[k + "bg_weight"])
        np.testing.assert_array_equal(np.nonzero(bw.any(1))[0], g[k + "inside_rows"])
        np.testing.assert_allclose(bt[g[k + "rows_idx"]], g[k + "bbox_targets_rows"], rtol=RTOL, atol=ATOL)


@pytest.mark.parametrize("name", ["rand1k", "rand3k", "clustered2k", "intgrid600"])
@pytest.mark.parametrize("thr", [0.5, 0.3, 0.7])
def test_nms(golden, name, thr):
    g = golden("nms")
    d = g[name]
    np.testing.assert_array_equal(O.nms(d[:, :4], d[:, 4], thr, O.NMS_GE_F64), g["%s_keep_ge_%g" % (name, thr)])
    np.testing.assert_array_equal(O.nms(d[:, :4], d[:, 4], thr, O.NMS_GT_F32), g["%s_keep_gt_%g" % (name, thr)])
    np.testing.assert_array_equal(O.nms(d[:, :4], d[:, 4], thr, O.NMS_GE_F64), g["%s_keep_cynms_%g" % (name, thr)])


def test_nms_comparators_differ_on_integer_grid(golden):
    g = golden("nms")
    assert len(g["intgrid600_keep_ge_0.5"]) != len(g["intgrid600_keep_gt_0.5"])


def test_roialign_crop_and_resize(golden):
    g = golden("roialign")
    img, boxes, ind = g["image"], g["boxes"], g["box_ind"]
    np.testing.assert_array_equal(O.crop_and_resize_fwd(img, boxes, ind, 0.0, 7, 7), g["crops7"])
    np.testing.assert_array_equal(O.crop_and_resize_fwd(img, boxes, ind, 0.0, 3, 5), g["crops3x5"])
    np.testing.assert_array_equal(O.crop_and_resize_fwd(img, boxes, ind, 0.0, 1, 1), g["crops1"])
    np.testing.assert_array_equal(O.crop_and_resize_bwd(g["grad"], boxes, ind, img.shape), g["grad_image7"])


def test_roialign_kat5(golden):
    g = golden("roialign")
    ramp = np.arange(16, dtype=np.float32).reshape(1, 1, 4, 4)
    rois = np.asarray([[0, 0, 0, 3, 3]], np.float32)
    boxes, ind = O.tf_boxes(rois, 1.0, 2, 2, 4, 4)
    np.testing.assert_array_equal(boxes, g["kat5_box"])
    out = O.roi_align_tf_fwd(ramp, rois, 1.0, 2, 2)
    np.testing.assert_array_equal(out, g["kat5_fwd"])
    np.testing.assert_allclose(out.ravel(), [1.25, 2.75, 7.25, 8.75], atol=1e-6)
    gi = O.roi_align_tf_bwd(np.ones((1, 1, 2, 2), np.float32), rois, 1.0, ramp.shape)
    np.testing.assert_array_equal(gi, g["kat5_bwd"])
    np.testing.assert_allclose(gi[0, 0, 0], [.5625, .375, .5625, 0], atol=1e-6)


def _roi_case(g, c):
    ngt = g["c%d_ngt" % c]
    gts = np.split(g["c%d_gt" % c], np.cumsum(ngt)[:-1]) if ngt.size else []
    return g["c%d_boxes" % c], g["c%d_ids" % c], [x.reshape(-1, 5) for x in gts]


@pytest.mark.parametrize("case", range(6))
def test_roi_matching_and_sampling(golden, case):
    """libs/layers/box.py matching_box / sample_rois outputs recorded from the reference (make_golden.py)."""
    g = golden("roi_sampling")
    b, ids, gts = _roi_case(g, case)
    lab, match, mbox = O.matching_box(b, ids, gts, 0.4, 0.6)
    np.testing.assert_array_equal(lab, g["c%d_match_labels" % case])
    np.testing.assert_array_equal(match, g["c%d_match_inds" % case])
    np.testing.assert_array_equal(mbox, g["c%d_match_boxes" % case])
    sb, sl, si = O.sample_rois(b, ids, gts, 0.5, 16, 0.25, 0.2, rng=np.random.RandomState(100 + case))
    np.testing.assert_array_equal(sb, g["c%d_sample_boxes" % case])
    np.testing.assert_array_equal(sl, g["c%d_sample_labels" % case])
    np.testing.assert_array_equal(si, g["c%d_sample_ids" % case])


@pytest.mark.parametrize("seed", range(4))
@pytest.mark.parametrize("thr", [0.3, 0.5])
def test_nms_variants(golden, seed, thr):
    g = golden("nms_variants")
    d = g["d%d" % seed]
    np.testing.assert_array_equal(O.nms_new(d, thr), g["d%d_new_%g" % (seed, thr)])
    keep, clusters = O.nms_multianchor(d, thr)
    np.testing.assert_array_equal(keep, g["d%d_multi_keep_%g" % (seed, thr)])
    np.testing.assert_array_equal(clusters, g["d%d_multi_clusters_%g" % (seed, thr)])


def test_dets_cluster(golden):
    """dets_cluster (single_shot.py:401-415): the oracle against outputs of the reference's own function.  The sums are
    NumPy's on both sides, so the comparison is exact with the NumPy that recorded the vectors and within float32
    rounding of the pairwise order otherwise."""
    g = golden("dets_cluster")
    nv = golden("nms_variants")
    exact = str(g["numpy_version"]) == np.__version__
    cases = [(nv["d%d" % s], nv["d%d_multi_keep_%g" % (s, t)], nv["d%d_multi_clusters_%g" % (s, t)], g["d%d_voted_%g" % (s, t)])
             for s in range(4) for t in (0.3, 0.5)]
    cases.append((g["synth_dets"], g["synth_keep"], g["synth_clusters"], g["synth_voted"]))
    for d, keep, clusters, want in cases:
        got = O.dets_cluster(d, keep, clusters)
        assert got.shape == want.shape and got.dtype == np.float32
        if exact:
            np.testing.assert_array_equal(got, want)
        else:
            np.testing.assert_allclose(got, want, rtol=1e-6, atol=1e-4)
    assert O.dets_cluster(cases[0][0], [], cases[0][2]) is not None and O.dets_cluster(cases[0][0], [], cases[0][2]).shape == cases[0][0].shape


def test_to_dets(golden):
    """to_Dets* conversions (libs/nets/model.py:357-469): outputs recorded from the reference's own methods."""
    from oracle.make_golden import to_dets_cases, to_dets_inputs
    g = golden("to_dets")
    for case in to_dets_cases():
        name, act, multi, C, thr, B, _ = case
        b, p, i = to_dets_inputs(case)
        assert np.array_equal(b, g[name + "_boxes"]) and np.array_equal(p, g[name + "_probs"]) and np.array_equal(i, g[name + "_ids"])
        d = O.to_dets(b, p, i, B, act, multi, thr)
        assert np.array_equal(np.cumsum([0] + [x.shape[0] for x in d]), g[name + "_offsets"])
        assert np.array_equal(np.concatenate(d, 0), g[name + "_dets"])


def test_anchor_target_real_gt64_counts(golden):
    """The oracle against the compiled reference's recorded counts on real CityPersons GT (tests/golden/real_gt64.npz):
    the densest image, an image without GT and two more, at full CP-R size."""
    g = golden("real_gt64")
    cfg = O.citypersons_retinanet_cfg()
    anc = O.anchor_pyramid(cfg.strides, 1024, 2048, cfg.anchor_scales, cfg.anchor_ratios, cfg.anchor_base).astype(np.float32)
    off = g["offsets"]
    n = off[1:] - off[:-1]
    for i in sorted({int(n.argmax()), int(n.argmin()), 7, 33}):
        lab, lw, _, _ = O.anchor_encode(g["gt"][off[i]:off[i + 1]], anc, cfg)
        assert [(lab >= 1).sum(), (lab == 0).sum(), (lab < 0).sum()] == list(g["counts"][i])
        assert (lw[lab == 0] == g["bg_weight"][i]).all()
