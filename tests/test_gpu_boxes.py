"""GPU parity (through the C ABI) for anchors, IoU matrices, encode/decode and the fused
anchor-target assignment, against the CPU oracle and the committed golden vectors.

Bars: anchors, IoU (float64), labels, label weights, inside weights BIT-EXACT; float32 targets and
decoded boxes within 1e-5 relative (tests/util.py), the tolerance BASELINE.json's north_star states.
"""
import numpy as np
import pytest
import torch

import detbox
from oracle import oracle as O
from oracle.make_golden import synth_gt
from util import ATOL, RTOL, assert_boxes_close

pytestmark = pytest.mark.gpu
CPR = O.citypersons_retinanet_cfg()


def cu(a, dtype=None):
    t = torch.from_numpy(np.ascontiguousarray(a))
    if dtype is not None:
        t = t.to(dtype)
    return t.cuda()


@pytest.fixture(scope="module")
def cpr_anchors():
    ref = O.anchor_pyramid(CPR.strides, 1024, 2048, CPR.anchor_scales, CPR.anchor_ratios, 16)
    got = detbox.anchor_pyramid(CPR.strides, 1024, 2048, CPR.anchor_scales, CPR.anchor_ratios, 16)
    return ref, got


def test_anchor_pyramid_cpr(cpr_anchors, golden):
    ref, got = cpr_anchors
    assert got.shape == (261888, 4) and got.dtype == torch.float32
    np.testing.assert_array_equal(got.cpu().numpy(), ref.astype(np.float32))
    g = golden("anchors")
    np.testing.assert_array_equal(got.cpu().numpy()[g["cpr_pyramid_rows_idx"]], g["cpr_pyramid_rows"].astype(np.float32))
    got64 = detbox.anchor_pyramid(CPR.strides, 1024, 2048, CPR.anchor_scales, CPR.anchor_ratios, 16, dtype=torch.float64)
    np.testing.assert_array_equal(got64.cpu().numpy(), ref)


def test_anchor_pyramid_shifted(golden):
    g = golden("anchors")
    got = detbox.anchor_pyramid([4, 8, 16], 64, 96, [[2, 3], [4, 6], [8, 12]], [0.5, 1, 2.0], 16,
                                [[0.0, 0.0], [0.5, 0.25]], dtype=torch.float64)
    np.testing.assert_array_equal(got.cpu().numpy(), g["small_pyramid"])
    for lvl in range(5):
        np.testing.assert_array_equal(detbox.base_anchors(CPR.anchor_scales[lvl], CPR.anchor_ratios, 16), g["base_cpr_l%d" % lvl])
    np.testing.assert_array_equal(detbox.base_anchors([1.5, 3], [0.5, 1, 2.0, 3.0], 15), g["base_odd15"])


def test_iou_golden(golden):
    g = golden("iou")
    for b, q, want in ((g["fix_anchors"], g["fix_gt"][:, :4], g["fix_ov"]), (g["boxes"], g["query"], g["ov"]),
                       (g["boxes_int"], g["query_int"], g["ov_int"])):
        got = detbox.bbox_overlaps(cu(b, torch.float64), cu(q, torch.float64)).cpu().numpy()
        np.testing.assert_array_equal(got, want)
    got = detbox.bbox_intersections(cu(g["boxes"]).double(), cu(g["query"]).double()).cpu().numpy()
    np.testing.assert_array_equal(got, g["inter"])
    got = detbox.bbox_exclude_ignored_areas(cu(g["boxes"]).double(), cu(g["query"]).double(), 0.2).cpu().numpy()
    np.testing.assert_array_equal(got, g["excl"])


@pytest.mark.parametrize("n,k", [(1, 1), (257, 65), (5000, 130), (3, 0)])
def test_iou_random(n, k):
    rng = np.random.RandomState(n + k)
    b = rng.uniform(0, 300, (n, 4)); b[:, 2:] = b[:, :2] + rng.uniform(0, 90, (n, 2))
    q = rng.uniform(0, 300, (k, 4)); q[:, 2:] = q[:, :2] + rng.uniform(0, 90, (k, 2))
    b, q = b.astype(np.float32).astype(np.float64), q.astype(np.float32).astype(np.float64)
    np.testing.assert_array_equal(detbox.bbox_overlaps(cu(b), cu(q)).cpu().numpy(), O.bbox_overlaps(b, q))
    np.testing.assert_array_equal(detbox.bbox_intersections(cu(b), cu(q)).cpu().numpy(), O.bbox_intersections(b, q))


def test_transform_golden(golden):
    g = golden("transform")
    a, gt, d = cu(g["anchors"]), cu(g["gt"]), cu(g["deltas"])
    np.testing.assert_allclose(detbox.box_encode(a, gt).cpu().numpy(), g["enc"], rtol=RTOL, atol=ATOL)
    np.testing.assert_allclose(detbox.box_encode(a, gt, "linear").cpu().numpy(), g["enc_linear"], rtol=RTOL, atol=ATOL)
    assert_boxes_close(detbox.box_decode(a, d).cpu().numpy(), g["dec"])
    assert_boxes_close(detbox.box_decode(a, d, im_shape=(1024, 2048)).cpu().numpy(), g["dec_clip"])
    assert_boxes_close(detbox.box_decode(a, d, "linear").cpu().numpy(), g["dec_linear"])
    clipped = detbox.clip_boxes_(detbox.box_decode(a, d), (1024, 2048)).cpu().numpy()
    assert_boxes_close(clipped, g["dec_clip"])


def test_encode_divisions_are_ieee_exact(cpr_anchors):
    """box_math.cuh replaces the 8 IEEE divisions of the encode by FMA-corrected reciprocals
    (tests/tools/div_exact_check.c is the exhaustive host-side proof for the constants).  dx and dy
    involve no transcendental, so they must equal the oracle's plain-division result BIT FOR BIT; dw
    and dh differ only through logf (<= 1 ulp either side)."""
    _, anc = cpr_anchors
    a = anc.cpu().numpy()
    rng = np.random.RandomState(17)
    gt = synth_gt(rng, a.shape[0], 1024, 2048, 0.0)[:, :4]
    gt[::7] = a[::7]                                   # identical boxes: zero numerators
    gt[1::11, :2] += rng.uniform(-1e-3, 1e-3, gt[1::11, :2].shape).astype(np.float32)   # tiny numerators
    got = detbox.box_encode(anc, cu(gt)).cpu().numpy()
    want = O.bbox_transform(a, gt)
    assert np.array_equal(got[:, :2].view(np.uint32), want[:, :2].view(np.uint32))
    np.testing.assert_allclose(got[:, 2:], want[:, 2:], rtol=RTOL, atol=ATOL)


def test_transform_full_size_roundtrip(cpr_anchors):
    """Full CP-R size: decode(encode(gt)) == gt (size-independent property) + oracle on a sample."""
    _, anc = cpr_anchors
    rng = np.random.RandomState(5)
    gt = synth_gt(rng, anc.shape[0], 1024, 2048, 0.0)[:, :4]
    enc = detbox.box_encode(anc, cu(gt))
    dec = detbox.box_decode(anc, enc).cpu().numpy()
    w = np.maximum(gt[:, 2] - gt[:, 0], gt[:, 3] - gt[:, 1]) + 1
    assert_boxes_close(dec, gt, rtol=2e-5, extra_scale=np.maximum(w, anc.cpu().numpy().__abs__().max(1)))
    idx = rng.choice(anc.shape[0], 4000, replace=False)
    np.testing.assert_allclose(enc.cpu().numpy()[idx], O.bbox_transform(anc.cpu().numpy()[idx], gt[idx]), rtol=RTOL, atol=ATOL)


def run_target(anchors_np, gts, cfg, pyramid=None, format="reference"):
    off = np.zeros(len(gts) + 1, np.int32)
    off[1:] = np.cumsum([g.shape[0] for g in gts])
    gt_all = np.concatenate(gts, 0).astype(np.float32) if off[-1] else np.zeros((0, 5), np.float32)
    tc = detbox.make_target_cfg(cfg.rpn_bg_threshold, cfg.rpn_fg_threshold, cfg.rpn_clobber_positives,
                                cfg.ignored_area_intersection_fraction, cfg.bbweights, cfg.rpn_box_encoding)
    if format == "compact":
        out = detbox.anchor_target(cu(anchors_np), cu(gt_all), cu(off), tc, pyramid=pyramid, format="compact")
    else:
        out = detbox.anchor_target(cu(anchors_np), cu(gt_all), cu(off), tc, return_counts=True, pyramid=pyramid)
    return [o.cpu().numpy() for o in out]


def check_target(got, b, gt, anchors_np, cfg):
    lab, lw, bt, bw, mo, asg, isum = O.anchor_encode(gt, anchors_np, cfg, diagnostics=True)
    bad = np.nonzero(got[0][b] != lab)[0]
    if bad.size:   # north_star: report IoUs within 1e-6 of a threshold instead of failing on them
        thr = np.asarray([cfg.rpn_bg_threshold, cfg.rpn_fg_threshold, 0.5])
        near = (np.abs(mo[bad, None] - thr[None]).min(1) <= 1e-6) | (np.abs(isum[bad] - cfg.ignored_area_intersection_fraction) <= 1e-6)
        assert near.all(), "label mismatch away from any threshold at anchors %s" % bad[~near][:10]
        print("near-threshold label differences:", bad.tolist())
    np.testing.assert_array_equal(got[0][b], lab)
    np.testing.assert_array_equal(got[1][b], lw)
    np.testing.assert_array_equal(got[3][b], bw)
    np.testing.assert_allclose(got[2][b], bt, rtol=RTOL, atol=ATOL)
    np.testing.assert_array_equal(got[4][b], [(lab >= 1).sum(), (lab == 0).sum()])


CASES = ["fixture", "small_a", "small_noclobber", "small_zero_col", "small_zero_col_noclobber", "small_nogt", "small_dup_gt"]


@pytest.mark.parametrize("name", CASES)
def test_anchor_target_golden(golden, name):
    g = golden("anchor_target")
    bg, fg, clob = g[name + "_cfg"]
    cfg = O.Cfg(rpn_bg_threshold=float(bg), rpn_fg_threshold=float(fg), rpn_clobber_positives=bool(clob))
    got = run_target(g[name + "_anchors"], [g[name + "_gt"]], cfg)
    np.testing.assert_array_equal(got[0][0], g[name + "_labels"])
    np.testing.assert_array_equal(got[1][0], g[name + "_label_weights"])
    np.testing.assert_array_equal(got[3][0], g[name + "_bbox_inside_weights"])
    np.testing.assert_allclose(got[2][0], g[name + "_bbox_targets"], rtol=RTOL, atol=ATOL)


def test_anchor_target_cpr_batch(cpr_anchors, golden):
    """Full CP-R size, a ragged batch: G in {12 (golden image), 0, 1, 32, 99, 150 (> one GT chunk)}."""
    ref, _ = cpr_anchors
    anc = ref.astype(np.float32)
    g = golden("anchor_target")
    gts = [g["cpr_gt"], np.zeros((0, 5), np.float32)]
    for i, G in enumerate([1, 32, 99, 150]):
        gt = synth_gt(np.random.RandomState(40 + i), G, 1024, 2048, 0.15)
        if i % 2:
            gt[:, :4] = np.round(gt[:, :4])
        gts.append(gt)
    got = run_target(anc, gts, CPR)
    want = np.zeros(anc.shape[0], np.float32)
    want[g["cpr_labels_nz_idx"]] = g["cpr_labels_nz_val"]
    np.testing.assert_array_equal(got[0][0], want)           # golden labels of the reference itself
    for b, gt in enumerate(gts):
        check_target(got, b, gt, anc, CPR)


def test_anchor_target_real_citypersons_gt(cpr_anchors, golden):
    """six real CityPersons validation GT sets as ONE batch against the CP-R pyramid; labels, label
    weights and the rows carrying inside weights bit-exact against the compiled reference's recorded outputs, the
    sampled target rows within the float tolerance."""
    ref, _ = cpr_anchors
    anc = ref.astype(np.float32)
    g = golden("anchor_target_real")
    ids = [int(i) for i in g["image_ids"]]
    got = run_target(anc, [g["img%d_gt" % i] for i in ids], CPR)
    for b, i in enumerate(ids):
        k = "img%d_" % i
        want = np.zeros(anc.shape[0], np.float32)
        want[g[k + "labels_nz_idx"]] = g[k + "labels_nz_val"]
        np.testing.assert_array_equal(got[0][b], want)
        np.testing.assert_array_equal(got[4][b], g[k + "counts"][:2])
        np.testing.assert_array_equal(got[1][b][want == 0][:1], g[k + "bg_weight"])
        assert (got[1][b][want >= 1] == 1).all() and (got[1][b][want < 0] == 0).all()
        np.testing.assert_array_equal(np.nonzero(got[3][b].any(1))[0], g[k + "inside_rows"])
        np.testing.assert_allclose(got[2][b][g[k + "rows_idx"]], g[k + "bbox_targets_rows"], rtol=RTOL, atol=ATOL)


@pytest.mark.parametrize("seed", range(6))
def test_anchor_target_random_small(seed):
    rng = np.random.RandomState(100 + seed)
    scales = [[2, 2.52, 3.175], [4, 5.04, 6.349], [8, 10.079, 12.7]]
    anc = O.anchor_pyramid([8, 16, 32], 192, 320, scales, [1, 2.5], 16).astype(np.float32)
    bg, fg = [(0.5, 0.6), (0.4, 0.6), (0.3, 0.7)][seed % 3]
    cfg = O.Cfg(rpn_bg_threshold=bg, rpn_fg_threshold=fg, rpn_clobber_positives=seed != 4,
                rpn_box_encoding="linear" if seed == 5 else "fastrcnn", bbweights=[1., 2., 0.5, 1.5] if seed == 3 else [1.] * 4)
    gts = []
    for G in [1, 3, 12, 33, 60, 99, 0, 200]:
        gt = synth_gt(rng, G, 192, 320, 0.2, ncls=3)
        if (G + seed) % 2:
            gt[:, :4] = np.round(gt[:, :4])
        gts.append(gt)
    gts.append(np.asarray([[10, 10, 60, 90, 1], [400, 300, 420, 330, 2], [30, 20, 80, 100, -1]], np.float32))  # zero column
    got = run_target(anc, gts, cfg)
    for b, gt in enumerate(gts):
        check_target(got, b, gt, anc, cfg)


def test_anchor_target_large_coordinates():
    """Coordinates beyond 2^15 switch the column-max pass's float32 screen off (its error bound no longer holds);
    sub-pixel intersections and far-apart boxes take its `unsure` route.  Labels stay bit-exact either way."""
    rng = np.random.RandomState(321)
    scales = [[2, 2.52, 3.175], [4, 5.04, 6.349], [8, 10.079, 12.7]]
    base = O.anchor_pyramid([8, 16, 32], 192, 320, scales, [1, 2.5], 16).astype(np.float32)
    cfg = O.Cfg(rpn_bg_threshold=0.4, rpn_fg_threshold=0.6)
    for shift, mul in ((0.0, 1.0), (40000.0, 1.0), (0.0, 300.0)):
        anc = (base * np.float32(mul) + np.float32(shift)).astype(np.float32)
        gts = []
        for G in [2, 12, 40]:
            gt = synth_gt(rng, G, 192, 320, 0.2, ncls=3)
            gt[:, :4] = gt[:, :4] * np.float32(mul) + np.float32(shift)
            gts.append(gt.astype(np.float32))
        # a GT that only grazes anchors: intersections a fraction of a pixel wide
        g = anc[rng.choice(anc.shape[0], 6, replace=False)].copy()
        w = g[:, 2] - g[:, 0]
        g[:, 0] = g[:, 2] + np.float32(1.0) - np.float32(3e-4) * np.maximum(np.float32(1.0), np.abs(g[:, 2]) * np.float32(1e-3))
        g[:, 2] = g[:, 0] + w
        gts.append(np.concatenate([g, np.ones((6, 1), np.float32)], 1))
        got = run_target(anc, gts, cfg)
        for b, gt in enumerate(gts):
            check_target(got, b, gt, anc, cfg)


def test_rcp_inrange_exhaustive():
    """box_math.cuh::rcp_rn_inrange == the IEEE round-to-nearest reciprocal for EVERY float32 in [2^-20, 2^20]
    (3.4e8 values, the range the encode's single test admits)."""
    import torch
    lo = int(np.float32(2.0 ** -20).view(np.uint32))
    hi = int(np.float32(2.0 ** 20).view(np.uint32))
    bad = torch.zeros(1, dtype=torch.int64, device="cuda")
    rc = detbox.lib().detbox_selftest_rcp(lo, hi, bad.data_ptr(), torch.cuda.current_stream().cuda_stream)
    assert rc == 0
    assert int(bad.item()) == 0


def test_anchor_target_encode_paths_bit_exact():
    """The assign kernel's encode takes FMA-corrected reciprocals when every numerator and divisor is inside
    [2^-20, 2^20] and the plain IEEE divisions otherwise (box_math.cuh::encode_fastrcnn_pre).  Either way dx, dy must
    equal the oracle's divisions BIT FOR BIT; dw, dh differ only through logf.  GT rows here force both paths:
    a GT equal to an anchor (zero numerators), a huge GT, a width below 2^-20, an inverted GT (log of a negative)."""
    scales = [[2, 2.52, 3.175], [4, 5.04, 6.349], [8, 10.079, 12.7]]
    anc = O.anchor_pyramid([8, 16, 32], 192, 320, scales, [1, 2.5], 16).astype(np.float32)
    tiny = [0.5, 5.0, np.float32(-0.5 + 1.2e-7), 40.0, 1]            # gw = x2 - x1 + 1 ~ 1.2e-7 < 2^-20
    special = [
        np.r_[anc[1234], 1], np.r_[anc[17], 2],                     # zero numerators / log(1) = +0
        [-3e6, -2e6, 4e6, 3e6, 1],                                   # gw, gh > 2^20
        tiny,
        [90, 80, 40, 30, 2],                                         # inverted: gw, gh < 0 -> NaN targets, as NumPy
        [33.25, 41.5, 97.75, 160.125, 3],
        [7, 5, 6, 40, 1],                                            # gw = 0: log(0) = -inf
    ]
    gts = [np.asarray(special, np.float32)]                          # (the huge GT overlaps every anchor)
    for first in range(len(special)):                                # each special row as GT 0: the target of every
        rows = [special[first], special[5], special[0]]              # anchor that overlaps nothing (argmax of a zero row)
        gts.append(np.asarray(rows, np.float32))
    got = run_target(anc, gts, CPR)
    for b, gt in enumerate(gts):
        with np.errstate(all="ignore"):
            lab, lw, bt, bw = O.anchor_encode(gt, anc, CPR)[:4]
        np.testing.assert_array_equal(got[0][b], lab)
        g, w = got[2][b], bt
        nan_g, nan_w = np.isnan(g), np.isnan(w)
        np.testing.assert_array_equal(nan_g, nan_w)
        ok = ~nan_w[:, :2]
        assert np.array_equal(g[:, :2].view(np.uint32)[ok], w[:, :2].view(np.uint32)[ok])
        np.testing.assert_allclose(g[:, 2:], w[:, 2:], rtol=RTOL, atol=ATOL, equal_nan=True)


def test_anchor_target_host_entry(cpr_anchors, golden):
    """detbox_anchor_target_host: host arrays in/out through the C ABI (the end-to-end boundary)."""
    import ctypes
    ref, _ = cpr_anchors
    anc = ref.astype(np.float32)
    gt = golden("anchor_target")["cpr_gt"]
    A = anc.shape[0]
    lab, lw = np.empty(A, np.float32), np.empty(A, np.float32)
    bt, bw = np.empty((A, 4), np.float32), np.empty((A, 4), np.float32)
    tc = detbox.make_target_cfg(CPR.rpn_bg_threshold, CPR.rpn_fg_threshold)
    p = lambda a: a.ctypes.data_as(ctypes.c_void_p)
    rc = detbox.lib().detbox_anchor_target_host(p(anc), A, p(gt), gt.shape[0], ctypes.byref(tc), p(lab), p(lw), p(bt), p(bw), 0)
    assert rc == 0
    want = O.anchor_encode(gt, anc, CPR)
    np.testing.assert_array_equal(lab, want[0])
    np.testing.assert_array_equal(lw, want[1])
    np.testing.assert_allclose(bt, want[2], rtol=RTOL, atol=ATOL)
    np.testing.assert_array_equal(bw, want[3])


def test_no_cpu_fallback():
    with pytest.raises(detbox.DetboxError):
        detbox.box_decode(torch.zeros(4, 4), torch.zeros(4, 4))


# ---- round 2: separable GT-0 tables (pyramid descriptor) and the compact output format -------------------------------
def ragged_gts(seed, ih, iw, counts=(12, 0, 1, 32, 99, 150, 5)):
    gts = []
    for i, G in enumerate(counts):
        gt = synth_gt(np.random.RandomState(seed + i), G, ih, iw, 0.15) if G else np.zeros((0, 5), np.float32)
        if i % 2:
            gt[:, :4] = np.round(gt[:, :4])
        gts.append(gt)
    return gts


def test_anchor_target_tables_bit_identical_cpr(cpr_anchors, golden):
    """The table path (pyramid descriptor given) must reproduce the generic path BIT FOR BIT at CP-R size — all four
    arrays and the counts — for ragged G (0 ... 150), real CityPersons GT, and a GT row 0 placed so that one axis falls
    outside the fast path's range (entry marker -> full encode)."""
    ref, _ = cpr_anchors
    anc = ref.astype(np.float32)
    pyr = detbox.pyramid_desc(CPR.strides, 1024, 2048, CPR.anchor_scales, CPR.anchor_ratios, 16)
    g = golden("anchor_target_real")
    gts = ragged_gts(70, 1024, 2048) + [g["img%d_gt" % int(i)] for i in g["image_ids"][:3]]
    # GT 0 centred exactly on anchor centres of level 0 (nx == 0 for a column of anchors -> plain-division path there)
    a0 = anc[6 * (256 * 40 + 100)]
    cx, cy = (a0[0] + a0[2]) / 2, (a0[1] + a0[3]) / 2
    gts.append(np.asarray([[cx - 20, cy - 50, cx + 20, cy + 50, 1], [300, 300, 340, 400, 1]], np.float32))
    plain = run_target(anc, gts, CPR)
    tabled = run_target(anc, gts, CPR, pyramid=pyr)
    for a, b, name in zip(plain, tabled, ("labels", "label_weights", "bbox_targets", "bbox_inside_weights", "counts")):
        assert np.array_equal(a.view(np.uint32) if a.dtype == np.float32 else a, b.view(np.uint32) if b.dtype == np.float32 else b), name


@pytest.mark.parametrize("seed", range(4))
def test_anchor_target_tables_odd_pyramids(seed):
    """Grids that are not powers of two, several shifts per cell, tiles straddling levels, a partial last tile."""
    rng = np.random.RandomState(300 + seed)
    strides = [8, 16, 32][: 2 + seed % 2]
    ih, iw = [(200, 328), (136, 520), (96, 96), (264, 184)][seed]
    scales = [[2, 2.52, 3.175], [4, 5.04, 6.349], [8, 10.079, 12.7]][: len(strides)]
    shift = [[0.0, 0.0], [0.5, 0.5]] if seed % 2 else [[0.0, 0.0]]
    args = (strides, ih, iw, scales, [1, 2.5, 0.4][: 2 + seed // 2], 16, shift)
    anc_t = detbox.anchor_pyramid(*args)
    anc = anc_t.cpu().numpy()
    pyr = detbox.pyramid_desc(*args)
    cfg = O.Cfg(rpn_bg_threshold=0.4, rpn_fg_threshold=0.6, bbweights=[1., 2., 0.5, 1.5] if seed == 3 else [1.] * 4)
    gts = [synth_gt(rng, G, ih, iw, 0.2, ncls=3) if G else np.zeros((0, 5), np.float32) for G in (3, 0, 40, 1, 140)]
    plain = run_target(anc, gts, cfg)
    tabled = run_target(anc, gts, cfg, pyramid=pyr)
    for a, b in zip(plain, tabled):
        assert np.array_equal(a.view(np.uint32) if a.dtype == np.float32 else a, b.view(np.uint32) if b.dtype == np.float32 else b)
    for b, gt in enumerate(gts):
        check_target(tabled, b, gt, anc, cfg)


def test_anchor_target_tables_reject_non_separable_anchors():
    """A descriptor that does not describe the anchors (rows perturbed / permuted) must not change any result: the
    device-side separability check fails and the generic path runs."""
    args = ([8, 16], 128, 256, [[2, 3], [4, 6]], [1, 2.5], 16)
    anc = detbox.anchor_pyramid(*args).cpu().numpy()
    pyr = detbox.pyramid_desc(*args)
    rng = np.random.RandomState(5)
    gts = [synth_gt(rng, 6, 128, 256, 0.2), synth_gt(rng, 2, 128, 256, 0.0)]
    for mode in ("jitter_one", "permute"):
        a2 = anc.copy()
        if mode == "jitter_one":
            a2[1234, 2] += 0.25                      # one x2 off its column's value
        else:
            a2 = a2[rng.permutation(a2.shape[0])]
        plain = run_target(a2, gts, CPR)
        tabled = run_target(a2, gts, CPR, pyramid=pyr)
        for a, b in zip(plain, tabled):
            assert np.array_equal(a.view(np.uint32) if a.dtype == np.float32 else a, b.view(np.uint32) if b.dtype == np.float32 else b), mode
    bad = detbox.make_pyramid_desc([4, 4], [16, 8], [32, 15])          # describes the wrong number of rows
    with pytest.raises(detbox.DetboxError):
        run_target(anc, gts, CPR, pyramid=bad)


@pytest.mark.parametrize("with_pyr", [False, True])
def test_anchor_target_compact_format(cpr_anchors, with_pyr):
    """Compact format: packed byte + counts expand (detbox_expand_targets_host) to the reference-format arrays bit for
    bit; bbox_targets are the same bits; a partial last tile and an unaligned batch stride take the byte-store path."""
    ref, _ = cpr_anchors
    for anc, ih, iw, pyr in ((ref.astype(np.float32), 1024, 2048,
                              detbox.pyramid_desc(CPR.strides, 1024, 2048, CPR.anchor_scales, CPR.anchor_ratios, 16)),
                             (detbox.anchor_pyramid([8, 16], 104, 200, [[2, 3, 4], [4, 6, 7]], [1, 2.5, 0.5], 16).cpu().numpy(), 104, 200,
                              detbox.pyramid_desc([8, 16], 104, 200, [[2, 3, 4], [4, 6, 7]], [1, 2.5, 0.5], 16))):
        gts = ragged_gts(90, ih, iw, counts=(12, 0, 3, 40))
        full = run_target(anc, gts, CPR, pyramid=pyr if with_pyr else None)
        packed, bt, counts = run_target(anc, gts, CPR, pyramid=pyr if with_pyr else None, format="compact")
        assert np.array_equal(bt.view(np.uint32), full[2].view(np.uint32))
        assert np.array_equal(counts, full[4])
        labels, lw, bw = detbox.expand_targets(packed, counts, CPR.bbweights)
        assert np.array_equal(labels.numpy(), full[0]) and np.array_equal(lw.numpy(), full[1]) and np.array_equal(bw.numpy(), full[3])


def test_anchor_target_compact_rejects_unpackable_class():
    anc = detbox.anchor_pyramid([8], 64, 64, [[2, 3]], [1], 16).cpu().numpy()
    gt = np.asarray([[10, 10, 40, 50, 300]], np.float32)          # class 300 does not fit the byte
    packed, bt, counts = run_target(anc, [gt], CPR, format="compact")
    assert ((packed & 0x7f) == 127).any()
    with pytest.raises(detbox.DetboxError):
        detbox.expand_targets(packed, counts)


def test_anchor_target_64_real_citypersons_images(cpr_anchors, golden):
    """64 real anno_val.mat GT sets (G = 0 ... 46, 475 ignore regions) as ONE batch against the CP-R
    pyramid — (#fg, #bg, #ignored) and the background label weight of every image bit-exact against what the COMPILED
    reference's anchor.encode returned (tests/golden/real_gt64.npz, oracle/make_real_gt64.py); generic path, table path
    and compact format agree."""
    ref, _ = cpr_anchors
    anc = ref.astype(np.float32)
    g = golden("real_gt64")
    off = g["offsets"]
    gts = [g["gt"][off[i]:off[i + 1]] for i in range(64)]
    got = run_target(anc, gts, CPR)
    lab = got[0]
    np.testing.assert_array_equal(np.stack([(lab >= 1).sum(1), (lab == 0).sum(1), (lab < 0).sum(1)], 1), g["counts"])
    np.testing.assert_array_equal(got[4], g["counts"][:, :2])
    for b in range(64):
        bg = got[1][b][lab[b] == 0]
        assert bg.size and (bg == g["bg_weight"][b]).all()
    pyr = detbox.pyramid_desc(CPR.strides, 1024, 2048, CPR.anchor_scales, CPR.anchor_ratios, 16)
    packed, bt, counts = run_target(anc, gts, CPR, pyramid=pyr, format="compact")
    assert np.array_equal(bt.view(np.uint32), got[2].view(np.uint32)) and np.array_equal(counts, got[4])
    labels, lw, bw = detbox.expand_targets(packed, counts, CPR.bbweights)
    assert np.array_equal(labels.numpy(), lab) and np.array_equal(lw.numpy(), got[1]) and np.array_equal(bw.numpy(), got[3])


@pytest.mark.parametrize("fill", [0, 30, 110])
@pytest.mark.parametrize("seed", range(3))
def test_anchor_target_decide_screen_adversarial(seed, fill):
    """The float32 screen of the decide phase must hand every undecidable case to the exact float64 path.  GT sets built
    FROM anchors so that the cases sit exactly on the screen's boundaries: a GT equal to an anchor (IoU 1 = its column
    maximum: promotion), the same GT twice (argmax tie: first maximum wins), GT that touch an anchor exactly (iw = 0),
    overlap it by exactly one pixel column (iw = 1), by 1e-4 (inside the undecided band) and by -1e-4, GT scaled so that
    the IoU with an anchor lands on the thresholds 0.5 / 0.6 up to rounding, ignore regions whose intersection fraction
    lands on 0.3.  Labels / weights bit-exact against the oracle, as everywhere.  fill > 0 appends random GT rows so that
    the image has 33 .. 128 GT and takes the several-mask-words route (decide_multi) with the same boundary cases."""
    rng = np.random.RandomState(900 + seed)
    scales = [[2, 2.52, 3.175], [4, 5.04, 6.349], [8, 10.079, 12.7]]
    anc = O.anchor_pyramid([8, 16, 32], 160, 256, scales, [1, 2.5], 16).astype(np.float32)
    cfg = O.Cfg(rpn_bg_threshold=0.5, rpn_fg_threshold=0.6)
    gts = []
    for _ in range(6):
        a = anc[rng.choice(anc.shape[0], 6, replace=False)].astype(np.float64)
        w, h = a[:, 2] - a[:, 0] + 1, a[:, 3] - a[:, 1] + 1
        rows = [a[0], a[0],                                                     # equal to an anchor, twice
                [a[1, 2] + 1, a[1, 1], a[1, 2] + 1 + w[1], a[1, 3]],              # touches: iw = 0
                [a[1, 2], a[1, 1], a[1, 2] + w[1], a[1, 3]],                      # iw = 1
                [a[2, 2] + 1 - 1e-4, a[2, 1], a[2, 2] + w[2], a[2, 3]],           # iw = 1e-4
                [a[2, 2] + 1 + 1e-4, a[2, 1], a[2, 2] + 2 * w[2], a[2, 3]]]       # iw = -1e-4
        for t, k in ((0.5, 3), (0.6, 4)):                                       # same corner, area ratio t: IoU = t up to rounding
            rows.append([a[k, 0], a[k, 1], a[k, 0] + w[k] * t - 1, a[k, 3]])
            rows.append([a[k, 0], a[k, 1], a[k, 2], a[k, 1] + h[k] / t - 1])
        gt = np.concatenate([np.asarray(rows, np.float64), np.ones((len(rows), 1))], 1)
        ign = [a[5, 0], a[5, 1], a[5, 0] + w[5] * 0.3 - 1, a[5, 3], -1]           # covers 0.3 of anchor 5
        gt = np.concatenate([gt, np.asarray([ign])], 0).astype(np.float32)
        if fill:
            gt = np.concatenate([gt, synth_gt(rng, fill, 160, 256, 0.2, ncls=2)], 0).astype(np.float32)
        gts.append(gt[rng.permutation(gt.shape[0])] if seed else gt)
    got = run_target(anc, gts, cfg)
    for b, gt in enumerate(gts):
        check_target(got, b, gt, anc, cfg)


def test_anchor_target_float32_anchor_contract(cpr_anchors, golden):
    """the reference has two anchor dtypes on its own paths (float64 pyramid per sample, data_layer.py:99-102;
    float32 per batch, :138-140).  This library implements the float32 one.  tests/golden/anchor_f64_delta.npz (from the
    COMPILED reference, oracle/make_golden_f64_anchors.py) holds both paths' labels on the 12 densest real CityPersons images:
    ours must equal the float32 path everywhere — including the 2 anchors (of 3.1 M) where the reference's own float64 path
    labels differently."""
    ref, _ = cpr_anchors
    anc = ref.astype(np.float32)
    g = golden("anchor_f64_delta")
    ids = [int(i) for i in g["image_ids"]]
    got = run_target(anc, [g["img%d_gt" % i] for i in ids], CPR)
    n_diff = 0
    for b, i in enumerate(ids):
        k = "img%d_" % i
        want = np.zeros(anc.shape[0], np.float32)
        want[g[k + "labels32_nz_idx"]] = g[k + "labels32_nz_val"]
        np.testing.assert_array_equal(got[0][b], want)
        d = g[k + "diff_idx"]
        if d.size:
            assert (got[0][b][d] != g[k + "labels64_at_diff"]).all()       # the documented delta to the float64-anchor path
        n_diff += d.size
    assert n_diff == int(g["total_label_diffs"]) == 2
