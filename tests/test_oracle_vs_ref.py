"""The CPU restatement against the live, UNMODIFIED reference (oracle/_ref, compiled from
/root/reference by oracle/build_ref.py) on fresh seeded inputs.  Skipped when oracle/_ref has
not been built.  CPU only."""
import numpy as np
import pytest

from oracle import oracle as O
from oracle import ref_loader
from oracle.make_golden import synth_gt, set_cfg
from util import assert_boxes_close

pytestmark = pytest.mark.skipif(not ref_loader.available(), reason="oracle/_ref not built")
RTOL, ATOL = 1e-5, 1e-6


@pytest.fixture(scope="module")
def ref():
    return ref_loader.load()


def rand_boxes(rng, n, size=300, maxwh=90):
    b = rng.uniform(0, size, (n, 4))
    b[:, 2:] = b[:, :2] + rng.uniform(0, maxwh, (n, 2))
    return b.astype(np.float32)


@pytest.mark.parametrize("seed", range(4))
def test_iou_random(ref, seed):
    rng = np.random.RandomState(seed)
    b, q = rand_boxes(rng, 700).astype(np.float64), rand_boxes(rng, 31).astype(np.float64)
    np.testing.assert_array_equal(O.bbox_overlaps(b, q), ref.cython_bbox.bbox_overlaps(b, q))
    np.testing.assert_array_equal(O.bbox_intersections(b, q), ref.cython_bbox.bbox_intersections(b, q))
    np.testing.assert_array_equal(O.bbox_exclude_ignored_areas(b, q, 0.3),
                                  ref.cython_bbox.bbox_exclude_ignored_areas(b, q, 0.3))


@pytest.mark.parametrize("args", [
    ([8, 16, 32, 64, 128], 896, 1408, [[2, 2.52, 3.17], [4, 5.04, 6.35], [8, 10.08, 12.7], [16, 20.16, 25.4], [32, 40.32, 50.8]], [0.5, 1, 2.0], 16, [[0.0, 0.0]]),
    ([4, 8, 16, 32], 256, 512, [[2, 4], [4, 8], [8, 16], [16, 32]], [0.5, 1, 2.0], 16, [[0, 0], [0.5, 0.5]]),
    ([8, 16], 100, 130, [2., 3.3, 7.1], [0.37, 1.9], 24, [[0.0, 0.0]]),
])
def test_anchor_pyramid_random(ref, args):
    np.testing.assert_array_equal(O.anchor_pyramid(*args), ref.boxes_anchor.anchor_pyramid(*args))


@pytest.mark.parametrize("seed", range(6))
def test_anchor_encode_random(ref, seed):
    rng = np.random.RandomState(100 + seed)
    scales = [[2, 2.52, 3.175], [4, 5.04, 6.349], [8, 10.079, 12.7]]
    anc = ref.boxes_anchor.anchor_pyramid([8, 16, 32], 192, 320, scales, [1, 2.5], 16).astype(np.float32)
    G = [1, 3, 12, 33, 60, 99][seed]
    gt = synth_gt(rng, G, 192, 320, 0.2, ncls=3)
    if seed % 2:
        gt[:, :4] = np.round(gt[:, :4])           # integer GT -> exact IoU ties
    bg, fg = [(0.5, 0.6), (0.4, 0.6), (0.3, 0.7)][seed % 3]
    clob = seed != 4
    set_cfg(ref.cfg, bg, fg, clob)
    want = ref.layers_anchor.encode(gt.copy(), anc.copy())
    got = O.anchor_encode(gt, anc, O.Cfg(rpn_bg_threshold=bg, rpn_fg_threshold=fg, rpn_clobber_positives=clob))
    np.testing.assert_array_equal(got[0], want[0])
    np.testing.assert_array_equal(got[1], want[1])
    np.testing.assert_allclose(got[2], want[2], rtol=RTOL, atol=ATOL)
    np.testing.assert_array_equal(got[3], want[3])


@pytest.mark.parametrize("seed", range(3))
def test_decode_random(ref, seed):
    rng = np.random.RandomState(200 + seed)
    anc = rand_boxes(rng, 2000, 1000, 300)
    d = rng.normal(0, 0.5, (2000, 4)).astype(np.float32)
    probs = rng.rand(2000, 2).astype(np.float32)
    ref.cfg.rpn_box_encoding = "fastrcnn"
    fb, cls, _ = ref.layers_anchor.decode(d.copy(), probs, anc, 600, 1000, 2)
    got, gcls, _ = O.anchor_decode(d, probs, anc, 600, 1000, 2)
    assert_boxes_close(got, fb)
    np.testing.assert_array_equal(gcls, cls)


@pytest.mark.parametrize("seed", range(3))
@pytest.mark.parametrize("thr", [0.5, 0.7])
def test_nms_random(ref, seed, thr):
    rng = np.random.RandomState(300 + seed)
    n = 1500
    b = rand_boxes(rng, n, 600, 120)
    s = rng.permutation(np.linspace(0, 1, n)).astype(np.float32)
    d = np.concatenate([b, s[:, None]], 1)
    np.testing.assert_array_equal(O.cpu_nms(d, thr), ref.cpu_nms.cpu_nms(d, thr))
    np.testing.assert_array_equal(O.gpu_nms(d, thr), ref.py_cpu_nms.py_cpu_nms(d, thr))


def test_crop_and_resize_random(ref):
    rng = np.random.RandomState(7)
    img = rng.normal(0, 1, (3, 8, 25, 31)).astype(np.float32)
    rois = np.concatenate([rng.randint(0, 3, (50, 1)).astype(np.float32),
                           rand_boxes(rng, 50, 200, 120)], 1)
    boxes, ind = O.tf_boxes(rois, 0.125, 7, 7, 25, 31)
    np.testing.assert_array_equal(O.crop_and_resize_fwd(img, boxes, ind, 0.0, 7, 7),
                                  ref.crop_and_resize_forward(img, boxes, ind, 0.0, 7, 7))
    g = rng.normal(0, 1, (50, 8, 7, 7)).astype(np.float32)
    np.testing.assert_array_equal(O.crop_and_resize_bwd(g, boxes, ind, img.shape),
                                  ref.crop_and_resize_backward(g, boxes, ind, img.shape))


def test_tf_boxes_matches_torch_wrapper_arithmetic():
    """roi_align_tf.py:24-40 re-evaluated with torch CPU float32 ops (the wrapper module itself
    cannot be imported: it needs the torch-0.4 ffi backend)."""
    import torch
    rng = np.random.RandomState(8)
    rois = np.concatenate([np.zeros((64, 1), np.float32), rand_boxes(rng, 64, 900, 300)], 1)
    ch = cw = 7
    H, W, s = 64, 128, 1.0 / 16
    t = torch.from_numpy(rois)
    bx = float(s) * t[:, 1:5].contiguous()
    x1, y1, x2, y2 = torch.split(bx, 1, dim=1)
    sw, sh = (x2 - x1) / float(cw), (y2 - y1) / float(ch)
    nx0 = (x1 + sw / 2 - 0.5) / float(W - 1)
    ny0 = (y1 + sh / 2 - 0.5) / float(H - 1)
    nw = sw * float(cw - 1) / float(W - 1)
    nh = sh * float(ch - 1) / float(H - 1)
    want = torch.cat((ny0, nx0, ny0 + nh, nx0 + nw), 1).numpy()
    got, _ = O.tf_boxes(rois, s, ch, cw, H, W)
    np.testing.assert_array_equal(got, want)


@pytest.mark.parametrize("seed", range(12))
def test_matching_box_and_sample_rois_random(ref, seed):
    """oracle.matching_box / sample_rois == the reference's libs/layers/box.py:12-69, 175-292 (its own code on
    CPU torch tensors; np.random seeded identically on both sides)."""
    torch = pytest.importorskip("torch")
    from oracle.make_golden import synth_rois
    if ref.layers_box is None:
        pytest.skip("reference box.py not importable")
    b, ids, gts = synth_rois(50 + seed, sorted_ids=seed % 2 == 0)
    want = ref.layers_box.matching_box(torch.from_numpy(b), torch.from_numpy(ids), gts, 0.4, 0.6)
    got = O.matching_box(b, ids, gts, 0.4, 0.6)
    for w, g in zip(want, got):
        np.testing.assert_array_equal(w.numpy(), g)
    np.random.seed(seed)
    want = ref.layers_box.sample_rois(torch.from_numpy(b), torch.from_numpy(ids), gts, 0.5, 16, 0.25, 0.2)
    got = O.sample_rois(b, ids, gts, 0.5, 16, 0.25, 0.2, rng=np.random.RandomState(seed))
    for w, g in zip(want, got):
        np.testing.assert_array_equal(w.numpy(), g)


@pytest.mark.parametrize("seed", range(4, 8))
def test_nms_variants_random(ref, seed):
    """oracle.nms_new / nms_multianchor == the reference's compiled libs/boxes/nms.pyx:70-195."""
    from oracle.make_golden import variant_dets
    d = variant_dets(seed)
    for thr in (0.3, 0.6):
        assert [int(i) for i in ref.cython_nms.nms_new(d, thr)] == O.nms_new(d, thr)
        keep, clusters = ref.cython_nms.nms_multianchor(d, thr)
        k2, c2 = O.nms_multianchor(d, thr)
        assert [int(i) for i in keep] == k2
        np.testing.assert_array_equal(np.asarray(clusters), c2)
