"""GPU parity for the post-processing half of the path: stable sort, objectness top-k + decode,
NMS (device and host-buffer entry points), against the CPU oracle and the golden keep lists that
came from the reference's own cpu_nms / py_cpu_nms.  Index results are BIT-EXACT."""
import ctypes

import numpy as np
import pytest
import torch

import detbox
from oracle import oracle as O
from util import assert_boxes_close

pytestmark = pytest.mark.gpu


def cu(a, dtype=None):
    t = torch.from_numpy(np.ascontiguousarray(a))
    if dtype is not None:
        t = t.to(dtype)
    return t.cuda()


def keep_list(keep, n_keep):
    nk = n_keep.cpu().numpy()
    return keep.cpu().numpy()[:nk[-1]], nk


# ------------------------------------------------------------------------------------- sort ----
@pytest.mark.parametrize("n", [1, 31, 2048, 2049, 70001, 1500000])
def test_sort_desc(n):
    rng = np.random.RandomState(n)
    s = rng.randn(n).astype(np.float32)
    if n > 100:
        s[rng.randint(0, n, n // 3)] = s[rng.randint(0, n, n // 3)]       # ties
        s[:7] = [0.0, -0.0, np.inf, -np.inf, 1e-40, -1e-40, 3.0]
    got = detbox.sort_desc(cu(s)).cpu().numpy()
    srt = s[got]
    assert (srt[:-1] >= srt[1:]).all()                                      # sortedness
    assert np.array_equal(np.sort(got), np.arange(n))                       # a permutation
    eq = srt[:-1] == srt[1:]
    signs_equal = np.signbit(srt[:-1]) == np.signbit(srt[1:])
    assert (got[:-1][eq & signs_equal] < got[1:][eq & signs_equal]).all()   # stable: ties by index
    if n <= 70001:
        s2 = s.copy(); s2[s2 == 0] = 0.0
        want = O.stable_desc_order(np.where(np.signbit(s) & (s == 0), -1e-45, s).astype(np.float64))
        # (-0.0 sorts after +0.0 in the key order: the documented total order on bit patterns)
        np.testing.assert_array_equal(got, want)


# ------------------------------------------------------------------------------------ top-k ----
def make_head_outputs(rng, B, A, C, activation, ties=False):
    deltas = rng.normal(0, 0.5, (B * A, 4)).astype(np.float32)
    logits = rng.normal(-4.6, 1.0, (B * A, C)).astype(np.float32)
    hot = rng.rand(B * A) < 0.01
    logits[hot, (1 if activation == "softmax" else 0)] += 6.0
    if activation == "softmax":
        e = np.exp(logits - logits.max(1, keepdims=True)); p = (e / e.sum(1, keepdims=True)).astype(np.float32)
    else:
        p = (1.0 / (1.0 + np.exp(-logits))).astype(np.float32)
    if ties:
        p = np.round(p * 64).astype(np.float32) / 64       # heavy ties: the tie rule decides
    return deltas, p


@pytest.mark.parametrize("B,A,C,act,top_n,per_image,ties", [
    (1, 5000, 2, "softmax", 900, False, False),
    (3, 4097, 2, "softmax", 1000, False, True),
    (4, 3000, 2, "softmax", 300, True, True),
    (2, 6000, 80, "sigmoid", 900, False, False),
    (5, 2500, 80, "sigmoid", 400, True, True),
    (2, 3000, 3, "sigmoid", 7000, False, False),      # top_n > B*A: everything, sorted
    (2, 300, 81, "softmax", 50, True, False),
    (1, 261888, 2, "softmax", 900, False, False),     # CP-R size, eval setting (test.py:148)
    (300, 64, 2, "softmax", 10, True, True),          # > 256 images: two segment passes
    (3, 9000, 2, "softmax", 3000, True, True),        # per-image K > 2048: global radix sort + image pass
    (2, 20000, 2, "softmax", 6000, True, False),      # MaskRCNN test-time budget per image (maskrcnn.py:97)
    (16, 132450, 80, "sigmoid", 14400, False, False), # BASELINE configs[3]: COCO-R batch 16, global 900*B
    (16, 132450, 80, "sigmoid", 1000, True, False),   # ... and 1000 per image
    (5, 2501, 2, "softmax", 300, True, True),         # segment starts at every alignment mod 4 (vectorised partition: head / body / tail)
    (7, 3, 2, "softmax", 2, True, True),              # segments shorter than one vector
    (3, 4099, 80, "sigmoid", 500, True, False),       # 4099 % 4 == 3
    (2, 40000, 2, "softmax", 20000, False, True),     # K > 16384: three-pass radix select + global radix sort (the old K > 2048 route)
    (2, 30000, 2, "softmax", 16384, True, False),     # exactly the one-block sort's capacity, per image
])
def test_topk_decode(B, A, C, act, top_n, per_image, ties):
    rng = np.random.RandomState(B * 7 + A)
    anc = rng.uniform(0, 500, (A, 4)).astype(np.float32); anc[:, 2:] += anc[:, :2]
    deltas, p = make_head_outputs(rng, B, A, C, act, ties)
    got = detbox.topk_decode(cu(deltas), cu(p), cu(anc), top_n, act, "fastrcnn", per_image,
                             want=("probs", "img_ids", "anchors", "flat_idx", "scores"))
    wb, wp, wi, wa, widx = O.decode_and_choose_top_n(deltas, p, anc, top_n, act, "fastrcnn", per_image)
    np.testing.assert_array_equal(got["flat_idx"].cpu().numpy(), widx)       # bit-exact selection + order
    np.testing.assert_array_equal(got["img_ids"].cpu().numpy(), wi)
    np.testing.assert_array_equal(got["probs"].cpu().numpy(), wp)
    np.testing.assert_array_equal(got["anchors"].cpu().numpy(), wa)
    np.testing.assert_array_equal(got["scores"].cpu().numpy(), O.objectness(p, act)[widx])
    assert_boxes_close(got["boxes"].cpu().numpy(), wb, extra_scale=np.abs(wa).max(1))


def test_topk_full_batch_property():
    """BASELINE size (64 CP-R images, 16.7M anchors): size-independent checks — the selected
    scores are sorted, every unselected score is <= the k-th, indices are unique."""
    B, A, k = 64, 261888, 900
    g = torch.Generator(device="cuda"); g.manual_seed(0)
    p0 = torch.rand((B * A,), device="cuda", generator=g)
    prob = torch.stack([p0, 1 - p0], 1).contiguous()
    deltas = torch.zeros((B * A, 4), device="cuda")
    anc = torch.rand((A, 4), device="cuda") * 100
    for per_image in (False, True):
        out = detbox.topk_decode(deltas, prob, anc, k, "softmax", "fastrcnn", per_image, want=("flat_idx", "scores", "img_ids"))
        idx, sc = out["flat_idx"], out["scores"]
        obj = 1.0 - prob[:, 0]
        assert torch.equal(sc, obj[idx])
        assert idx.unique().numel() == idx.numel()
        if per_image:
            assert idx.numel() == B * k
            assert torch.equal(out["img_ids"], torch.arange(B, device="cuda").repeat_interleave(k))
            s2 = sc.view(B, k)
            assert (s2[:, :-1] >= s2[:, 1:]).all()
            kth = s2[:, -1]
            mask = torch.ones(B * A, dtype=torch.bool, device="cuda"); mask[idx] = False
            assert (obj.view(B, A) <= kth[:, None])[mask.view(B, A)].all()
        else:
            assert (sc[:-1] >= sc[1:]).all()
            assert torch.equal(torch.sort(sc, descending=True)[0], torch.topk(obj, k)[0])


# -------------------------------------------------------------------------------------- NMS ----
MODES = [("ge", O.NMS_GE_F64, detbox.NMS_GE_F64), ("gt", O.NMS_GT_F32, detbox.NMS_GT_F32)]


@pytest.mark.parametrize("name", ["rand1k", "rand3k", "clustered2k", "intgrid600"])
@pytest.mark.parametrize("thr", [0.5, 0.3, 0.7])
def test_nms_golden(golden, name, thr):
    g = golden("nms")
    d = g[name]
    for tag, _, mode in MODES:
        keep, nk = keep_list(*detbox.nms(cu(d[:, :4]), cu(d[:, 4]), thr, mode))
        np.testing.assert_array_equal(keep, g["%s_keep_%s_%g" % (name, tag, thr)])
    keep, _ = keep_list(*detbox.nms(cu(d), cu(d[:, 4]), thr, detbox.NMS_GE_F32))     # stride-5 rows
    np.testing.assert_array_equal(keep, O.nms(d[:, :4], d[:, 4], thr, O.NMS_GE_F32))


def sweep_boxes(rng, n):
    h = np.exp(rng.uniform(np.log(12), np.log(264), n)); w = h / 2.44
    cx, cy = rng.uniform(0, 2048, n), rng.uniform(0, 1024, n)
    b = np.stack([cx - w / 2, cy - h / 2, cx + w / 2, cy + h / 2], 1)
    b[:, 0::2] = np.clip(b[:, 0::2], 0, 2047); b[:, 1::2] = np.clip(b[:, 1::2], 0, 1023)
    return b.astype(np.float32), rng.permutation(np.linspace(0, 1, n)).astype(np.float32)


@pytest.mark.parametrize("n", [1, 2, 63, 64, 65, 1000, 5000, 20000])
def test_nms_sweep_vs_oracle(n):
    b, s = sweep_boxes(np.random.RandomState(2), n)
    for _, omode, mode in MODES:
        keep, nk = keep_list(*detbox.nms(cu(b), cu(s), 0.5, mode))
        np.testing.assert_array_equal(keep, O.nms(b, s, 0.5, omode))
        assert nk[0] == nk[1] == keep.size


def test_nms_100k_properties():
    """100k boxes (BASELINE sweep top end; 1.25 GB mask): greedy-NMS invariants that do not need the
    O(n^2) CPU oracle — kept boxes are mutually below threshold, every dropped box is suppressed by
    a higher-scored kept box, keep order is score-descending; plus idempotence."""
    n = 100000
    b, s = sweep_boxes(np.random.RandomState(3), n)
    keep, nk = keep_list(*detbox.nms(cu(b), cu(s), 0.5, detbox.NMS_GT_F32))
    assert (s[keep][:-1] > s[keep][1:]).all()
    kb = cu(b[keep]).double()
    iou = detbox.bbox_overlaps(kb, kb)
    iou.fill_diagonal_(0)
    assert float(iou.max()) <= 0.5 + 1e-6
    dropped = np.setdiff1d(np.arange(n), keep)
    sub = dropped[np.random.RandomState(0).choice(dropped.size, 3000, replace=False)]
    ov = detbox.bbox_overlaps(cu(b[sub]).double(), kb).cpu().numpy()
    higher = s[keep][None, :] > s[sub][:, None]
    assert ((ov > 0.5 - 1e-6) & higher).any(1).all()
    keep2, _ = keep_list(*detbox.nms(cu(b[keep]), cu(s[keep]), 0.5, detbox.NMS_GT_F32))
    np.testing.assert_array_equal(keep2, np.arange(keep.size))            # idempotent


def test_nms_100k_exact_vs_oracle():
    """BASELINE configs[1] top end: keep indices at n = 100,000 bit-exact against the C oracle (pinned to the compiled
    cpu_nms / py_cpu_nms), both comparators: `>=` in float64 (libs/nms/cpu_nms.pyx:17-68) and `>` in float32
    (libs/nms/nms_kernel.cu:46-90)."""
    n = 100000
    b, s = sweep_boxes(np.random.RandomState(3), n)
    db, ds = cu(b), cu(s)
    for _, omode, mode in MODES:
        keep, nk = keep_list(*detbox.nms(db, ds, 0.5, mode))
        want = O.nms(b, s, 0.5, omode)
        assert keep.size == want.size and nk[0] == want.size
        np.testing.assert_array_equal(keep, want)


def test_nms_segments_match_per_image_loop():
    """_apply_nms_in_batch (model.py:343-355): segments == the reference's per-image Python loop."""
    rng = np.random.RandomState(9)
    n, B = 6000, 7
    b, s = sweep_boxes(rng, n)
    b = (b / 4).astype(np.float32)                      # denser -> real suppression
    img = rng.randint(0, B, n).astype(np.int32)
    img[img == 3] = 4                                   # an empty segment
    keep, nk = keep_list(*detbox.nms(cu(b), cu(s), 0.5, detbox.NMS_GT_F32, seg_ids=cu(img), n_seg=B))
    want, counts = [], []
    for i in range(B):
        inds = np.nonzero(img == i)[0]
        k = O.nms(b[inds], s[inds], 0.5, O.NMS_GT_F32) if inds.size else np.zeros(0, np.int64)
        want.append(inds[k]); counts.append(k.size)
    np.testing.assert_array_equal(keep, np.concatenate(want))
    np.testing.assert_array_equal(nk[:B], counts)
    assert nk[B] == sum(counts)
    # probs-driven variant through the oracle's composition
    probs = np.stack([1 - s, s], 1).astype(np.float32)
    np.testing.assert_array_equal(keep, O.apply_nms_in_batch(b, probs, img, "softmax", 0.5, O.NMS_GT_F32))


def test_nms_many_segments():
    rng = np.random.RandomState(11)
    n, B = 20000, 300
    b, s = sweep_boxes(rng, n)
    b = (b / 8).astype(np.float32)
    img = rng.randint(0, B, n).astype(np.int32)
    keep, nk = keep_list(*detbox.nms(cu(b), cu(s), 0.5, detbox.NMS_GE_F64, seg_ids=cu(img), n_seg=B))
    want = []
    for i in range(B):
        inds = np.nonzero(img == i)[0]
        want.append(inds[O.nms(b[inds], s[inds], 0.5, O.NMS_GE_F64)] if inds.size else np.zeros(0, np.int64))
    np.testing.assert_array_equal(keep, np.concatenate(want))


def test_nms_unbalanced_segments():
    """Segment sizes from 0 to 5000: the per-segment resolve with rows in shared memory (<= ~1200 boxes),
    rows read from global memory (larger), and empty segments, in one call."""
    rng = np.random.RandomState(21)
    sizes = [5000, 0, 1000, 37, 1, 1900, 64, 0]
    n, B = sum(sizes), len(sizes)
    b, s = sweep_boxes(rng, n)
    b = (b / 4).astype(np.float32)
    img = rng.permutation(np.repeat(np.arange(B), sizes)).astype(np.int32)
    for _, omode, mode in MODES:
        keep, nk = keep_list(*detbox.nms(cu(b), cu(s), 0.4, mode, seg_ids=cu(img), n_seg=B))
        want = []
        for i in range(B):
            inds = np.nonzero(img == i)[0]
            want.append(inds[O.nms(b[inds], s[inds], 0.4, omode)] if inds.size else np.zeros(0, np.int64))
        np.testing.assert_array_equal(keep, np.concatenate(want))
        np.testing.assert_array_equal(nk[:B], [w.size for w in want])


@pytest.mark.parametrize("thr", [0.5, 1.0 / 3.0, 0.25, 0.2, 0.6, 1.0])
def test_nms_quotients_on_the_threshold(thr):
    """Small integer boxes: thousands of pairs whose IoU is exactly thr (or one ulp off after float32
    rounding) — the approximate-quotient shortcut must hand every one of them to the IEEE division."""
    rng = np.random.RandomState(4)
    n = 3000
    x1 = rng.randint(0, 24, n); y1 = rng.randint(0, 24, n)
    b = np.stack([x1, y1, x1 + rng.randint(0, 8, n), y1 + rng.randint(0, 8, n)], 1).astype(np.float32)
    s = rng.permutation(n).astype(np.float32)
    for omode, mode in ((O.NMS_GE_F64, detbox.NMS_GE_F64), (O.NMS_GE_F32, detbox.NMS_GE_F32), (O.NMS_GT_F32, detbox.NMS_GT_F32)):
        keep, _ = keep_list(*detbox.nms(cu(b), cu(s), thr, mode))
        np.testing.assert_array_equal(keep, O.nms(b, s, thr, omode))
        img = (np.arange(n) % 5).astype(np.int32)
        keep, _ = keep_list(*detbox.nms(cu(b), cu(s), thr, mode, seg_ids=cu(img), n_seg=5))
        want = [np.nonzero(img == i)[0] for i in range(5)]
        want = np.concatenate([w[O.nms(b[w], s[w], thr, omode)] for w in want])
        np.testing.assert_array_equal(keep, want)


def test_nms_edge_cases():
    keep, nk = detbox.nms(torch.zeros((0, 4), device="cuda"), torch.zeros((0,), device="cuda"), 0.5)
    assert int(nk[-1]) == 0
    # degenerate boxes (x2 < x1): area <= 0, possibly NaN IoU -> never suppressed
    b = np.asarray([[10, 10, 5, 5], [10, 10, 5, 5], [0, 0, -1, -1], [0, 0, -1, -1], [1, 1, 8, 8], [1, 1, 8, 8]], np.float32)
    s = np.asarray([.9, .8, .7, .6, .5, .4], np.float32)
    for _, omode, mode in MODES:
        keep, _ = keep_list(*detbox.nms(cu(b), cu(s), 0.5, mode))
        np.testing.assert_array_equal(keep, O.nms(b, s, 0.5, omode))
    # thr <= 0 with '>=': everything after the first is suppressed
    bb, ss = sweep_boxes(np.random.RandomState(1), 500)
    for thr in (0.0, -1.0):
        for _, omode, mode in MODES:
            keep, _ = keep_list(*detbox.nms(cu(bb), cu(ss), thr, mode))
            np.testing.assert_array_equal(keep, O.nms(bb, ss, thr, omode))


def test_reference_ffi_symbol_nms():
    """`_nms` with the exact signature of libs/nms/gpu_nms.hpp:1-2 (sorted host boxes in, indices
    into the sorted array out, '>' comparator) and detbox_nms_host (unsorted dets)."""
    b, s = sweep_boxes(np.random.RandomState(5), 3000)
    b = (b / 3).astype(np.float32)
    dets = np.concatenate([b, s[:, None]], 1).astype(np.float32)
    order = O.stable_desc_order(s)
    sorted_dets = np.ascontiguousarray(dets[order])
    keep = np.zeros(3000, np.int32); num = ctypes.c_int(0)
    p = lambda a: a.ctypes.data_as(ctypes.c_void_p)
    detbox.lib()._nms(p(keep), ctypes.cast(ctypes.byref(num), ctypes.c_void_p), p(sorted_dets), 3000, 5, ctypes.c_float(0.5), 0)
    want = O.nms(b, s, 0.5, O.NMS_GT_F32)
    np.testing.assert_array_equal(order[keep[:num.value]], want)          # gpu_nms.pyx:31 order[keep]
    keep2 = np.zeros(3000, np.int32); num2 = ctypes.c_int(0)
    rc = detbox.lib().detbox_nms_host(p(dets), 3000, 0.5, detbox.NMS_GE_F64, p(keep2),
                                      ctypes.cast(ctypes.byref(num2), ctypes.c_void_p), 0)
    assert rc == 0
    np.testing.assert_array_equal(keep2[:num2.value], O.nms(b, s, 0.5, O.NMS_GE_F64))


# ------------------------------------------------------------------------------ NMS variants ----
@pytest.mark.parametrize("seed", range(4))
@pytest.mark.parametrize("thr", [0.3, 0.5])
def test_nms_variants_golden(golden, seed, thr):
    """nms_new / nms_multianchor (libs/boxes/nms.pyx:70-195) through the mirror: keep lists and cluster ids
    bit-exact against the vectors recorded from the compiled reference."""
    import importlib
    cn = importlib.import_module(detbox.libs.__name__ + ".boxes.cython_nms")
    g = golden("nms_variants")
    d = g["d%d" % seed]
    np.testing.assert_array_equal(cn.nms_new(d, float(thr)), g["d%d_new_%g" % (seed, thr)])
    keep, clusters = cn.nms_multianchor(d, float(thr))
    np.testing.assert_array_equal(keep, g["d%d_multi_keep_%g" % (seed, thr)])
    np.testing.assert_array_equal(clusters, g["d%d_multi_clusters_%g" % (seed, thr)])


def test_nms_variants_segmented_vs_oracle():
    from oracle.make_golden import variant_dets
    d = variant_dets(7)
    img = (np.arange(d.shape[0]) % 3).astype(np.int32)
    for variant, fn in (("new", lambda x: (O.nms_new(x, 0.4), None)), ("multianchor", lambda x: O.nms_multianchor(x, 0.4))):
        keep, nk, cl = detbox.nms(cu(d[:, :4]), cu(d[:, 4]), 0.4, detbox.NMS_GE_F64, seg_ids=cu(img), n_seg=3,
                                  variant=variant, want_clusters=True)
        keep = keep[:int(nk[-1])].cpu().numpy(); cl = cl.cpu().numpy()
        want_keep, want_cl = [], np.arange(d.shape[0])
        for i in range(3):
            inds = np.nonzero(img == i)[0]
            k, c = fn(d[inds])
            want_keep.append(inds[np.asarray(k, np.int64)])
            if c is not None:
                want_cl[inds] = inds[c]
        np.testing.assert_array_equal(keep, np.concatenate(want_keep))
        if variant == "multianchor":
            np.testing.assert_array_equal(cl, want_cl)


# ------------------------------------------------------------------------------ dets_cluster ----
def test_dets_cluster_golden(golden):
    """dets_cluster (single_shot.py:401-415) through the mirror against the reference's own outputs: float32 sums in
    NumPy's order (pairwise score sum, sequential row sum), so rows agree to the last bit with the NumPy that recorded
    the vectors; the stated bound is 1e-5 relative (north_star's float tolerance)."""
    import importlib
    ss = importlib.import_module(detbox.libs.__name__ + ".postprocessings.single_shot")
    g = golden("dets_cluster"); nv = golden("nms_variants")
    cases = [(nv["d%d" % s], nv["d%d_multi_keep_%g" % (s, t)], nv["d%d_multi_clusters_%g" % (s, t)], g["d%d_voted_%g" % (s, t)])
             for s in range(4) for t in (0.3, 0.5)]
    cases.append((g["synth_dets"], g["synth_keep"], g["synth_clusters"], g["synth_voted"]))
    n_exact = 0
    for d, keep, clusters, want in cases:
        got = ss.dets_cluster(d, keep, clusters)
        assert got.shape == want.shape and got.dtype == np.float32
        np.testing.assert_allclose(got, want, rtol=1e-5, atol=1e-5 * 2048)
        n_exact += int(np.array_equal(got, want))
    assert n_exact == len(cases), "%d of %d cases bit-exact" % (n_exact, len(cases))
    assert ss.dets_cluster(cases[0][0], [], cases[0][2]) is cases[0][0]          # upstream: no keeps -> dets


def test_dets_cluster_after_device_nms():
    """nms(variant='multianchor') -> dets_cluster without leaving the device (n_keep stays a device scalar)."""
    from oracle.make_golden import variant_dets
    d = variant_dets(5)
    keep, nk, cl = detbox.nms(cu(d[:, :4]), cu(d[:, 4]), 0.4, detbox.NMS_GE_F64, variant="multianchor", want_clusters=True)
    out = detbox.dets_cluster(cu(d), keep, cl, n_keep=nk[-1:])
    m = int(nk[-1])
    want = O.dets_cluster(d, keep[:m].cpu().numpy(), cl.cpu().numpy())
    np.testing.assert_array_equal(out[:m].cpu().numpy(), want)
    assert not out[m:].any()


def test_nms_out_of_range_segment_id_is_memory_safe():
    """a segment id outside [0, n_seg) (here == n_seg and -3, int32 and int64 ids) must never be used as an
    index.  The call completes, and every segment that holds no bad row keeps exactly what it keeps without them."""
    rng = np.random.RandomState(17)
    n, n_seg = 1200, 4
    b, s = sweep_boxes(rng, n)
    seg = rng.randint(0, n_seg, n).astype(np.int32)
    bad = np.nonzero(seg == n_seg - 1)[0][:5]                    # rows of the LAST segment get id n_seg, rows of segment 0 get -3
    bad0 = np.nonzero(seg == 0)[0][:5]
    seg_bad = seg.copy(); seg_bad[bad] = n_seg; seg_bad[bad0] = -3
    keep_ref, nk_ref = keep_list(*detbox.nms(cu(b), cu(s), 0.5, detbox.NMS_GT_F32, seg_ids=cu(seg), n_seg=n_seg))
    for dt in (torch.int32, torch.int64):
        keep, nk = keep_list(*detbox.nms(cu(b), cu(s), 0.5, detbox.NMS_GT_F32, seg_ids=cu(seg_bad).to(dt), n_seg=n_seg))
        torch.cuda.synchronize()
        np.testing.assert_array_equal(nk, nk_ref)                # clamped to the nearest segment == their own here
        np.testing.assert_array_equal(keep, keep_ref)
