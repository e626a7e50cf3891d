"""GPU parity for RoIAlign forward / backward over FPN levels against the CPU oracle
(crop_and_resize.c restated and pinned bit-exactly to the compiled reference; the grid-point kernel
restated from roi_align_kernel.cu).  Tolerance 1e-5 relative to the feature magnitude — forward is
in fact expected bit-exact in bin-centre mode; backward sums float32 atomics in a different order."""
import numpy as np
import pytest
import torch

import detbox
from oracle import oracle as O

pytestmark = pytest.mark.gpu
STRIDES = [4, 8, 16, 32]


def cu(a, dtype=None):
    t = torch.from_numpy(np.ascontiguousarray(a))
    if dtype is not None:
        t = t.to(dtype)
    return t.cuda()


def make_rois(rng, R, B, ih, iw, small=12, large=400):
    h = np.exp(rng.uniform(np.log(small), np.log(large), R)); w = h / rng.uniform(0.5, 2.8, R)
    cx, cy = rng.uniform(0, iw, R), rng.uniform(0, ih, R)
    b = np.stack([cx - w / 2, cy - h / 2, cx + w / 2, cy + h / 2], 1)      # partly outside the canvas
    return np.concatenate([rng.randint(0, B, (R, 1)), b], 1).astype(np.float32)


def close(got, want, scale=1.0, rtol=1e-5):
    err = np.abs(got.astype(np.float64) - want.astype(np.float64)).max() if want.size else 0.0
    assert err <= rtol * max(scale, float(np.abs(want).max()) if want.size else 1.0), err


U = 2.0 ** -24                                            # float32 unit roundoff


def tap_stats(rois, scale, ph, pw, shape, mode, g, lv_sel=None):
    """Per element of a level's gradient map: n = number of (RoI, bin, tap) contributions and S = sum of their absolute
    values, from an independent float64 restatement of the tap positions and weights (roi_align_tf.py:24-40 +
    crop_and_resize.c:44-72,258-264, or roi_align_kernel.cu:32-56,147-150 — whose weights can be negative in the last
    row / column).  They feed the summation-error bound below; slack there covers taps that a float32 rounding moves
    across a pixel boundary (such a tap carries a weight below 1e-6)."""
    B, C, H, W = shape
    cnt = np.zeros((B, H, W), np.int64)
    S = np.zeros((B, C, H, W), np.float64)
    rs, gs = (rois, g) if lv_sel is None else (rois[lv_sel], g[lv_sel])
    for r, gr in zip(rs, np.abs(gs.astype(np.float64))):
        b = int(r[0])
        if not 0 <= b < B:
            continue
        def axis(lo, hi, P, L):
            if mode == "tf":
                b1, b2 = scale * lo, scale * hi
                sp = (b2 - b1) / P
                v = b1 + sp / 2 - 0.5 + np.arange(P) * sp
                ok = (v >= -1e-4) & (v <= L - 1 + 1e-4)
                v = np.clip(v, 0, L - 1)
                i0 = np.floor(v).astype(int); i1 = np.ceil(v).astype(int)
            else:
                s0, e0 = lo * scale, hi * scale
                v = np.arange(P) * (max(e0 - s0 + 1, 0) / (P - 1)) + s0
                ok = (v >= -1e-4) & (v < L + 1e-4)
                v = np.clip(v, 0, L)
                i0 = np.maximum(np.minimum(np.floor(v).astype(int), L - 2), 0); i1 = np.minimum(i0 + 1, L - 1)
            w = v - i0
            return i0, i1, w, ok
        y0, y1, wy, oky = axis(float(r[2]), float(r[4]), ph, H)
        x0, x1, wx, okx = axis(float(r[1]), float(r[3]), pw, W)
        m = oky[:, None] & okx[None, :]
        for ys, wys in ((y0, 1 - wy), (y1, wy)):
            for xs, wxs in ((x0, 1 - wx), (x1, wx)):
                yy, xx = np.broadcast_to(ys[:, None], m.shape)[m], np.broadcast_to(xs[None, :], m.shape)[m]
                np.add.at(cnt[b], (yy, xx), 1)
                t = gr[:, m] * np.abs(wys[:, None] * wxs[None, :])[m][None, :]
                for c in range(C):
                    np.add.at(S[b, c], (yy, xx), t[c])
    return cnt, S


def assert_bwd_within_bound(got, want, stats, gmax):
    """|got - want| <= 2 * gamma_(n + 8) * sum|terms| per element: both sides add the SAME float32 terms
    (crop_and_resize.c:258-264 / roi_align_kernel.cu:147-150) in different orders, and each is within
    gamma_(n-1) * sum|terms| of the exact sum (Higham, Accuracy and Stability of Numerical Algorithms, eq. 4.4);
    n and sum|terms| from tap_stats."""
    counts, abs_sum = stats
    n = counts[:, None, :, :].astype(np.float64) + 8
    bound = 2 * (n * U / (1 - n * U)) * (abs_sum * (1 + 1e-4) + 1e-5 * gmax)
    err = np.abs(got.astype(np.float64) - want.astype(np.float64))
    bad = err > bound
    assert not bad.any(), "backward outside the summation bound at %d elements, worst err/bound %.3g" % (bad.sum(), (err[bad] / bound[bad]).max())
    assert float(err.max()) <= 1e-5 * max(1.0, float(np.abs(want).max())), "backward above 1e-5 relative"      # north_star's tolerance


def as_layout(t, nhwc):
    return t.contiguous(memory_format=torch.channels_last) if nhwc else t.contiguous()


def test_kat5_bin_centre(golden):
    g = golden("roialign")
    ramp = np.arange(16, dtype=np.float32).reshape(1, 1, 4, 4)
    rois = np.asarray([[0, 0, 0, 3, 3]], np.float32)
    out = detbox.roi_align_forward([cu(ramp)], cu(rois), [1.0], 2, 2, detbox.ROI_BIN_CENTRE)
    np.testing.assert_array_equal(out.cpu().numpy(), g["kat5_fwd"])
    gi = detbox.roi_align_backward(torch.ones((1, 1, 2, 2), device="cuda"), [ramp.shape], cu(rois), [1.0], detbox.ROI_BIN_CENTRE)
    np.testing.assert_allclose(gi[0].cpu().numpy(), g["kat5_bwd"], atol=1e-6)


@pytest.mark.parametrize("nhwc", [False, True])
@pytest.mark.parametrize("mode", ["tf", "grid"])
@pytest.mark.parametrize("ph,pw", [(7, 7), (3, 5), (1, 1), (14, 14)])
def test_single_level(nhwc, mode, ph, pw):
    if mode == "grid" and ph == 1:
        pytest.skip("aligned size 1 divides by zero in the reference (roi_align_kernel.cu:40)")
    rng = np.random.RandomState(ph * 10 + pw)
    B, C, H, W = 3, 8, 25, 31
    feat = rng.normal(0, 1, (B, C, H, W)).astype(np.float32)
    rois = make_rois(rng, 60, B, H * 8, W * 8)
    rois[0, 0] = 7                                         # batch index out of range -> zeros
    m = detbox.ROI_BIN_CENTRE if mode == "tf" else detbox.ROI_GRID_POINT
    fwd = O.roi_align_tf_fwd if mode == "tf" else O.roi_align_grid_fwd
    bwd = O.roi_align_tf_bwd if mode == "tf" else O.roi_align_grid_bwd
    out = detbox.roi_align_forward([as_layout(cu(feat), nhwc)], cu(rois), [0.125], ph, pw, m)
    assert out.shape == (60, C, ph, pw)
    want = fwd(feat, rois, np.float32(0.125), ph, pw)
    if mode == "tf":
        np.testing.assert_array_equal(out.cpu().numpy(), want)            # same float32 op sequence
    else:
        close(out.cpu().numpy(), want)
    g = rng.normal(0, 1, want.shape).astype(np.float32)
    gi = detbox.roi_align_backward(as_layout(cu(g), nhwc), [feat.shape], cu(rois), [0.125], m, channels_last=nhwc)
    assert_bwd_within_bound(gi[0].cpu().numpy(), bwd(g, rois, np.float32(0.125), feat.shape),
                            tap_stats(rois, 0.125, ph, pw, feat.shape, mode, g), float(np.abs(g).max()))


@pytest.mark.parametrize("kind", ["avg", "max"])
def test_grid_point_pooled(kind):
    """RoIAlignAvg / RoIAlignMax (modules/roi_align.py:23-46): (P+1)^2 grid samples + 2x2/1 pool, fused."""
    rng = np.random.RandomState(3)
    B, C, H, W = 2, 6, 40, 52
    feat = rng.normal(0, 1, (B, C, H, W)).astype(np.float32)
    rois = make_rois(rng, 40, B, H * 16, W * 16)
    out = detbox.roi_align_forward([cu(feat)], cu(rois), [1.0 / 16], 7, 7, detbox.ROI_GRID_POINT, pool=kind)
    want = O.pool2x2(O.roi_align_grid_fwd(feat, rois, np.float32(1.0 / 16), 8, 8), kind)
    close(out.cpu().numpy(), want)


def pyramid(rng, B, C, ih, iw):
    return [rng.normal(0, 1, (B, C, ih // s, iw // s)).astype(np.float32) for s in STRIDES]


@pytest.mark.parametrize("nhwc", [False, True])
@pytest.mark.parametrize("mode", ["tf", "grid"])
def test_pyramid_level_assigned(nhwc, mode):
    """PyramidRoIAlign (pyramid_roi_align.py:30-60): in-kernel level choice, output in RoI order."""
    rng = np.random.RandomState(21)
    B, C, ih, iw = 2, 16, 256, 384
    feats = pyramid(rng, B, C, ih, iw)
    cfg = O.Cfg(strides=tuple(STRIDES), anchor_scales=[[2, 2.52, 3.17], [4, 5.04, 6.35], [8, 10.08, 12.7], [16, 20.16, 25.4]])
    ref_areas = O.pyramid_ref_areas(cfg)
    rois = make_rois(rng, 300, B, ih, iw, 8, 500)
    lv_want = O.assign_levels(rois[:, 1:], ref_areas)
    lv = detbox.assign_levels(cu(rois[:, 1:]), ref_areas).cpu().numpy()
    np.testing.assert_array_equal(lv, lv_want)
    assert len(np.unique(lv)) == 4
    m = detbox.ROI_BIN_CENTRE if mode == "tf" else detbox.ROI_GRID_POINT
    scales = [np.float32(1.0 / s) for s in STRIDES]
    fl = [as_layout(cu(f), nhwc) for f in feats]
    out = detbox.roi_align_forward(fl, cu(rois), scales, 7, 7, m, ref_areas=ref_areas)
    want = O.pyramid_roi_align(feats, rois[:, 1:], rois[:, 0], STRIDES, ref_areas, 7, 7, mode)
    close(out.cpu().numpy(), want)
    out2 = detbox.roi_align_forward(fl, cu(rois), scales, 7, 7, m, levels=cu(lv_want))      # explicit levels
    assert torch.equal(out, out2)
    # backward: per-level oracle on the RoIs of that level
    g = rng.normal(0, 1, want.shape).astype(np.float32)
    gi = detbox.roi_align_backward(as_layout(cu(g), nhwc), [f.shape for f in feats], cu(rois), scales, m,
                                   ref_areas=ref_areas, channels_last=nhwc)
    bwd = O.roi_align_tf_bwd if mode == "tf" else O.roi_align_grid_bwd
    for l, f in enumerate(feats):
        sel = lv_want == l
        assert_bwd_within_bound(gi[l].cpu().numpy(), bwd(g[sel], rois[sel], scales[l], f.shape),
                                tap_stats(rois, float(scales[l]), 7, 7, f.shape, mode, g, sel), float(np.abs(g).max()))


def test_pyramid_all_levels():
    """PyramidRoIAlign2 (pyramid_roi_align.py:82-91): every RoI on every level, channel concat."""
    rng = np.random.RandomState(22)
    B, C, ih, iw = 2, 8, 128, 192
    feats = pyramid(rng, B, C, ih, iw)
    rois = make_rois(rng, 100, B, ih, iw, 8, 200)
    scales = [np.float32(1.0 / s) for s in STRIDES]
    out = detbox.roi_align_forward([cu(f) for f in feats], cu(rois), scales, 7, 7, detbox.ROI_BIN_CENTRE, all_levels=True)
    want = O.pyramid_roi_align(feats, rois[:, 1:], rois[:, 0], STRIDES, None, 7, 7, "tf", all_levels=True)
    assert out.shape == (100, 4 * C, 7, 7)
    np.testing.assert_array_equal(out.cpu().numpy(), want)
    g = rng.normal(0, 1, want.shape).astype(np.float32)
    gi = detbox.roi_align_backward(cu(g), [f.shape for f in feats], cu(rois), scales, detbox.ROI_BIN_CENTRE, all_levels=True)
    for l, f in enumerate(feats):
        gl = np.ascontiguousarray(g[:, l * C:(l + 1) * C])
        assert_bwd_within_bound(gi[l].cpu().numpy(), O.roi_align_tf_bwd(gl, rois, scales[l], f.shape),
                                tap_stats(rois, float(scales[l]), 7, 7, f.shape, "tf", gl), float(np.abs(g).max()))


def test_linearity_and_constant_plane_at_baseline_shape():
    """Config-3 shape (B=8, C=256, P2-P5 of 1024x2048, 4096 RoIs): size-independent properties.
    fwd(a*x + y) == a*fwd(x) + fwd(y) (bilinear sampling is linear in the features), a constant
    plane samples to that constant wherever the sample is inside, and <fwd(x), g> == <x, bwd(g)>
    (backward is the adjoint of forward)."""
    g = torch.Generator(device="cuda"); g.manual_seed(3)
    B, C = 8, 256
    shapes = [(B, C, 1024 // s, 2048 // s) for s in STRIDES]
    x = [torch.randn(s, device="cuda", generator=g) for s in shapes]
    rng = np.random.RandomState(4)
    rois = cu(make_rois(rng, 4096, B, 1024, 2048, 12, 400))
    cfg = O.Cfg(strides=tuple(STRIDES), anchor_scales=[[2, 2.52, 3.17], [4, 5.04, 6.35], [8, 10.08, 12.7], [16, 20.16, 25.4]])
    ref_areas = O.pyramid_ref_areas(cfg)
    scales = [1.0 / s for s in STRIDES]
    f = lambda feats: detbox.roi_align_forward(feats, rois, scales, 7, 7, detbox.ROI_BIN_CENTRE, ref_areas=ref_areas)
    fx = f(x)
    f2 = f([2.0 * t for t in x])
    assert torch.allclose(f2, 2.0 * fx, rtol=1e-5, atol=1e-5)
    ones = f([torch.ones_like(t) for t in x])
    assert ((ones - 1).abs() < 1e-5).logical_or(ones == 0).all()
    go = torch.randn(fx.shape, device="cuda", generator=g)
    gi = detbox.roi_align_backward(go, shapes, rois, scales, detbox.ROI_BIN_CENTRE, ref_areas=ref_areas)
    lhs = float((fx.double() * go.double()).sum())
    rhs = float(sum((a.double() * b.double()).sum() for a, b in zip(x, gi)))
    assert abs(lhs - rhs) <= 1e-5 * max(abs(lhs), 1.0) + 1e-2


@pytest.mark.parametrize("mode", ["tf", "grid"])
def test_backward_is_the_gradient_finite_differences(mode):
    """The reference's commented-out gradcheck (roi_align_tf/test.py:101-120: RoIAlign(3,3) on 2x2x10x10) as a
    central finite-difference test in float64 arithmetic on the host: d<fwd(x), g>/dx == bwd(g)."""
    rng = np.random.RandomState(5)
    B, C, H, W, ph, pw = 2, 2, 10, 10, 3, 3
    feat = rng.normal(0, 1, (B, C, H, W)).astype(np.float32)
    rois = np.asarray([[0, 1.3, 2.2, 7.9, 8.4], [1, 0.0, 0.0, 9.0, 9.0], [1, 3.5, 4.1, 5.2, 6.6], [0, -2.0, 1.0, 4.0, 12.0]], np.float32)
    m = detbox.ROI_BIN_CENTRE if mode == "tf" else detbox.ROI_GRID_POINT
    g = rng.normal(0, 1, (rois.shape[0], C, ph, pw)).astype(np.float32)
    gi = detbox.roi_align_backward(cu(g), [feat.shape], cu(rois), [1.0], m)[0].cpu().numpy()
    f = lambda x: float((detbox.roi_align_forward([cu(x)], cu(rois), [1.0], ph, pw, m).cpu().numpy().astype(np.float64) * g).sum())
    eps = 0.5                                             # the op is linear in the features: any step gives the exact slope
    for _ in range(25):
        idx = tuple(rng.randint(0, s) for s in feat.shape)
        xp, xm = feat.copy(), feat.copy()
        xp[idx] += eps; xm[idx] -= eps
        num = (f(xp) - f(xm)) / (2 * eps)
        assert abs(num - gi[idx]) <= 1e-4 * max(1.0, abs(num)), (idx, num, gi[idx])


# ---- round 2: the shared-memory staged NCHW kernels -----------------------------------------------------------------
@pytest.mark.parametrize("mode", ["tf", "grid"])
@pytest.mark.parametrize("W", [128, 126, 61])
def test_nchw_staged_window_sizes(mode, W):
    """Windows of every class the staged kernels distinguish: one pass (small RoIs), several passes (a channel's window
    is a few thousand floats), the direct fallback (window above the 40 KB budget: a RoI covering the whole 128 x 128
    map), 16-byte row pieces (W % 4 == 0) and scalar staging (W = 126, 61).  Forward bit-exact against crop_and_resize.c
    in bin-centre mode; backward inside the summation bound."""
    rng = np.random.RandomState(W)
    B, C, H = 2, 40, 128                                   # C = 40: a partial second channel chunk
    feat = rng.normal(0, 1, (B, C, H, W)).astype(np.float32)
    rois = make_rois(rng, 24, B, H * 4, W * 4, 20, 200)
    big = np.asarray([[0, -10, -10, W * 4 + 10, H * 4 + 10], [1, 0, 0, W * 4 - 1, H * 4 - 1], [1, 8, 4, W * 3, H * 3.5],
                      [0, 30, 20, 30.5, 20.5], [1, W * 4 + 50, 10, W * 4 + 90, 60]], np.float32)      # whole map, mid, sub-pixel, outside
    rois = np.concatenate([rois, big], 0)
    m = detbox.ROI_BIN_CENTRE if mode == "tf" else detbox.ROI_GRID_POINT
    fwd = O.roi_align_tf_fwd if mode == "tf" else O.roi_align_grid_fwd
    bwd = O.roi_align_tf_bwd if mode == "tf" else O.roi_align_grid_bwd
    out = detbox.roi_align_forward([cu(feat)], cu(rois), [0.25], 7, 7, m).cpu().numpy()
    want = fwd(feat, rois, np.float32(0.25), 7, 7)
    if mode == "tf":
        np.testing.assert_array_equal(out, want)
    else:
        close(out, want)
    g = rng.normal(0, 1, want.shape).astype(np.float32)
    gi = detbox.roi_align_backward(cu(g), [feat.shape], cu(rois), [0.25], m)[0].cpu().numpy()
    assert_bwd_within_bound(gi, bwd(g, rois, np.float32(0.25), feat.shape), tap_stats(rois, 0.25, 7, 7, feat.shape, mode, g),
                            float(np.abs(g).max()))


def test_config3_values_vs_openmp_reference():
    """BASELINE configs[2] shape (B=8, C=256, P2-P5 of 1024x2048, 4096 RoIs, NCHW): VALUES of 320 RoIs of that batch
    against the oracle of crop_and_resize.c (bit-exact: same float32 operation sequence), and the backward of those
    RoIs into the dense config-3 gradient maps inside the summation bound."""
    g = torch.Generator(device="cuda"); g.manual_seed(11)
    B, C = 8, 256
    shapes = [(B, C, 1024 // s, 2048 // s) for s in STRIDES]
    x = [torch.randn(s, device="cuda", generator=g) for s in shapes]
    rng = np.random.RandomState(12)
    rois_np = make_rois(rng, 4096, B, 1024, 2048, 12, 400)
    cfg = O.Cfg(strides=tuple(STRIDES), anchor_scales=[[2, 2.52, 3.17], [4, 5.04, 6.35], [8, 10.08, 12.7], [16, 20.16, 25.4]])
    ref_areas = O.pyramid_ref_areas(cfg)
    scales = [np.float32(1.0 / s) for s in STRIDES]
    out = detbox.roi_align_forward(x, cu(rois_np), scales, 7, 7, detbox.ROI_BIN_CENTRE, ref_areas=ref_areas)
    lv = O.assign_levels(rois_np[:, 1:], ref_areas)
    pick = np.sort(rng.choice(4096, 320, replace=False))
    out_np = out[torch.from_numpy(pick).cuda()].cpu().numpy()
    go = torch.zeros(out.shape, device="cuda")
    go_pick = rng.normal(0, 1, (320, C, 7, 7)).astype(np.float32)
    go[torch.from_numpy(pick).cuda()] = cu(go_pick)
    gi = detbox.roi_align_backward(go, shapes, cu(rois_np), scales, detbox.ROI_BIN_CENTRE, ref_areas=ref_areas)
    for l in range(4):
        sel = lv[pick] == l
        if not sel.any():
            continue
        f = x[l].cpu().numpy()
        want = O.roi_align_tf_fwd(f, rois_np[pick][sel], scales[l], 7, 7)
        np.testing.assert_array_equal(out_np[sel], want)
        del f
        rs, gs = rois_np[pick][sel], go_pick[sel]
        assert_bwd_within_bound(gi[l].cpu().numpy(), O.roi_align_tf_bwd(gs, rs, scales[l], shapes[l]),
                                tap_stats(rs, float(scales[l]), 7, 7, shapes[l], "tf", gs), float(np.abs(gs).max()))
    assert len(np.unique(lv[pick])) == 4


def test_nchw_staged_backward_opt_in_subprocess():
    """The default NCHW backward is the separable kernel in bin-centre mode (PW <= 8) and the exact-term gather kernel in
    grid-point mode (one red.global.add.v4 per touched 16-byte piece either way).  The other paths stay selectable per
    process — DETBOX_ROI_BWD_GATHER=1 (exact-term gather in bin-centre mode too), DETBOX_ROI_BWD_STAGED=1 (sums in a
    shared-memory window), DETBOX_ROI_BWD_DIRECT=1 (4 scalar reds per output element); all measured slower on B200 — so
    each is exercised in a process of its own against the same bound as the default path."""
    import os, subprocess, sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    code = (
        "import sys; sys.path.insert(0, %r); sys.path.insert(0, %r)\n"
        "import numpy as np, torch, detbox\n"
        "from oracle import oracle as O\n"
        "import test_gpu_roi as T\n"
        "rng = np.random.RandomState(7)\n"
        "B, C, H, W = 2, 40, 96, 128\n"
        "feat_shape = (B, C, H, W)\n"
        "rois = T.make_rois(rng, 40, B, H * 4, W * 4, 20, 200)\n"
        "g = rng.normal(0, 1, (40, C, 7, 7)).astype(np.float32)\n"
        "for mode, m, bwd in (('tf', detbox.ROI_BIN_CENTRE, O.roi_align_tf_bwd), ('grid', detbox.ROI_GRID_POINT, O.roi_align_grid_bwd)):\n"
        "    gi = detbox.roi_align_backward(T.cu(g), [feat_shape], T.cu(rois), [0.25], m)[0].cpu().numpy()\n"
        "    T.assert_bwd_within_bound(gi, bwd(g, rois, np.float32(0.25), feat_shape), T.tap_stats(rois, 0.25, 7, 7, feat_shape, mode, g), float(np.abs(g).max()))\n"
        "print('staged backward ok')\n" % (root, os.path.join(root, "tests")))
    # (DETBOX_ROI_BWD_NOTILED=1: without it the bin-centre case takes the tile-centric dense backward, which has a test of
    #  its own below; alone it selects the separable scatter kernel)
    for var in ("DETBOX_ROI_BWD_STAGED", "DETBOX_ROI_BWD_DIRECT", "DETBOX_ROI_BWD_GATHER", "DETBOX_ROI_BWD_NOTILED"):
        env = dict(os.environ, **{var: "1", "DETBOX_ROI_BWD_NOTILED": "1"})
        r = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True, timeout=600)
        assert r.returncode == 0 and "staged backward ok" in r.stdout, var + "\n" + r.stdout + r.stderr


def _bwd_both(g, shapes, rois, scales, mode, **kw):
    """roi_align_backward through the tile-centric dense kernel and through the RoI-centric scatter kernels."""
    import sys
    mod = sys.modules[detbox.roi_align_backward.__module__]
    before = mod.tiled_backward_calls
    tiled = [t.cpu().numpy() for t in detbox.roi_align_backward(cu(g), shapes, cu(rois), scales, mode, deterministic=True, **kw)]
    assert mod.tiled_backward_calls == before + 1, "the tiled entry point did not take this configuration"
    tiled2 = [t.cpu().numpy() for t in detbox.roi_align_backward(cu(g), shapes, cu(rois), scales, mode, deterministic=True, **kw)]
    scat = [t.cpu().numpy() for t in detbox.roi_align_backward(cu(g), shapes, cu(rois), scales, mode, deterministic=False, **kw)]
    for a, b in zip(tiled, tiled2):
        np.testing.assert_array_equal(a, b)                  # fixed summation order: bit-identical from run to run
    return tiled, scat


@pytest.mark.parametrize("case", ["pyramid_odd", "tiny_rois", "crowded_tile", "crop_norm", "no_rois", "partial_chunk"])
def test_tiled_backward(case):
    """detbox_roi_align_backward_tiled (the default for NCHW / bin-centre or normalised crop / no pool / pooled size <= 8 x 8):
    every gradient map written once, tile by tile.  Against the float64-derived bound of the oracle (crop_and_resize.c:207-268) and against the
    scatter kernels; map sizes that are not multiples of the 8 x 32 tile, RoIs smaller than their 7 x 7 bins (more than two
    bin columns per pixel: the generic column loop), more RoIs on one tile than one list round holds (accumulating rounds),
    channel counts that are not a multiple of the 32-channel chunk, and R = 0 (all zeros)."""
    rng = np.random.RandomState(sum(map(ord, case)))
    mode, name = detbox.ROI_BIN_CENTRE, "tf"
    B, C = 2, 40
    if case == "pyramid_odd":
        shapes = [(B, C, 61, 93), (B, C, 31, 47), (B, C, 16, 24), (B, C, 8, 12)]
        scales = [np.float32(1.0 / s) for s in STRIDES]
        rois = make_rois(rng, 300, B, 244, 372, 8, 300)
        kw = dict(levels=cu(rng.randint(0, 4, 300).astype(np.int32)))
    elif case == "tiny_rois":
        shapes, scales = [(B, C, 40, 64)], [np.float32(0.25)]
        rois = make_rois(rng, 200, B, 160, 256, 1.5, 14)          # 0.4 .. 3.5 feature pixels for 7 bins
        kw = {}
    elif case == "crowded_tile":
        shapes, scales = [(1, 8, 24, 64)], [np.float32(0.25)]
        rois = make_rois(rng, 700, 1, 40, 60, 20, 60)              # all 700 on the first tiles: 256-entry list rounds
        kw = {}
    elif case == "crop_norm":
        mode, name = detbox.ROI_CROP_NORM, "crop"
        shapes, scales = [(B, C, 50, 70)], [np.float32(1.0)]
        y1, x1 = rng.uniform(-0.2, 0.8, 120), rng.uniform(-0.2, 0.8, 120)
        rois = np.stack([rng.randint(0, B, 120), y1, x1, y1 + rng.uniform(0.05, 0.6, 120), x1 + rng.uniform(0.05, 0.6, 120)], 1).astype(np.float32)
        kw = {}
    elif case == "no_rois":
        shapes, scales = [(B, C, 20, 40)], [np.float32(0.25)]
        rois = np.zeros((0, 5), np.float32)
        kw = {}
    else:
        shapes, scales = [(B, 37, 33, 65)], [np.float32(0.125)]
        rois = make_rois(rng, 150, B, 264, 520, 16, 300)
        kw = {}
    Cc = shapes[0][1]
    g = rng.normal(0, 1, (rois.shape[0], Cc, 7, 7)).astype(np.float32)
    tiled, scat = _bwd_both(g, shapes, rois, scales, mode, **kw)
    if case == "no_rois":
        assert not tiled[0].any() and not scat[0].any()
        return
    lv = kw["levels"].cpu().numpy() if "levels" in kw else np.zeros(rois.shape[0], np.int64)
    gmax = float(np.abs(g).max())
    for l, shp in enumerate(shapes):
        sel = lv == l
        if name == "crop":
            want = O.crop_and_resize_bwd(g[sel], rois[sel, 1:], rois[sel, 0].astype(np.int32), shp)
            close(tiled[l], want, gmax * 8)
            close(tiled[l], scat[l], gmax * 8)
            continue
        want = O.roi_align_tf_bwd(g[sel], rois[sel], scales[l], shp)
        stats = tap_stats(rois, float(scales[l]), 7, 7, shp, name, g, sel)
        assert_bwd_within_bound(tiled[l], want, stats, gmax)
        assert_bwd_within_bound(scat[l], want, stats, gmax)
        # untouched pixels are exact zeros in both (the tiled kernel writes them, the scatter path memsets them)
        odd = (tiled[l] == 0) != (scat[l] == 0)
        assert not odd.any() or max(np.abs(tiled[l][odd]).max(), np.abs(scat[l][odd]).max()) < 1e-6 * gmax


def test_results_do_not_depend_on_the_roi_processing_order():
    """detbox_roi_cfg.roi_order only changes which block handles which RoI (an L2-friendly order by (level, image, row
    band, column band), detbox_roi_locality_order): the forward must be bit-identical for the RoI order, the automatic
    order and a random permutation; the backward (float32 atomics: only the order of additions moves) stays inside the
    summation bound.  The order itself must be a permutation sorted by the documented key."""
    rng = np.random.RandomState(41)
    B, C, ih, iw = 2, 24, 256, 384
    feats = pyramid(rng, B, C, ih, iw)
    cfg = O.Cfg(strides=tuple(STRIDES), anchor_scales=[[2, 2.52, 3.17], [4, 5.04, 6.35], [8, 10.08, 12.7], [16, 20.16, 25.4]])
    ref_areas = O.pyramid_ref_areas(cfg)
    rois = make_rois(rng, 700, B, ih, iw, 8, 300)                      # >= ORDER_MIN_ROIS: order="auto" computes one
    scales = [np.float32(1.0 / s) for s in STRIDES]
    fl = [cu(f) for f in feats]
    shapes = [f.shape for f in feats]
    od = detbox.ops_roi.roi_locality_order(cu(rois), scales, shapes, None, ref_areas)
    o = od.cpu().numpy()
    assert np.array_equal(np.sort(o), np.arange(700))
    lv = O.assign_levels(rois[:, 1:], ref_areas)
    key = ((lv[o].astype(np.int64) << 8 | rois[o, 0].astype(np.int64)) << 11)
    assert (np.diff(key) <= 0).all()                                   # (level, image) major, descending
    perm = cu(rng.permutation(700).astype(np.int32))
    g = rng.normal(0, 1, (700, C, 7, 7)).astype(np.float32)
    for mode, m in (("tf", detbox.ROI_BIN_CENTRE), ("grid", detbox.ROI_GRID_POINT)):
        base = detbox.roi_align_forward(fl, cu(rois), scales, 7, 7, m, ref_areas=ref_areas, order=False)
        for order in ("auto", od, perm):
            assert torch.equal(detbox.roi_align_forward(fl, cu(rois), scales, 7, 7, m, ref_areas=ref_areas, order=order), base)
        bwd = O.roi_align_tf_bwd if mode == "tf" else O.roi_align_grid_bwd
        for order in (False, "auto", perm):
            gi = detbox.roi_align_backward(cu(g), shapes, cu(rois), scales, m, ref_areas=ref_areas, order=order)
            for l, f in enumerate(feats):
                sel = lv == l
                assert_bwd_within_bound(gi[l].cpu().numpy(), bwd(g[sel], rois[sel], scales[l], f.shape),
                                        tap_stats(rois, float(scales[l]), 7, 7, f.shape, mode, g, sel), float(np.abs(g).max()))
