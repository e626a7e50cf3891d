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
    close(gi[0].cpu().numpy(), bwd(g, rois, np.float32(0.125), feat.shape), scale=float(np.abs(g).max()) * 4)


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
        close(gi[l].cpu().numpy(), bwd(g[sel], rois[sel], scales[l], f.shape), scale=float(np.abs(g).max()) * 8)


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
        close(gi[l].cpu().numpy(), O.roi_align_tf_bwd(g[:, l * C:(l + 1) * C], rois, scales[l], f.shape), scale=float(np.abs(g).max()) * 8)


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
