"""Same-box cross-check against the reference's OWN CUDA kernels.

`oracle/build_ref.py --cuda` compiles libs/nms/nms_kernel.cu, libs/layers/roi_align/src/cuda/roi_align_kernel.cu,
libs/layers/roi_align_tf/src/cuda/crop_and_resize_kernel.cu and libs/layers/roi_target/src/cuda/roi_target_kernel.cu
unmodified for sm_100a into oracle/_ref/libref_cuda_*.so (git-ignored, travels to the GPU box).  These pieces have
no CPU implementation upstream, so this is what pins the `>` NMS comparator, the grid-point RoIAlign, the GPU
crop-and-resize and the RoI-target kernel to the reference itself rather than to a restatement.  Skipped when the
prebuilt libraries are absent."""
import ctypes
import os

import numpy as np
import pytest
import torch

import detbox

pytestmark = pytest.mark.gpu
REF = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle", "_ref")
c_int, c_float, c_void_p = ctypes.c_int, ctypes.c_float, ctypes.c_void_p


def ref_lib(name):
    path = os.path.join(REF, "libref_cuda_%s.so" % name)
    if not os.path.exists(path):
        pytest.skip("%s not built (oracle/build_ref.py --cuda)" % path)
    return ctypes.CDLL(path)


def cu(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def sweep_boxes(rng, n, shrink=1.0):
    h = np.exp(rng.uniform(np.log(12), np.log(264), n)); w = h / 2.44
    cx, cy = rng.uniform(0, 2048, n) / shrink, rng.uniform(0, 1024, n) / shrink
    b = np.stack([cx - w / 2, cy - h / 2, cx + w / 2, cy + h / 2], 1)
    return b.astype(np.float32)


# ------------------------------------------------------------------------------------------ NMS ----
@pytest.mark.parametrize("kind,n", [("sparse", 3000), ("dense", 3000), ("intgrid", 2500), ("sparse", 20000), ("tiny", 5)])
@pytest.mark.parametrize("thr", [0.5, 0.3])
def test_nms_gt_comparator_vs_reference_kernel(kind, n, thr):
    """_nms (libs/nms/nms_kernel.cu:103-156): sorted host dets in, indices into the sorted array out."""
    L = ref_lib("nms")
    fn = L._Z4_nmsPiS_PKfiifi
    fn.restype = None
    fn.argtypes = [c_void_p, c_void_p, c_void_p, c_int, c_int, c_float, c_int]
    rng = np.random.RandomState(n + int(thr * 10))
    if kind == "intgrid":
        x1 = rng.randint(0, 40, n); y1 = rng.randint(0, 40, n)
        b = np.stack([x1, y1, x1 + rng.randint(0, 10, n), y1 + rng.randint(0, 10, n)], 1).astype(np.float32)
    else:
        b = sweep_boxes(rng, n, 6.0 if kind == "dense" else 1.0)
    s = np.sort(rng.permutation(n).astype(np.float32) / n)[::-1]
    dets = np.ascontiguousarray(np.concatenate([b, s[:, None]], 1), dtype=np.float32)       # score-descending
    keep = np.zeros(n, np.int32); num = np.zeros(1, np.int32)
    fn(keep.ctypes.data, num.ctypes.data, dets.ctypes.data, n, 5, thr, 0)
    want = keep[:num[0]].astype(np.int64)
    got, nk = detbox.nms(cu(dets[:, :4]), cu(dets[:, 4]), thr, detbox.NMS_GT_F32)
    got = got[:int(nk[-1])].cpu().numpy()
    np.testing.assert_array_equal(got, want)
    # the library's own drop-in for the same symbol
    keep2 = np.zeros(n, np.int32); num2 = np.zeros(1, np.int32)
    detbox.lib()._nms(keep2.ctypes.data, num2.ctypes.data, dets.ctypes.data, n, 5, c_float(thr), 0)
    np.testing.assert_array_equal(keep2[:num2[0]], keep[:num[0]])


def test_nms_100k_exact_vs_reference_kernel():
    """BASELINE configs[1] top end: n = 100,000 boxes, keep list bit-exact against the reference's own nms_kernel.cu
    (1.25 GB mask on the device, scanned on the host by the reference's launcher, nms_kernel.cu:103-156)."""
    L = ref_lib("nms")
    fn = L._Z4_nmsPiS_PKfiifi
    fn.restype = None
    fn.argtypes = [c_void_p, c_void_p, c_void_p, c_int, c_int, c_float, c_int]
    n = 100000
    rng = np.random.RandomState(3)
    b = sweep_boxes(rng, n)
    s = np.sort(rng.permutation(n).astype(np.float32) / n)[::-1]
    dets = np.ascontiguousarray(np.concatenate([b, s[:, None]], 1), dtype=np.float32)
    keep = np.zeros(n, np.int32); num = np.zeros(1, np.int32)
    fn(keep.ctypes.data, num.ctypes.data, dets.ctypes.data, n, 5, 0.5, 0)
    want = keep[:num[0]].astype(np.int64)
    got, nk = detbox.nms(cu(dets[:, :4]), cu(dets[:, 4]), 0.5, detbox.NMS_GT_F32)
    np.testing.assert_array_equal(got[:int(nk[-1])].cpu().numpy(), want)
    keep2 = np.zeros(n, np.int32); num2 = np.zeros(1, np.int32)
    detbox.lib()._nms(keep2.ctypes.data, num2.ctypes.data, dets.ctypes.data, n, 5, c_float(0.5), 0)
    np.testing.assert_array_equal(keep2[:num2[0]], keep[:num[0]])


# ------------------------------------------------------------------------- grid-point RoIAlign ----
def make_rois(rng, R, B, ih, iw):
    h = np.exp(rng.uniform(np.log(12), np.log(400), R)); w = h / rng.uniform(0.5, 2.8, R)
    cx, cy = rng.uniform(0, iw, R), rng.uniform(0, ih, R)
    b = np.stack([cx - w / 2, cy - h / 2, cx + w / 2, cy + h / 2], 1)
    return np.concatenate([rng.randint(0, B, (R, 1)), b], 1).astype(np.float32)


@pytest.mark.parametrize("ph,pw", [(7, 7), (8, 8), (3, 5)])
def test_grid_point_roi_align_vs_reference_kernel(ph, pw):
    """ROIAlignForward / ROIAlignBackward (roi_align_kernel.cu:15-68, 97-153) through their launchers."""
    L = ref_lib("roi_align")
    fwd, bwd = L.ROIAlignForwardLaucher, L.ROIAlignBackwardLaucher
    fwd.argtypes = [c_void_p, c_float, c_int, c_int, c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p]
    bwd.argtypes = [c_void_p, c_float, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p]
    rng = np.random.RandomState(ph)
    B, C, H, W, R, scale = 3, 16, 50, 68, 200, 1.0 / 16
    feat = cu(rng.normal(0, 1, (B, C, H, W)).astype(np.float32))
    rois = cu(make_rois(rng, R, B, H * 16, W * 16))
    want = torch.zeros((R, C, ph, pw), device="cuda")
    torch.cuda.synchronize()
    fwd(feat.data_ptr(), scale, R, H, W, C, ph, pw, rois.data_ptr(), want.data_ptr(), None)
    torch.cuda.synchronize()
    got = detbox.roi_align_forward([feat], rois, [scale], ph, pw, detbox.ROI_GRID_POINT)
    err = (got - want).abs().max().item()
    assert err <= 1e-5 * max(1.0, want.abs().max().item()), err
    g = cu(rng.normal(0, 1, (R, C, ph, pw)).astype(np.float32))
    want_g = torch.zeros((B, C, H, W), device="cuda")
    torch.cuda.synchronize()
    bwd(g.data_ptr(), scale, B, R, H, W, C, ph, pw, rois.data_ptr(), want_g.data_ptr(), None)
    torch.cuda.synchronize()
    got_g = detbox.roi_align_backward(g, [tuple(feat.shape)], rois, [scale], detbox.ROI_GRID_POINT)[0]
    err = (got_g - want_g).abs().max().item()
    assert err <= 1e-5 * max(1.0, want_g.abs().max().item()), err          # north_star's 1e-5 relative; float32 atomics on both sides, different orders


# ---------------------------------------------------------------------------- crop_and_resize ----
@pytest.mark.parametrize("ch,cw,ext", [(7, 7, 0.0), (14, 14, -1.5), (1, 1, 0.0), (3, 6, 2.0)])
def test_crop_and_resize_vs_reference_kernel(ch, cw, ext):
    """CropAndResizeKernel / CropAndResizeBackpropImageKernel (crop_and_resize_kernel.cu:10-165)."""
    L = ref_lib("crop_and_resize")
    fwd, bwd = L.CropAndResizeLaucher, L.CropAndResizeBackpropImageLaucher
    fwd.restype = bwd.restype = None
    fwd.argtypes = [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_float, c_void_p, c_void_p]
    bwd.argtypes = [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_void_p, c_void_p]
    rng = np.random.RandomState(ch * 7 + cw)
    B, C, H, W, R = 3, 8, 40, 56, 150
    img = cu(rng.normal(0, 1, (B, C, H, W)).astype(np.float32))
    y1, x1 = rng.uniform(-0.2, 0.9, R), rng.uniform(-0.2, 0.9, R)
    boxes = np.stack([y1, x1, y1 + rng.uniform(0.02, 0.6, R), x1 + rng.uniform(0.02, 0.6, R)], 1).astype(np.float32)
    ind = rng.randint(0, B, R).astype(np.int32)
    d_boxes, d_ind = cu(boxes), cu(ind)
    want = torch.zeros((R, C, ch, cw), device="cuda")
    torch.cuda.synchronize()
    fwd(img.data_ptr(), d_boxes.data_ptr(), d_ind.data_ptr(), R, B, H, W, ch, cw, C, ext, want.data_ptr(), None)
    torch.cuda.synchronize()
    rois = cu(np.concatenate([ind[:, None].astype(np.float32), boxes], 1))
    got = detbox.roi_align_forward([img], rois, [1.0], ch, cw, detbox.ROI_CROP_NORM, extrapolation=ext)
    err = (got - want).abs().max().item()
    assert err <= 1e-5 * max(1.0, want.abs().max().item()), err
    g = cu(rng.normal(0, 1, (R, C, ch, cw)).astype(np.float32))
    want_g = torch.zeros((B, C, H, W), device="cuda")
    torch.cuda.synchronize()
    bwd(g.data_ptr(), d_boxes.data_ptr(), d_ind.data_ptr(), R, B, H, W, ch, cw, C, want_g.data_ptr(), None)
    torch.cuda.synchronize()
    got_g = detbox.roi_align_backward(g, [tuple(img.shape)], rois, [1.0], detbox.ROI_CROP_NORM)[0]
    err = (got_g - want_g).abs().max().item()
    assert err <= 1e-5 * max(1.0, want_g.abs().max().item()), err
