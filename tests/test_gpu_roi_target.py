"""GPU parity for RoI -> GT matching, RoI sampling and RoI targets through the C ABI:
bit-exact labels / match indices / matched boxes / sampled rows against the golden vectors recorded from the
reference's libs/layers/box.py and against the oracle on seeded inputs; RoI-target labels bit-exact and
deltas within 1e-5 against the oracle, and against the reference's own CUDA kernel in test_gpu_ref_cuda.py."""
import ctypes
import os

import numpy as np
import pytest
import torch

import detbox
from oracle import oracle as O
from oracle.make_golden import synth_rois

pytestmark = pytest.mark.gpu


def mirror():
    import importlib
    return importlib.import_module(detbox.libs.__name__ + ".layers.box"), importlib.import_module(detbox.libs.__name__ + ".layers.roi_target")


def cu(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def _roi_case(g, c):
    ngt = g["c%d_ngt" % c]
    gts = np.split(g["c%d_gt" % c], np.cumsum(ngt)[:-1]) if ngt.size else []
    return g["c%d_boxes" % c], g["c%d_ids" % c], [x.reshape(-1, 5) for x in gts]


@pytest.mark.parametrize("case", range(6))
def test_golden_matching_and_sampling(golden, case):
    B, _ = mirror()
    g = golden("roi_sampling")
    b, ids, gts = _roi_case(g, case)
    lab, match, mbox = B.matching_box(cu(b), cu(ids), gts, 0.4, 0.6)
    assert lab.dtype == torch.int64 and match.dtype == torch.int32 and mbox.dtype == torch.float32 and lab.is_cuda
    np.testing.assert_array_equal(lab.cpu().numpy(), g["c%d_match_labels" % case])
    np.testing.assert_array_equal(match.cpu().numpy(), g["c%d_match_inds" % case])
    np.testing.assert_array_equal(mbox.cpu().numpy(), g["c%d_match_boxes" % case])
    sb, sl, si = B.sample_rois(cu(b), cu(ids), gts, 0.5, 16, 0.25, 0.2, rng=np.random.RandomState(100 + case))
    assert sl.dtype == torch.int64 and si.dtype == torch.int64 and sb.is_cuda
    np.testing.assert_array_equal(sb.cpu().numpy(), g["c%d_sample_boxes" % case])
    np.testing.assert_array_equal(sl.cpu().numpy(), g["c%d_sample_labels" % case])
    np.testing.assert_array_equal(si.cpu().numpy(), g["c%d_sample_ids" % case])


@pytest.mark.parametrize("seed,n,Bn", [(7, 5000, 8), (8, 20000, 16), (9, 1, 1), (10, 3000, 64)])
def test_matching_and_sampling_vs_oracle(seed, n, Bn):
    """MaskRCNN training sizes: 2000-6000 proposals x 8-16 images (maskrcnn.py:91-97)."""
    B, _ = mirror()
    b, ids, gts = synth_rois(seed, n=n, B=Bn, sorted_ids=seed % 2 == 1)
    got = B.matching_box(cu(b), cu(ids), gts)
    want = O.matching_box(b, ids, gts)
    for gt_, w in zip(got, want):
        np.testing.assert_array_equal(gt_.cpu().numpy(), w)
    got = B.sample_rois(cu(b), cu(ids), gts, rois_per_image=64, rng=np.random.RandomState(seed))
    want = O.sample_rois(b, ids, gts, rois_per_image=64, rng=np.random.RandomState(seed))
    for gt_, w in zip(got, want):
        np.testing.assert_array_equal(gt_.cpu().numpy(), w)


def test_sampling_edge_cases():
    B, _ = mirror()
    # no GT anywhere: every (large enough) proposal is background
    b, ids, _ = synth_rois(3, n=200, B=2)
    gts = [np.zeros((0, 5), np.float32)] * 2
    got = B.sample_rois(cu(b), cu(ids), gts, rois_per_image=16, rng=np.random.RandomState(0))
    want = O.sample_rois(b, ids, gts, rois_per_image=16, rng=np.random.RandomState(0))
    for g_, w in zip(got, want):
        np.testing.assert_array_equal(g_.cpu().numpy(), w)
    # nothing survives (all proposals tiny, no GT): the guard row (box.py:282-286)
    tiny = np.asarray([[5, 5, 6, 6], [9, 9, 11, 12]], np.float32)
    sb, sl, si = B.sample_rois(cu(tiny), cu(np.zeros(2, np.int64)), [np.zeros((0, 5), np.float32)], rng=np.random.RandomState(0))
    np.testing.assert_array_equal(sb.cpu().numpy(), tiny[:1]); assert sl.tolist() == [-1] and si.tolist() == [0]
    with pytest.raises(detbox.DetboxError):
        B.sample_rois(torch.zeros((2, 4)), torch.zeros(2, dtype=torch.int64), [np.zeros((0, 5), np.float32)])


@pytest.mark.parametrize("seed,R,G,Bn", [(0, 300, 12, 3), (1, 4096, 80, 8), (2, 10, 0, 1), (3, 700, 5, 1)])
def test_roi_target_vs_oracle_and_reference_kernel(seed, R, G, Bn):
    _, T = mirror()
    rng = np.random.RandomState(seed)
    wh = np.exp(rng.uniform(np.log(8), np.log(200), (R, 2))); c = rng.uniform(0, 400, (R, 2))
    rois = np.concatenate([c - wh / 2, c + wh / 2], 1).astype(np.float32)
    rid = rng.randint(0, Bn, R).astype(np.int64)
    gwh = np.exp(rng.uniform(np.log(10), np.log(200), (G, 2))); gc = rng.uniform(0, 400, (G, 2))
    gt = np.concatenate([gc - gwh / 2, gc + gwh / 2, rng.choice([-1, 0, 1, 2, 3], G)[:, None]], 1).astype(np.float32)
    gid = rng.randint(0, Bn, G).astype(np.int64)
    for j in range(G):                                          # plant overlapping rois (and exact duplicates: IoU 1 ties)
        for k in rng.randint(0, R, 4):
            rois[k] = gt[j, :4] + (rng.uniform(-4, 4, 4) if rng.rand() < 0.8 else np.zeros(4)).astype(np.float32); rid[k] = gid[j]
    labels, deltas, bbw = T.RoITarget(0.5)(cu(rois), cu(rid), cu(gt), cu(gid))
    assert labels.dtype == torch.int64
    wl, wd, ww = O.roi_target(rois, rid, gt, gid, 0.5)
    np.testing.assert_array_equal(labels.cpu().numpy(), wl)
    np.testing.assert_array_equal(bbw.cpu().numpy(), ww)
    np.testing.assert_allclose(deltas.cpu().numpy(), wd, rtol=1e-5, atol=1e-5)
    # the reference's kernel, compiled unmodified (oracle/build_ref.py --cuda), on the same device buffers
    path = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle", "_ref", "libref_cuda_roi_target.so")
    if not os.path.exists(path) or G == 0:
        return
    L = ctypes.CDLL(path)
    vp, ci, cf = ctypes.c_void_p, ctypes.c_int, ctypes.c_float
    L.ROITargetForwardLaucher.argtypes = [vp, vp, vp, vp, cf, cf, ci, ci, vp, vp, vp, vp]
    d_r, d_ri, d_g, d_gi = cu(rois), cu(rid), cu(gt), cu(gid)
    rl = torch.full((R,), -1, dtype=torch.int64, device="cuda"); rd = torch.zeros((R, 4), device="cuda"); rw = torch.zeros((R, 4), device="cuda")
    torch.cuda.synchronize()
    L.ROITargetForwardLaucher(d_r.data_ptr(), d_ri.data_ptr(), d_g.data_ptr(), d_gi.data_ptr(), 0.0, 0.5, R, G,
                              rl.data_ptr(), rd.data_ptr(), rw.data_ptr(), None)
    torch.cuda.synchronize()
    assert torch.equal(labels, rl) and torch.equal(bbw, rw)
    assert (deltas - rd).abs().max().item() <= 1e-5 * max(1.0, rd.abs().max().item())


def test_roi_target_reference_fixture():
    """The reference's own smoke fixture (libs/layers/roi_target/modules/roi_target_test.py:15-40: 10 RoIs in two
    images, 6 GT of class 0 / -1 / 3 / 4, RoITarget(0.55)); upstream only prints.  Checked against the oracle and,
    when present, the reference kernel compiled unmodified."""
    _, T = mirror()
    rois = np.asarray([[120, 120, 140, 150], [95, 100, 120, 130], [120, 120, 125, 125], [40, 50, 50, 60], [95, 100, 120, 130],
                       [0, 0, 9, 9], [1, 1, 8, 8], [10, 10, 20, 20], [16, 10, 26, 20], [70, 60, 80, 90]], np.float32)
    rid = np.asarray([0, 0, 0, 0, 1, 1, 1, 1, 1, 0], np.int64)
    gt = np.asarray([[121, 120, 140, 150, 0], [100, 100, 120, 130, 0], [121, 120, 140, 150, -1], [100, 100, 120, 130, -1],
                     [1, 1, 8, 8, 3], [13, 10, 23, 20, 4]], np.float32)
    gid = np.asarray([0, 0, 1, 1, 1, 1], np.int64)
    labels, deltas, bbw = T.RoITarget(0.55)(cu(rois), cu(rid), cu(gt), cu(gid))
    wl, wd, ww = O.roi_target(rois, rid, gt, gid, 0.55)
    np.testing.assert_array_equal(labels.cpu().numpy(), wl)
    np.testing.assert_array_equal(bbw.cpu().numpy(), ww)
    np.testing.assert_allclose(deltas.cpu().numpy(), wd, rtol=1e-5, atol=1e-6)
    # known answers: RoIs 0-2 and 4 sit inside class-0 / class -1 GT (inter / area(roi) >= 0.55 -> label -1: the kernel scores
    # ignore boxes by coverage of the RoI), RoI 6 IS the GT (1,1,8,8,3), RoI 5 contains it, RoIs 7 / 8 flank the class-4 GT
    assert labels.cpu().tolist() == [-1, -1, -1, 0, -1, 3, 3, 4, 4, 0]
    assert bbw.cpu().numpy()[:, 0].tolist() == [0, 0, 0, 0, 0, 1, 1, 1, 1, 0]
    assert np.abs(deltas.cpu().numpy()[6]).max() == 0
    np.testing.assert_allclose(deltas.cpu().numpy()[7:9, 0], [30.0 / 11, -30.0 / 11], rtol=1e-6)
