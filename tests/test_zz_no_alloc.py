"""detbox_nms_static (csrc/no_alloc.cu): the NMS entry point that allocates nothing — its scratch is a file-scope
__device__ array inside the library (no counterpart in the reference).

Called through the C ABI with raw pointers, on TWO calls in a row that share the array (a large single-segment problem,
then a smaller segmented one, then the first again), each compared with the existing detbox_nms entry point on the same
inputs: keep indices and counts are integers, so the bar is the one every NMS test here uses — bit-exact.  The file is
named so that pytest collects it LAST; it passes alone in a fresh process and after the
other tests."""
import ctypes

import numpy as np
import pytest
import torch

import detbox

pytestmark = pytest.mark.gpu
MAX_BOXES, MAX_SEGMENTS = 4096, 64          # DETBOX_NMS_STATIC_MAX_BOXES / _SEGMENTS (include/detbox.h)


def boxes_scores(seed, n):
    rng = np.random.RandomState(seed)
    h = np.exp(rng.uniform(np.log(12), np.log(264), n)); w = h / 2.44
    cx, cy = rng.uniform(0, 700, n), rng.uniform(0, 350, n)             # dense: plenty of suppression
    b = np.stack([cx - w / 2, cy - h / 2, cx + w / 2, cy + h / 2], 1).astype(np.float32)
    s = rng.permutation(np.linspace(0, 1, n)).astype(np.float32)
    return torch.from_numpy(b).cuda(), torch.from_numpy(s).cuda()


def run_both(b, s, seg, n_seg, thr, mode):
    """(keep, n_keep) from the existing entry point (caller-provided workspace) and from the static one."""
    L = detbox.lib()
    n = b.shape[0]
    st = torch.cuda.current_stream().cuda_stream
    segp = seg.data_ptr() if seg is not None else None
    ws = torch.empty(L.detbox_nms_workspace_bytes(n, n_seg), dtype=torch.uint8, device="cuda")
    out = []
    for static in (False, True):
        keep = torch.full((n,), -7, dtype=torch.int64, device="cuda")
        nk = torch.full((max(n_seg, 1) + 1,), -7, dtype=torch.int32, device="cuda")
        if static:
            rc = L.detbox_nms_static(b.data_ptr(), 4, s.data_ptr(), n, segp, n_seg, thr, mode, keep.data_ptr(), nk.data_ptr(), st)
        else:
            rc = L.detbox_nms(b.data_ptr(), 4, s.data_ptr(), n, segp, n_seg, thr, mode, keep.data_ptr(), nk.data_ptr(),
                              ws.data_ptr(), ws.numel(), st)
        assert rc == 0, L.detbox_error_string(rc)
        torch.cuda.synchronize()
        nkh = nk.cpu().numpy()
        out.append((keep.cpu().numpy()[:int(nkh[-1])], nkh))
    return out


def test_static_scratch_nms_matches_existing_entry_on_consecutive_calls():
    b1, s1 = boxes_scores(1, MAX_BOXES)                                  # call 1: the maximum size, one segment
    (k_ref, n_ref), (k_st, n_st) = run_both(b1, s1, None, 1, 0.5, detbox.NMS_GT_F32)
    assert 0 < k_ref.size < MAX_BOXES
    np.testing.assert_array_equal(k_st, k_ref); np.testing.assert_array_equal(n_st, n_ref)
    b2, s2 = boxes_scores(2, 1500)                                       # call 2 right after: smaller, 5 segments, other comparator
    seg = torch.from_numpy(np.random.RandomState(3).randint(0, 5, 1500).astype(np.int32)).cuda()
    (k_ref2, n_ref2), (k_st2, n_st2) = run_both(b2, s2, seg, 5, 0.3, detbox.NMS_GE_F64)
    assert 0 < k_ref2.size < 1500 and n_ref2[:5].sum() == n_ref2[5]
    np.testing.assert_array_equal(k_st2, k_ref2); np.testing.assert_array_equal(n_st2, n_ref2)
    (_, _), (k_st3, n_st3) = run_both(b1, s1, None, 1, 0.5, detbox.NMS_GT_F32)   # call 1 again over call 2's leftovers
    np.testing.assert_array_equal(k_st3, k_ref); np.testing.assert_array_equal(n_st3, n_ref)


def test_static_scratch_nms_rejects_sizes_beyond_the_limit():
    L = detbox.lib()
    b, s = boxes_scores(4, MAX_BOXES + 1)
    keep = torch.full((MAX_BOXES + 1,), -7, dtype=torch.int64, device="cuda")
    nk = torch.full((MAX_SEGMENTS + 3,), -7, dtype=torch.int32, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    rc = L.detbox_nms_static(b.data_ptr(), 4, s.data_ptr(), MAX_BOXES + 1, None, 1, 0.5, detbox.NMS_GT_F32, keep.data_ptr(), nk.data_ptr(), st)
    assert rc == -4 and b"limit" in L.detbox_error_string(rc)            # DETBOX_ERR_UNSUPPORTED
    seg = torch.zeros(100, dtype=torch.int32, device="cuda")
    rc = L.detbox_nms_static(b.data_ptr(), 4, s.data_ptr(), 100, seg.data_ptr(), MAX_SEGMENTS + 1, 0.5, detbox.NMS_GT_F32,
                             keep.data_ptr(), nk.data_ptr(), st)
    assert rc == -4
    torch.cuda.synchronize()
    assert (keep == -7).all() and (nk == -7).all()                       # nothing was enqueued
