"""Shared test helpers."""
import numpy as np

RTOL, ATOL = 1e-5, 1e-6


def assert_boxes_close(got, want, rtol=RTOL, extra_scale=None):
    """Decoded boxes agree within `rtol` RELATIVE TO THE BOX'S OWN COORDINATE MAGNITUDE
    (max |coord| of the row, at least 1 px).  A per-component relative bound is meaningless
    here: x1 = ctr - 0.5*w cancels, so a 1-ulp difference in expf (CUDA / glibc / NumPy's SIMD
    exp are each within 1 ulp, not identical) shows up as 1 ulp of `ctr`, not of `x1`."""
    got, want = np.asarray(got, np.float64), np.asarray(want, np.float64)
    assert got.shape == want.shape
    if want.size == 0:
        return
    scale = np.maximum(np.abs(want).reshape(want.shape[0], -1).max(axis=1), 1.0)
    if extra_scale is not None:
        scale = np.maximum(scale, extra_scale)
    err = np.abs(got - want).reshape(want.shape[0], -1).max(axis=1)
    bad = np.nonzero(~(err <= rtol * scale))[0]
    assert bad.size == 0, "rows %s differ: got %s want %s" % (bad[:5], got[bad[:5]], want[bad[:5]])
