"""NumPy <-> device helpers shared by the NumPy-facing mirrors."""
import numpy as np
import torch

from ..._lib import DetboxError


def require_cuda():
    if not torch.cuda.is_available():
        raise DetboxError("detbox has no CPU path: a CUDA device is required")


def to_dev(a, dtype):
    require_cuda()
    return torch.from_numpy(np.ascontiguousarray(a, dtype=dtype)).cuda()


def check_buffer(a, dtype, ndim, name):
    """The Cython signature check: np.ndarray[DTYPE_t, ndim=2] (bbox.pyx:16-17)."""
    if not isinstance(a, np.ndarray):
        raise TypeError("Argument '%s' has incorrect type (expected numpy.ndarray, got %s)" % (name, type(a).__name__))
    if a.ndim != ndim:
        raise ValueError("Buffer has wrong number of dimensions (expected %d, got %d)" % (ndim, a.ndim))
    if a.dtype != dtype:
        raise ValueError("Buffer dtype mismatch, expected '%s' but got '%s'" % (np.dtype(dtype).name, a.dtype.name))
