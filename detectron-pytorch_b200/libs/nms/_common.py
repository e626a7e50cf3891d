"""Shared implementation of the NumPy-facing NMS entry points: host dets in, Python list out,
through the host-buffer C entry point detbox_nms_host (copies inside, like the reference's _nms)."""
import ctypes

import numpy as np
import torch

from ..._lib import check, lib, DetboxError


def nms_host(dets, thresh, cmp_mode, device_id=None):
    if not isinstance(dets, np.ndarray):
        raise TypeError("Argument 'dets' has incorrect type (expected numpy.ndarray, got %s)" % type(dets).__name__)
    if dets.ndim != 2 or dets.dtype != np.float32:
        raise ValueError("Buffer dtype mismatch, expected 'float32' [n,5] but got '%s' ndim %d" % (dets.dtype.name, dets.ndim))
    if not isinstance(thresh, float):                       # the generated Cython wrapper checks the exact type
        raise TypeError("Argument 'thresh' has incorrect type (expected float, got %s)" % type(thresh).__name__)
    if not torch.cuda.is_available():
        raise DetboxError("detbox has no CPU path: a CUDA device is required")
    n = dets.shape[0]
    if n == 0:
        return []
    d = np.ascontiguousarray(dets[:, :5])
    keep = np.empty((n,), np.int32)
    num = ctypes.c_int(0)
    dev = torch.cuda.current_device() if device_id is None else int(device_id)
    check(lib().detbox_nms_host(d.ctypes.data, n, float(thresh), int(cmp_mode), keep.ctypes.data, ctypes.addressof(num), dev), "nms")
    return [int(i) for i in keep[:num.value]]
