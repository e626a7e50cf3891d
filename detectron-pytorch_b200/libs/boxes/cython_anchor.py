"""libs.boxes.cython_anchor (reference libs/boxes/cython_anchor.pyx)."""
import ctypes

import numpy as np
import torch

from ... import ops
from ..._lib import check, lib
from ._np import check_buffer, require_cuda


def anchors_plane(height, width, stride, anchors_base, anchor_shift):
    """cython_anchor.pyx:14-48: (height, width, A*NS, 4) float64; entry [y,x,k+ns*A] =
    anchors_base[k] + (x*stride, y*stride) + stride*anchor_shift[ns].  `stride` is an int upstream."""
    check_buffer(anchors_base, np.float64, 2, "anchors_base"); check_buffer(anchor_shift, np.float64, 2, "anchor_shift")
    require_cuda()
    stride = int(stride)
    sh = np.hstack((anchor_shift[:, :2], anchor_shift[:, :2])) * stride
    cell = np.vstack([anchors_base + s for s in sh])
    cell_dev = torch.from_numpy(np.ascontiguousarray(cell)).cuda()
    out = torch.empty((int(height) * int(width) * cell.shape[0], 4), dtype=torch.float64, device="cuda")
    one = lambda v: (ctypes.c_int * 1)(int(v))
    check(lib().detbox_anchor_pyramid(cell_dev.data_ptr(), 1, one(cell.shape[0]), one(height), one(width), one(stride),
                                      out.data_ptr(), 1, ops._stream()), "anchors_plane")
    ops._count(1)
    return out.cpu().numpy().reshape(int(height), int(width), cell.shape[0], 4)
