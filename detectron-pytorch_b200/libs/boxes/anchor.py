"""libs.boxes.anchor (reference libs/boxes/anchor.py) — anchor generation on the device.
Returns NEW float64 NumPy arrays like upstream; the level grids are expanded by
anchor_pyramid_kernel (anchors.cu), the <= 30 cell anchors by detbox_base_anchors_host."""
import numpy as np

from ... import ops
from . import cython_anchor
from ._np import require_cuda


def anchors(scales=[2., 4., 8., 16., 32.], ratios=[0.5, 1, 2.0], base=16):
    """anchor.py:101-103: scales are cast to float32 first (as upstream), result float64 [R*S,4],
    ratio-major / scale-minor."""
    return ops.base_anchors(np.asarray(scales, np.float32), ratios, base)


def anchor_one_plane(height, width, stride, anchor_scales=[8., 16., 32], anchor_ratios=[0.5, 1.0, 2.0],
                     anchor_base=16, anchor_shifts=[[0.0, 0.0]]):
    """anchor.py:47-70 -> float64 [H*W*NS*A, 4], order y -> x -> (shift, ratio, scale)."""
    require_cuda()
    import torch
    out = ops.anchor_pyramid([stride], height * stride, width * stride, [list(anchor_scales)], anchor_ratios, anchor_base,
                             anchor_shifts, dtype=torch.float64)
    return out.cpu().numpy()


def anchor_pyramid(strides, ih, iw, anchor_scales, anchor_ratios, anchor_base, anchor_shift=[[0.0, 0.0]]):
    """anchor.py:30-44: all levels, fine -> coarse, float64 [sum_l H_l*W_l*A, 4]."""
    require_cuda()
    import torch
    return ops.anchor_pyramid(list(strides), ih, iw, anchor_scales, anchor_ratios, anchor_base, anchor_shift,
                              dtype=torch.float64).cpu().numpy()


anchor_pyramid_old = anchor_pyramid        # anchor.py:9-27 produces the identical array


def anchors_plane(height, width, stride=1.0, scales=[2., 4., 8., 16., 32.], ratios=[0.5, 1, 2.0], base=16,
                  anchor_shift=[[0.0, 0.0]]):
    """anchor.py:106-121 -> float64 [H, W, NS*A, 4]."""
    anc = anchors(scales, ratios, base).astype(np.float64)
    return cython_anchor.anchors_plane(height, width, stride, anc, np.array(anchor_shift, dtype=np.float64))
