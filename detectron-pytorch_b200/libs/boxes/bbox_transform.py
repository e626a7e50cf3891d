"""libs.boxes.bbox_transform (reference libs/boxes/bbox_transform.py) — NumPy in, NumPy out; the
arithmetic runs in float32 on the device in the reference's operation order (box_math.cuh).
float64 inputs are accepted and returned as float64 (computed in float32: the hot path only ever
passes float32, libs/layers/anchor.py:179-189)."""
import numpy as np
import torch

from ... import ops
from ._np import to_dev


def _rows(a):
    a = np.asarray(a)
    return a, (a.dtype if a.dtype in (np.float32, np.float64) else np.float32)


def bbox_transform(ex_rois, gt_rois):
    """bbox_transform.py:11-41: (dx,dy)/w/0.1, log(gw/ew)/0.2 -> [n,4]."""
    ex, dt = _rows(ex_rois)
    if ex.shape[0] == 0:
        return np.zeros((0, 4), dt)
    return ops.box_encode(to_dev(ex[:, :4], np.float32), to_dev(np.asarray(gt_rois)[:, :4], np.float32), "fastrcnn").cpu().numpy().astype(dt, copy=False)


def _decode(boxes, deltas, enc):
    b, _ = _rows(boxes)
    d, dt = _rows(deltas)
    if b.shape[0] == 0:                                             # bbox_transform.py:44-45
        return np.zeros((0, d.shape[1]), dtype=d.dtype)
    k = d.shape[1] // 4                                            # deltas[:, 0::4] k-class layout
    anc = to_dev(b[:, :4], np.float32)
    if k > 1:
        anc = anc.repeat_interleave(k, dim=0)
    out = ops.box_decode(anc, to_dev(d, np.float32).view(-1, 4), enc)
    return out.view(b.shape[0], 4 * k).cpu().numpy().astype(dt, copy=False)


def bbox_transform_inv(boxes, deltas):
    """bbox_transform.py:43-76: x0.1 / x0.2, exp(dw + log(w)), -1 on x2 / y2; no clip."""
    return _decode(boxes, deltas, "fastrcnn")


def clip_boxes(boxes, im_shape):
    """bbox_transform.py:78-91: clip IN PLACE to [0, W-1] x [0, H-1] and return the same array."""
    if boxes.shape[0] == 0:
        return boxes
    t = to_dev(boxes, np.float32)
    ops.clip_boxes_(t, im_shape)
    boxes[...] = t.cpu().numpy().astype(boxes.dtype, copy=False)
    return boxes


def bbox_transform_linear(anchors, gt_boxes, alpha=10.0):
    """bbox_transform.py:93-116: (gt - anchor) / (w | h) * alpha.  alpha = 10 (every upstream call site) is one kernel;
    any other alpha is the 'simple' encode ((gt - anchor) / size, the same first two operations) times alpha in
    float32 — the reference's own operation order, so both are bit-exact for float32 inputs."""
    a, dt = _rows(anchors)
    if a.shape[0] == 0:
        return np.zeros((0, 4), dt)
    ten = float(alpha) == 10.0
    t = ops.box_encode(to_dev(a[:, :4], np.float32), to_dev(np.asarray(gt_boxes)[:, :4], np.float32), "linear" if ten else "simple").cpu().numpy()
    if not ten:
        t = t * np.float32(alpha)
    return t.astype(dt, copy=False)


def bbox_transform_inv_linear(anchors, deltas, alpha=10.0):
    """bbox_transform.py:118-144: deltas * (w | h) / alpha + anchor.  alpha = 10 is one kernel in the reference's
    operation order; any other alpha runs the 'simple' decode on deltas / alpha — (d / alpha) * w instead of
    (d * w) / alpha, one rounding apart, inside the 1e-5 relative bar for decoded boxes."""
    if float(alpha) == 10.0:
        return _decode(anchors, deltas, "linear")
    d = np.asarray(deltas)
    return _decode(anchors, (d.astype(np.float32) / np.float32(alpha)).astype(d.dtype if d.dtype in (np.float32, np.float64) else np.float32), "simple")
