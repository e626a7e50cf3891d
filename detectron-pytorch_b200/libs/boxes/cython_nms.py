"""libs.boxes.cython_nms (reference libs/boxes/nms.pyx).  `nms` (:17-68) is identical to cpu_nms.
`nms_new` (:70-123) and `nms_multianchor` (:125-195) run on the same device pipeline as every other NMS
here (suppressor bitmask + fixed-point resolve) with their own pair rule."""
import numpy as np
import torch

from ... import ops, ops_post
from ..._lib import DetboxError
from ..nms.cpu_nms import cpu_nms as nms  # noqa: F401


def _variant(dets, thresh, variant, want_clusters):
    if not isinstance(dets, np.ndarray):
        raise TypeError("Argument 'dets' has incorrect type (expected numpy.ndarray, got %s)" % type(dets).__name__)
    if dets.ndim != 2 or dets.dtype != np.float32:
        raise ValueError("Buffer dtype mismatch, expected 'float32' [n,5] but got '%s' ndim %d" % (dets.dtype.name, dets.ndim))
    if not isinstance(thresh, float):
        raise TypeError("Argument 'thresh' has incorrect type (expected float, got %s)" % type(thresh).__name__)
    if not torch.cuda.is_available():
        raise DetboxError("detbox has no CPU path: a CUDA device is required")
    n = dets.shape[0]
    if n == 0:
        return [], np.arange(0)
    d = torch.from_numpy(np.ascontiguousarray(dets[:, :5])).cuda()
    out = ops_post.nms(d, d[:, 4], float(thresh), ops.NMS_GE_F64, variant=variant, want_clusters=want_clusters)
    keep = out[0][:int(out[1][-1].item())].cpu().numpy()
    return [int(i) for i in keep], (out[2].cpu().numpy() if want_clusters else None)


def nms_new(dets, thresh):
    """nms.pyx:70-123: IoU >= thresh, or one box holds > 95 % of the other's area."""
    return _variant(dets, thresh, "new", False)[0]


def nms_multianchor(dets, thresh):
    """nms.pyx:125-195 -> (keep list, clusters int64 [n]): (inter + 0.1) / (union + 0.1) >= thresh, or the centres
    are closer than 0.2 * min(intersection w, h); clusters[j] = the kept box that absorbed j."""
    return _variant(dets, thresh, "multianchor", True)
