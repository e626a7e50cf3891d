"""libs.boxes.nms_wrapper (reference libs/boxes/nms_wrapper.py)."""
import numpy as np

from ..nms import cpu_nms as _cpu_nms, gpu_nms as _gpu_nms


def nms_mode(force_cpu=False):
    """The comparator nms() applies: gpu_nms's `iou > thresh` in float32 (libs/nms/nms_kernel.cu:83) unless force_cpu,
    then cpu_nms's `iou >= thresh` in float64 (libs/nms/cpu_nms.pyx:65).  For callers that batch several nms() calls
    into one segmented device launch."""
    from ...ops import NMS_GE_F64, NMS_GT_F32
    return NMS_GE_F64 if force_cpu else NMS_GT_F32


def nms(dets, thresh, force_cpu=False):
    """nms_wrapper.py:12-19: [] if empty; force_cpu selects cpu_nms's '>=' comparator, else gpu_nms's '>'."""
    if dets.shape[0] == 0:
        return []
    if force_cpu:
        return _cpu_nms.cpu_nms(dets, thresh)
    return _gpu_nms.gpu_nms(dets, thresh, device_id=0)


def nms_wrapper(scores, boxes, threshold=0.7, class_sets=None):
    """nms_wrapper.py:21-49: per-class NMS (thr hard-coded 0.3 upstream, :40) then score filter."""
    num_class = scores.shape[1] if class_sets is None else len(class_sets)
    assert num_class * 4 == boxes.shape[1], 'Detection scores and boxes dont match %d vs %d' % (num_class, boxes.shape[1])
    class_sets = ['class_' + str(i) for i in range(0, num_class)] if class_sets is None else class_sets
    res = []
    for ind, cls in enumerate(class_sets[1:]):
        ind += 1
        dets = np.hstack((boxes[:, 4 * ind: 4 * (ind + 1)], scores[:, ind][:, np.newaxis])).astype(np.float32)
        keep = nms(dets, thresh=0.3)
        dets = dets[keep, :]
        dets = dets[np.where(dets[:, 4] > threshold)]
        res.append({'class': cls, 'dets': dets if dets.shape[0] > 0 else None})
    return res
