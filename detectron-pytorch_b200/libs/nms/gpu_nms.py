"""libs.nms.gpu_nms (reference libs/nms/gpu_nms.pyx + nms_kernel.cu): comparator iou > (float)thresh."""
from ... import ops
from ._common import nms_host


def gpu_nms(dets, thresh, device_id=0):
    """gpu_nms.pyx:16-31.  Unlike upstream nothing is allocated per call and the mask never leaves the device."""
    return nms_host(dets, float(thresh), ops.NMS_GT_F32, device_id)
