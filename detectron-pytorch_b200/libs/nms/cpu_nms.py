"""libs.nms.cpu_nms (reference libs/nms/cpu_nms.pyx).  The NAME is the reference's; the work runs on
the GPU with cpu_nms's comparator: suppress when (double)iou >= thresh (cpu_nms.pyx:65)."""
from ... import ops
from ._common import nms_host


def cpu_nms(dets, thresh):
    """dets float32 [n,5] = (x1,y1,x2,y2,score), thresh a Python float -> list of kept indices into
    the UNSORTED input, by descending score (ties: lower index first)."""
    return nms_host(dets, thresh, ops.NMS_GE_F64)
