"""libs.nms.py_cpu_nms (reference libs/nms/py_cpu_nms.py:10-38): the '>' comparator baseline."""
from ... import ops
from ._common import nms_host


def py_cpu_nms(dets, thresh):
    return nms_host(dets, float(thresh), ops.NMS_GT_F32)
