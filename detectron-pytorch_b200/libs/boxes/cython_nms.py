"""libs.boxes.cython_nms (reference libs/boxes/nms.pyx).  `nms` (:17-68) is identical to cpu_nms.
The variants nms_new (:70-123) and nms_multianchor (:125-195) are SURVEY §8f "next" rows and are
not built yet: they raise instead of silently falling back to anything."""
from ..nms.cpu_nms import cpu_nms as nms  # noqa: F401


def nms_new(dets, thresh):
    raise NotImplementedError("cython_nms.nms_new is a SURVEY §8f next-row; not implemented on the device yet")


def nms_multianchor(dets, thresh):
    raise NotImplementedError("cython_nms.nms_multianchor is a SURVEY §8f next-row; not implemented on the device yet")
