"""libs.boxes.cython_bbox (reference libs/boxes/bbox.pyx) — float64 IoU matrices on the device."""
import numpy as np

from ... import ops
from ._np import check_buffer, to_dev


def bbox_overlaps(boxes, query_boxes):
    """bbox.pyx:15-55: boxes (N,4) float64, query_boxes (K,4) float64 -> (N,K) float64 IoU (+1 convention)."""
    check_buffer(boxes, np.float64, 2, "boxes"); check_buffer(query_boxes, np.float64, 2, "query_boxes")
    if boxes.shape[0] == 0 or query_boxes.shape[0] == 0:
        return np.zeros((boxes.shape[0], query_boxes.shape[0]), np.float64)
    return ops.bbox_overlaps(to_dev(boxes, np.float64), to_dev(query_boxes, np.float64)).cpu().numpy()


def bbox_intersections(boxes, query_boxes):
    """bbox.pyx:57-94: intersection area / area(query box)."""
    check_buffer(boxes, np.float64, 2, "boxes"); check_buffer(query_boxes, np.float64, 2, "query_boxes")
    if boxes.shape[0] == 0 or query_boxes.shape[0] == 0:
        return np.zeros((boxes.shape[0], query_boxes.shape[0]), np.float64)
    return ops.bbox_intersections(to_dev(boxes, np.float64), to_dev(query_boxes, np.float64)).cpu().numpy()


def bbox_exclude_ignored_areas(boxes, ignored_areas, threshold):
    """bbox.pyx:96-135: int64 [N] flag, 1 where inter/area(box) > threshold for any ignored area."""
    check_buffer(boxes, np.float64, 2, "boxes"); check_buffer(ignored_areas, np.float64, 2, "ignored_areas")
    if boxes.shape[0] == 0 or ignored_areas.shape[0] == 0:
        return np.zeros((boxes.shape[0],), np.int64)
    return ops.bbox_exclude_ignored_areas(to_dev(boxes, np.float64), to_dev(ignored_areas, np.float64), float(threshold)).cpu().numpy()
