"""CPU-side checks of the drop-in boundary (no GPU, no compute calls): the mirror package exposes the
reference's module paths and names, argument validation raises the reference's exception types
before anything touches the device, and there is no CPU fallback behind any entry."""
import inspect

import numpy as np
import pytest
import torch

import detbox


@pytest.fixture(scope="module", autouse=True)
def libs():
    yield detbox.libs.install()
    detbox.libs.uninstall()          # tests/test_oracle_vs_ref.py imports the compiled reference as `libs`


def test_reference_module_paths_resolve():
    import libs.boxes.anchor, libs.boxes.bbox_transform, libs.boxes.cython_anchor, libs.boxes.cython_bbox  # noqa: E401,F401
    import libs.boxes.cython_nms, libs.boxes.nms_wrapper, libs.configs.config, libs.layers.anchor, libs.layers.box  # noqa: E401,F401
    import libs.layers.data_layer, libs.nets.model, libs.nms.cpu_nms, libs.nms.gpu_nms, libs.postprocessings.single_shot  # noqa: E401,F401
    from libs.layers.anchor_target.modules.anchor_target import AnchorTarget  # noqa: F401
    from libs.layers.nms.nms_layer import NMSLayer  # noqa: F401
    from libs.layers.roi_align.functions.roi_align import RoIAlignFunction  # noqa: F401
    from libs.layers.roi_align.modules.pyramid_roi_align import PyramidRoIAlign as P1  # noqa: F401
    from libs.layers.roi_align.modules.roi_align import RoIAlign, RoIAlignAvg, RoIAlignMax  # noqa: F401
    from libs.layers.roi_align_tf.crop_and_resize import CropAndResize, CropAndResizeFunction  # noqa: F401
    from libs.layers.roi_align_tf.pyramid_roi_align import PyramidRoIAlign, PyramidRoIAlign2  # noqa: F401
    from libs.layers.roi_align_tf.roi_align_tf import RoIAlign as R2  # noqa: F401


@pytest.mark.parametrize("path,name,params", [
    ("libs.boxes.cython_bbox", "bbox_overlaps", ["boxes", "query_boxes"]),
    ("libs.boxes.cython_bbox", "bbox_exclude_ignored_areas", ["boxes", "ignored_areas", "threshold"]),
    ("libs.boxes.anchor", "anchor_pyramid", ["strides", "ih", "iw", "anchor_scales", "anchor_ratios", "anchor_base", "anchor_shift"]),
    ("libs.boxes.anchor", "anchors_plane", ["height", "width", "stride", "scales", "ratios", "base", "anchor_shift"]),
    ("libs.boxes.bbox_transform", "bbox_transform_inv", ["boxes", "deltas"]),
    ("libs.boxes.bbox_transform", "clip_boxes", ["boxes", "im_shape"]),
    ("libs.boxes.nms_wrapper", "nms", ["dets", "thresh", "force_cpu"]),
    ("libs.nms.cpu_nms", "cpu_nms", ["dets", "thresh"]),
    ("libs.nms.gpu_nms", "gpu_nms", ["dets", "thresh", "device_id"]),
    ("libs.layers.anchor", "encode", ["gt_boxes", "all_anchors"]),
    ("libs.layers.anchor", "decode", ["boxes", "scores", "all_anchors", "ih", "iw", "num_classes"]),
    ("libs.layers.box", "apply_nms", ["boxes", "scores", "overlap_threshold"]),
    ("libs.layers.box", "decoding_box", ["deltas", "anchors", "box_encoding"]),
    ("libs.layers.box", "encoding_box", ["gt_boxes", "anchors", "box_encoding"]),
    ("libs.layers.data_layer", "compute_rpn_targets_in_batch", ["gt_boxes_list", "anchors"]),
])
def test_signatures_follow_the_reference(path, name, params):
    import importlib
    fn = getattr(importlib.import_module(path), name)
    assert list(inspect.signature(fn).parameters)[:len(params)] == params


def test_argument_validation_and_no_cpu_fallback():
    import libs.boxes.cython_bbox as cython_bbox
    from libs.boxes.cython_nms import nms_new
    from libs.nms.cpu_nms import cpu_nms
    b = np.zeros((3, 4), np.float64)
    with pytest.raises(ValueError):
        cython_bbox.bbox_overlaps(b.astype(np.float32), b)          # Cython: Buffer dtype mismatch
    with pytest.raises(TypeError):
        cython_bbox.bbox_overlaps([[0, 0, 1, 1]], b)
    with pytest.raises(TypeError):
        cpu_nms(np.zeros((2, 5), np.float32), 1)                    # thresh must be a Python float
    with pytest.raises(TypeError):
        nms_new(np.zeros((2, 5), np.float32), 1)
    with pytest.raises(ValueError):
        nms_new(np.zeros((2, 5), np.float64), 0.5)
    assert cython_bbox.bbox_overlaps(np.zeros((0, 4)), b).shape == (0, 3)
    if not torch.cuda.is_available():
        with pytest.raises(detbox.DetboxError):
            cython_bbox.bbox_overlaps(b, b)                          # no CPU path: fails loudly
        with pytest.raises(detbox.DetboxError):
            cpu_nms(np.zeros((2, 5), np.float32), 0.5)
        with pytest.raises(detbox.DetboxError):
            nms_new(np.zeros((2, 5), np.float32), 0.5)


def test_config_defaults_are_the_references():
    import libs.configs.config as cfg
    assert (cfg.rpn_bg_threshold, cfg.rpn_fg_threshold, cfg.anchor_base) == (0.4, 0.6, 16)
    assert cfg.strides == (8, 16, 32, 64, 128) and cfg.anchor_ratios == [0.5, 1, 2.0]
    assert cfg.rpn_sample_strategy == 'simple' and cfg.ignored_area_intersection_fraction == 0.3
