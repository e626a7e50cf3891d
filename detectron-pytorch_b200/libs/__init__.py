"""Mirror of the reference's module paths for the box-and-RoI hot path.

Every callable here has the reference's name, argument order, dtypes, return types and index
conventions, and runs on the hand-written CUDA kernels of libdetbox.so — there is no CPU
implementation behind any of them.  NumPy-facing functions (the reference's Cython / NumPy layer)
copy their arguments to the current CUDA device and the results back; torch-facing ones stay on
the device.

`install()` aliases this package as the top-level name ``libs`` so that reference code doing
``import libs.boxes.cython_bbox as cython_bbox`` or ``from libs.layers.box import apply_nms``
runs unmodified against it.
"""
import importlib
import sys

_SUBMODULES = [
    "configs.config", "boxes.cython_bbox", "boxes.cython_anchor", "boxes.anchor", "boxes.bbox_transform",
    "boxes.cython_nms", "boxes.nms_wrapper", "nms.cpu_nms", "nms.gpu_nms", "nms.py_cpu_nms",
    "layers.anchor", "layers.box", "layers.data_layer", "layers.anchor_target", "layers.nms_layer",
    "layers.roi_align_tf", "layers.roi_align", "layers.roi_target", "nets.model", "postprocessings.single_shot",
]


_ALIASES = {
    "layers.anchor_target.modules.anchor_target": "layers.anchor_target",
    "layers.nms.nms_layer": "layers.nms_layer",
    "layers.roi_target.modules.roi_target": "layers.roi_target",
    "layers.roi_target.functions.roi_target": "layers.roi_target",
    "layers.roi_align_tf.roi_align_tf": "layers.roi_align_tf",
    "layers.roi_align_tf.crop_and_resize": "layers.roi_align_tf",
    "layers.roi_align_tf.pyramid_roi_align": "layers.roi_align_tf",
    "layers.roi_align.modules.roi_align": "layers.roi_align",
    "layers.roi_align.functions.roi_align": "layers.roi_align",
    "layers.roi_align.modules.pyramid_roi_align": "layers.roi_align",
}


def install(name: str = "libs"):
    """Register this package (and its sub-modules) in sys.modules under `name`."""
    pkg = sys.modules[__name__]
    if name in sys.modules and sys.modules[name] is not pkg:
        raise ImportError("a different %r package is already imported" % name)
    sys.modules[name] = pkg
    for pk in ("configs", "boxes", "nms", "layers", "nets", "postprocessings"):
        sys.modules["%s.%s" % (name, pk)] = importlib.import_module("%s.%s" % (__name__, pk))
    for sub in _SUBMODULES:
        sys.modules["%s.%s" % (name, sub)] = importlib.import_module("%s.%s" % (__name__, sub))
    for alias, sub in _ALIASES.items():          # the reference's deeper paths -> the flat mirror modules
        sys.modules["%s.%s" % (name, alias)] = sys.modules["%s.%s" % (name, sub)]
    return pkg


def uninstall(name: str = "libs"):
    """Remove the aliases install() created (the mirror itself stays importable as detbox.libs)."""
    pkg = sys.modules[__name__]
    if sys.modules.get(name) is pkg:
        for k in [k for k in sys.modules if k == name or k.startswith(name + ".")]:
            del sys.modules[k]
