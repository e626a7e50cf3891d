"""CPU-only checks of the drop-in boundary: libdetbox.so loads, exports every symbol that
include/detbox.h declares, the ctypes prototypes cover them all, and the host-side pieces
(base anchors, argument validation that returns before any CUDA call) behave."""
import ctypes
import os
import re

import numpy as np

import detbox
from detbox import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    text = open(os.path.join(ROOT, "include", "detbox.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    names = re.findall(r"^\s*(?:const\s+char\s*\*\s*|(?:int|size_t|void)\s+)([A-Za-z_]\w*)\s*\(", text, flags=re.M)
    return sorted(set(names))


def test_library_exports_every_declared_symbol():
    names = header_symbols()
    assert len(names) >= 25 and "_nms" in names and "detbox_anchor_target" in names
    L = ctypes.CDLL(_lib.LIB_PATH)
    missing = [n for n in names if not hasattr(L, n)]
    assert not missing, "declared in detbox.h but not exported: %s" % missing


def test_ctypes_prototypes_cover_the_header():
    assert sorted(_lib.SIGNATURES) == header_symbols()
    detbox.lib()        # attaches every prototype; raises if one cannot be resolved


def test_struct_layouts_match_header(tmp_path):
    """Every ctypes mirror of a struct in include/detbox.h against what the C compiler makes of the header itself: sizes and
    the offsets of the last fields (a field added to the header but not to the mirror, or the other way round, shows here)."""
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    src = tmp_path / "sizes.c"
    src.write_text('#include <stdio.h>\n#include <stddef.h>\n#include "detbox.h"\nint main(void) {\n'
                   'printf("%zu %zu %zu %zu %zu %zu %zu\\n", sizeof(detbox_target_cfg), sizeof(detbox_roi_cfg), sizeof(detbox_match_cfg),\n'
                   '       sizeof(detbox_pyramid_desc), offsetof(detbox_roi_cfg, extrapolation), offsetof(detbox_roi_cfg, roi_order),\n'
                   '       offsetof(detbox_target_cfg, bbweights));\nreturn 0; }\n')
    exe = tmp_path / "sizes"
    subprocess.check_call(["gcc", "-I", os.path.join(root, "include"), str(src), "-o", str(exe)])
    got = [int(x) for x in subprocess.check_output([str(exe)]).split()]
    want = [ctypes.sizeof(_lib.TargetCfg), ctypes.sizeof(_lib.RoiCfg), ctypes.sizeof(_lib.MatchCfg), ctypes.sizeof(_lib.PyramidDesc),
            _lib.RoiCfg.extrapolation.offset, _lib.RoiCfg.roi_order.offset, _lib.TargetCfg.bbweights.offset]
    assert got == want, (got, want)
    assert ctypes.sizeof(_lib.TargetCfg) == 3 * 8 + 2 * 4 + 4 * 4


def test_base_anchors_host(golden):
    g = golden("anchors")
    np.testing.assert_array_equal(detbox.base_anchors([2., 4., 8., 16., 32.], [0.5, 1, 2.0], 16), g["base_default"])
    np.testing.assert_array_equal(detbox.base_anchors([1.5, 3], [0.5, 1, 2.0, 3.0], 15), g["base_odd15"])


def test_argument_validation_without_gpu():
    L = detbox.lib()
    assert L.detbox_error_string(0) == b"ok"
    assert L.detbox_box_encode(None, None, -1, 0, None, None) == -1          # DETBOX_ERR_INVALID
    assert L.detbox_box_decode(None, None, 5, 9, 0, 0, None, None) == -1
    assert L.detbox_nms_workspace_bytes(0, 1) > 0
    n = 20000
    need = L.detbox_nms_workspace_bytes(n, 1)
    assert need >= n * ((n + 63) // 64) * 8                                  # the bitmask dominates
    assert L.detbox_topk_decode_workspace_bytes(64, 261888, 900, 1) >= 64 * 261888 * 4
    assert b"workspace" in L.detbox_error_string(-2)


def test_ops_refuse_cpu_tensors():
    import pytest
    import torch
    with pytest.raises(detbox.DetboxError):
        detbox.nms(torch.zeros(3, 4), torch.zeros(3), 0.5)
    with pytest.raises(detbox.DetboxError):
        detbox.bbox_overlaps(torch.zeros(3, 4, dtype=torch.float64), torch.zeros(2, 4, dtype=torch.float64))


def test_tiled_roi_backward_rejects_what_it_does_not_cover():
    """detbox_roi_align_backward_tiled decides from the configuration alone (before any CUDA call) whether it takes a case;
    DETBOX_ERR_UNSUPPORTED (-4) tells the caller to use detbox_roi_align_backward.  Workspace: 16 + 8 * (ph + pw) bytes per RoI,
    each array rounded up to 256 bytes."""
    L = detbox.lib()

    def cfg(**kw):
        c = _lib.RoiCfg()
        c.n_levels, c.batch, c.channels, c.pooled_h, c.pooled_w = 1, 2, 8, 7, 7
        c.grad_feat[0], c.height[0], c.width[0], c.spatial_scale[0] = 0x1000, 16, 32, 0.25      # never dereferenced here
        c.mode = detbox.ROI_BIN_CENTRE
        for k, v in kw.items():
            setattr(c, k, v)
        return c

    def call(c, R=4):
        return L.detbox_roi_align_backward_tiled(ctypes.byref(c), 0x2000, R, None, 0x3000, None, 0, None)

    for bad in (dict(layout=1, out_layout=1), dict(mode=detbox.ROI_GRID_POINT), dict(all_levels=1), dict(pooled_h=14, pooled_w=14),
                dict(mode=detbox.ROI_GRID_POINT, pool=1)):
        assert call(cfg(**bad)) == -4, bad
    assert call(cfg()) == -2                     # covered, but no workspace was given: DETBOX_ERR_WORKSPACE, still before any launch
    assert call(cfg(), R=-1) == -1               # DETBOX_ERR_INVALID
    up = lambda v: (v + 255) // 256 * 256
    assert L.detbox_roi_align_backward_tiled_workspace_bytes(ctypes.byref(cfg()), 1000) == up(1000 * 16) + up(1000 * 14 * 8)
