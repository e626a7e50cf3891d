#!/usr/bin/env python
"""TEST INFRASTRUCTURE (runs only where /root/reference exists): tests/golden/anchor_f64_delta.npz.

ADVICE r1: the reference has TWO anchor dtypes on its own paths — the per-sample dataset path hands the float64 pyramid to
libs/layers/anchor.py:17 encode (data_layer.py:99-102), the batch path casts it to float32 first (data_layer.py:138-140).  CP-R's
anchor scales (2^(k/3)) are not float32 numbers, so the two can label an anchor differently where `ov == gt_max` or a
threshold is met to the last bits.  This records, from the COMPILED reference, the labels of both paths on real CityPersons
GT sets x the CP-R pyramid: the float32-path labels (what this library implements, sparse) and the indices / values where the
float64 path differs."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import ref_loader                                             # noqa: E402
from oracle.make_golden import CPR_RATIOS, CPR_SCALES, CPR_STRIDES, GOLD, real_gt_sets, set_cfg   # noqa: E402


def main():
    r = ref_loader.load()
    pyr64 = r.boxes_anchor.anchor_pyramid(CPR_STRIDES, 1024, 2048, CPR_SCALES, CPR_RATIOS, 16)
    assert pyr64.dtype == np.float64
    pyr32 = pyr64.astype(np.float32)
    set_cfg(r.cfg, 0.5, 0.6)
    sets = real_gt_sets(picks=None)
    n = np.asarray([g.shape[0] for _, g in sets])
    picks = [int(i) for i in np.argsort(-n)[:12]]           # the 12 densest images: most chances for a last-bit decision
    out = {"image_ids": np.asarray(picks, np.int64), "anchors_differ": np.asarray((pyr64 != pyr32.astype(np.float64)).any(1).sum())}
    total = 0
    for i in picks:
        gt = sets[i][1]
        l32, w32, t32, b32 = r.layers_anchor.encode(gt.copy(), pyr32.copy())
        l64, w64, t64, b64 = r.layers_anchor.encode(gt.copy(), pyr64.copy())
        k = "img%d_" % i
        nz = np.nonzero(l32 != 0)[0]
        d = np.nonzero(l32 != l64)[0]
        out[k + "gt"] = gt
        out[k + "labels32_nz_idx"], out[k + "labels32_nz_val"] = nz, l32[nz]
        out[k + "diff_idx"], out[k + "labels64_at_diff"] = d, l64[d]
        out[k + "inside_diff"] = np.asarray((b32 != b64).any(1).sum())
        out[k + "max_target_delta"] = np.asarray(np.abs(t32.astype(np.float64) - t64).max())
        total += d.size
        print("image %3d  G=%2d  labels differing between the float64- and float32-anchor paths: %d; inside-weight rows differing: %d; "
              "max |target delta| %.2e" % (i, gt.shape[0], d.size, int(out[k + "inside_diff"]), float(out[k + "max_target_delta"])))
    out["total_label_diffs"] = np.asarray(total)
    np.savez_compressed(os.path.join(GOLD, "anchor_f64_delta.npz"), **out)
    print("anchors whose float32 rounding changes a coordinate: %d of %d; label differences over 12 images: %d"
          % (int(out["anchors_differ"]), pyr64.shape[0], total))


if __name__ == "__main__":
    main()
