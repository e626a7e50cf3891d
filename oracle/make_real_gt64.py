#!/usr/bin/env python
"""TEST / BENCH INFRASTRUCTURE (runs only where /root/reference exists): tests/golden/real_gt64.npz.

The first 64 CityPersons validation images (data/citypersons/annotations/anno_val.mat, file order) as GT sets
[x1, y1, x2, y2, class] — the conversion of oracle/make_golden.py::real_gt_sets (libs/datasets/citypersons2.py:85-89,
801-812) — concatenated with offsets, plus, for every image, what the COMPILED reference's own
libs/layers/anchor.py:17 encode() returns against the CP-R pyramid (rpn_bg 0.5 / rpn_fg 0.6): the (#fg, #bg, #ignored)
counts and the background label weight.  bench.py's `g_sweep.real_anno_val_64` runs the step on this batch (SURVEY 8d
config 1: "the 500 real anno_val.mat GT sets"), tests/test_gpu_boxes.py checks the counts bit-exactly."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import ref_loader                                             # noqa: E402
from oracle.make_golden import (CPR_RATIOS, CPR_SCALES, CPR_STRIDES, GOLD, real_gt_sets, set_cfg)   # noqa: E402


def main():
    r = ref_loader.load()
    pyr = r.boxes_anchor.anchor_pyramid(CPR_STRIDES, 1024, 2048, CPR_SCALES, CPR_RATIOS, 16).astype(np.float32)
    set_cfg(r.cfg, 0.5, 0.6)
    sets = real_gt_sets(picks=None)[:64]
    off = np.zeros(len(sets) + 1, np.int32)
    off[1:] = np.cumsum([g.shape[0] for _, g in sets])
    counts, bgw = [], []
    for _, gt in sets:
        lab, lw, _, _ = r.layers_anchor.encode(gt.copy(), pyr.copy())
        counts.append([(lab >= 1).sum(), (lab == 0).sum(), (lab < 0).sum()])
        bgw.append(lw[lab == 0][0] if (lab == 0).any() else 0.0)
    np.savez_compressed(os.path.join(GOLD, "real_gt64.npz"), image_ids=np.asarray([i for i, _ in sets], np.int64),
                        gt=np.concatenate([g for _, g in sets], 0).astype(np.float32), offsets=off,
                        counts=np.asarray(counts, np.int64), bg_weight=np.asarray(bgw, np.float32))
    g = off[1:] - off[:-1]
    print("64 real GT sets: G min %d / mean %.1f / max %d, %d ignore regions" % (g.min(), g.mean(), g.max(), sum((s[:, 4] < 0).sum() for _, s in sets)))


if __name__ == "__main__":
    main()
