#!/usr/bin/env python
"""TEST INFRASTRUCTURE (runs only where /root/reference exists): golden vectors of the round-2 coverage rows, produced by
the reference's OWN code — tests/golden/r2_rows.npz.

  roi_*          libs/layers/roi.py:16-100 encode (compiled in place, oracle/_ref) with np.random seeded: inputs, the seed,
                 and (labels, bbox_targets, bbox_inside_weights); with and without GT; roi.py:102-125 decode
  trad_*         libs/layers/anchor.py:17-145 encode with cfg.rpn_sample_strategy = 'traditional' (np.random seeded)
  assign_*       libs/layers/assign.py:10-43 assign_boxes (pure NumPy; imported from the reference tree as it is)
  lin_*          libs/boxes/bbox_transform.py:93-144 bbox_transform_linear / _inv_linear with alpha = 4 and 10
"""
import importlib.util
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import ref_loader                                             # noqa: E402
from oracle.make_golden import GOLD, set_cfg, synth_gt                     # noqa: E402


def main():
    r = ref_loader.load()
    cfg = r.cfg
    out = {}
    # ---- roi.encode / decode ------------------------------------------------------------------------------------
    cfg.fg_threshold, cfg.roi_bg_threshold, cfg.rois_per_image, cfg.roi_fg_fraction = 0.5, 0.3, 128, 0.25
    rng = np.random.RandomState(71)
    for name, G, R in (("roi_a", 9, 600), ("roi_b", 3, 90), ("roi_nogt", 0, 200)):
        gt = synth_gt(rng, G, 512, 768, 0.0, ncls=4) if G else np.zeros((0, 5), np.float32)
        base = gt[rng.randint(0, max(G, 1), R), :4] if G else np.zeros((R, 4), np.float32)
        jit = rng.normal(0, 12, (R, 4)).astype(np.float32) * (rng.rand(R, 1) < 0.7)
        rois = (base + jit).astype(np.float32)
        far = rng.rand(R) < 0.4
        rois[far] = synth_gt(rng, int(far.sum()), 512, 768, 0.0)[:, :4]
        rois[:, 2:] = np.maximum(rois[:, 2:], rois[:, :2] + 2)
        seed = 1000 + len(out)
        np.random.seed(seed)
        lab, bt, bw = r.layers_roi.encode(gt.copy(), rois.copy(), 5)
        out.update({name + "_gt": gt, name + "_rois": rois, name + "_seed": np.asarray(seed), name + "_labels": lab,
                    name + "_bbox_targets": bt, name + "_bbox_inside_weights": bw})
    K, R = 5, 300
    rois = synth_gt(rng, R, 512, 768, 0.0)[:, :4]
    deltas = rng.normal(0, 0.4, (R, 4 * K)).astype(np.float32)
    scores = rng.rand(R, K).astype(np.float32)
    fb, cls, sc = r.layers_roi.decode(deltas.copy(), scores.copy(), rois.copy(), 512, 768)
    out.update(dec_rois=rois, dec_deltas=deltas, dec_scores=scores, dec_boxes=fb, dec_classes=cls, dec_max=sc)
    # ---- 'traditional' anchor sampling --------------------------------------------------------------------------
    set_cfg(cfg, 0.4, 0.6)
    cfg.rpn_sample_strategy, cfg.rpn_batch_size, cfg.rpn_fg_fraction = 'traditional', 384, 0.25
    anc = r.boxes_anchor.anchor_pyramid([8, 16, 32], 256, 384, [[2, 2.52, 3.175], [4, 5.04, 6.349], [8, 10.079, 12.7]], [1, 2.5], 16).astype(np.float32)
    out["trad_anchors"] = anc
    for name, G, bs in (("trad_a", 14, 384), ("trad_b", 2, 384), ("trad_c", 14, 64)):   # trad_c: 16 fg allowed < fg found -> BOTH draws happen
        gt = synth_gt(rng, G, 256, 384, 0.1)
        seed = 2000 + G + bs
        cfg.rpn_batch_size = bs
        out[name + "_batch_size"] = np.asarray(bs)
        np.random.seed(seed)
        lab, lw, bt, bw = r.layers_anchor.encode(gt.copy(), anc.copy())
        out.update({name + "_gt": gt, name + "_seed": np.asarray(seed), name + "_labels": lab, name + "_label_weights": lw,
                    name + "_bbox_targets": bt, name + "_bbox_inside_weights": bw})
    cfg.rpn_sample_strategy, cfg.rpn_batch_size = 'simple', 384
    # ---- assign_boxes (pure NumPy module, imported from the reference tree as it is) ---------------------------------
    spec = importlib.util.spec_from_file_location("ref_assign", "/root/reference/libs/layers/assign.py")
    m = importlib.util.module_from_spec(spec); spec.loader.exec_module(m)
    side = np.exp(rng.uniform(np.log(6), np.log(1500), 400)).astype(np.float32)
    ar = rng.uniform(0.4, 2.5, 400).astype(np.float32)
    x1, y1 = rng.uniform(0, 500, 400).astype(np.float32), rng.uniform(0, 500, 400).astype(np.float32)
    boxes = np.stack([x1, y1, x1 + side * ar, y1 + side / ar, np.ones(400, np.float32)], 1).astype(np.float32)
    exact = np.asarray([[0, 0, 224 * s, 224 * s, 1] for s in (0.125, 0.25, 0.5, 1, 2, 4, 8)], np.float32)   # sqrt(area) / 224 = powers of two
    boxes = np.concatenate([boxes, exact], 0)
    out.update(assign_boxes=boxes, assign_default=m.assign_boxes(boxes.copy()), assign_k36=m.assign_boxes(boxes.copy(), 3, 6, 112.0))
    np.seterr(all='warn')                                    # assign_boxes leaves np.seterr(all='raise') behind
    # ---- linear transforms with a general alpha -----------------------------------------------------------------------
    a = synth_gt(rng, 500, 600, 800, 0.0)[:, :4]
    g = (a + rng.normal(0, 15, a.shape)).astype(np.float32)
    d = rng.normal(0, 1.5, a.shape).astype(np.float32)
    for al in (4.0, 10.0):
        out["lin_enc_%g" % al] = r.bbox_transform.bbox_transform_linear(a.copy(), g.copy(), al)
        out["lin_dec_%g" % al] = r.bbox_transform.bbox_transform_inv_linear(a.copy(), d.copy(), al)
    out.update(lin_anchors=a, lin_gt=g, lin_deltas=d)
    np.savez_compressed(os.path.join(GOLD, "r2_rows.npz"), **out)
    for k in ("roi_a", "roi_b", "roi_nogt"):
        l = out[k + "_labels"]
        print(k, "fg %d bg %d ignored %d" % ((l > 0).sum(), (l == 0).sum(), (l < 0).sum()))
    for k in ("trad_a", "trad_b", "trad_c"):
        l = out[k + "_labels"]
        print(k, "fg %d bg %d ignored %d, weights all zero: %s" % ((l >= 1).sum(), (l == 0).sum(), (l < 0).sum(), not out[k + "_label_weights"].any()))
    print("assign levels", np.bincount(out["assign_default"]), "file KB", os.path.getsize(os.path.join(GOLD, "r2_rows.npz")) // 1024)


if __name__ == "__main__":
    main()
