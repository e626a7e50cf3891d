"""GPU parity of the batched host-buffer entry (detbox.pipeline.BoxOpsSession) — the call bench.py's
end-to-end leg times — against the CPU oracle, image by image: anchor labels / weights and NMS keep
indices bit-exact, regression targets and decoded boxes within 1e-5 (util.assert_boxes_close)."""
from importlib import import_module

import numpy as np
import pytest
import torch

from oracle import oracle as O
from oracle.make_golden import synth_gt
from util import assert_boxes_close

pytestmark = pytest.mark.gpu


def make_case(B, ih, iw, seed):
    cfg = O.citypersons_retinanet_cfg()
    strides, scales = cfg.strides[:3], cfg.anchor_scales[:3]
    rng = np.random.RandomState(seed)
    anc = O.anchor_pyramid(strides, ih, iw, scales, cfg.anchor_ratios, cfg.anchor_base).astype(np.float32)
    A = anc.shape[0]
    gts = [synth_gt(rng, int(g), ih, iw, 0.2) if g else np.zeros((0, 5), np.float32) for g in rng.randint(0, 9, B)]
    deltas = rng.normal(0, 0.5, (B * A, 4)).astype(np.float32)
    p1 = rng.permutation(np.linspace(0.0, 1.0, B * A)).astype(np.float32)          # unique scores
    prob = np.stack([1 - p1, p1], 1).astype(np.float32)
    return cfg, strides, scales, anc, gts, deltas, prob


@pytest.mark.parametrize("B,chunk", [(5, 2), (4, 4), (3, 8)])
def test_run_host_matches_oracle(B, chunk):
    pipeline = import_module("detectron-pytorch_b200.pipeline")
    ih, iw, top_n = 256, 512, 300
    cfg, strides, scales, anc, gts, deltas, prob = make_case(B, ih, iw, 11 + B)
    A = anc.shape[0]
    sess = pipeline.BoxOpsSession(strides, ih, iw, scales, cfg.anchor_ratios, cfg.anchor_base, B, 2, top_n=top_n,
                                  nms_thresh=0.5, bg=cfg.rpn_bg_threshold, fg=cfg.rpn_fg_threshold, max_gt_total=16 * B)
    assert sess.A == A
    h_prob, h_box = torch.from_numpy(prob).pin_memory(), torch.from_numpy(deltas).pin_memory()
    out = sess.run_host(gts, h_prob, h_box, chunk=chunk)
    dev = sess.run_device(torch.from_numpy(np.concatenate(gts, 0)).cuda(), sess.h_off.cuda(), h_prob.cuda(), h_box.cuda())
    torch.cuda.synchronize()
    nk = out["n_keep"].numpy()
    assert np.array_equal(nk, dev["n_keep"].cpu().numpy())                      # chunked == one-shot
    assert np.array_equal(out["keep"].numpy()[:nk[B]], dev["keep"].cpu().numpy()[:nk[B]])
    assert np.array_equal(out["img_ids"].numpy(), dev["img_ids"].cpu().numpy())
    pos = 0
    for b in range(B):
        want = O.anchor_encode(gts[b], anc, cfg)
        assert np.array_equal(out["labels"][b].numpy(), want[0])
        assert np.array_equal(out["label_weights"][b].numpy(), want[1])
        assert_boxes_close(out["bbox_targets"][b].numpy(), want[2])
        assert np.array_equal(out["bbox_inside_weights"][b].numpy(), want[3])
        wb, wp, _, _, widx = O.decode_and_choose_top_n(deltas[b * A:(b + 1) * A], prob[b * A:(b + 1) * A], anc, top_n, "softmax")
        rows = slice(b * top_n, (b + 1) * top_n)
        assert_boxes_close(out["boxes"].numpy()[rows], wb)
        assert np.array_equal(out["scores"].numpy()[rows], 1.0 - wp[:, 0])
        assert (out["img_ids"].numpy()[rows] == b).all()
        wk = O.nms(out["boxes"].numpy()[rows], out["scores"].numpy()[rows], 0.5, O.NMS_GT_F32)
        got = out["keep"].numpy()[pos:pos + nk[b]]
        assert np.array_equal(got, wk + b * top_n)
        pos += nk[b]
    dets = sess.detections(out)
    assert len(dets) == B and all(d.shape[1] == 5 for d in dets)


def test_run_host_targets_on_device():
    """targets="device": the same numbers as the host-buffer mode, the four target tensors left in HBM."""
    pipeline = import_module("detectron-pytorch_b200.pipeline")
    B, ih, iw, top_n = 5, 256, 512, 200
    cfg, strides, scales, anc, gts, deltas, prob = make_case(B, ih, iw, 3)
    sess = pipeline.BoxOpsSession(strides, ih, iw, scales, cfg.anchor_ratios, cfg.anchor_base, B, 2, top_n=top_n,
                                  nms_thresh=0.5, bg=cfg.rpn_bg_threshold, fg=cfg.rpn_fg_threshold, max_gt_total=16 * B)
    h_prob, h_box = torch.from_numpy(prob).pin_memory(), torch.from_numpy(deltas).pin_memory()
    host = {k: v.clone() for k, v in sess.run_host(gts, h_prob, h_box, chunk=2).items() if torch.is_tensor(v)}
    d2h_host = sess.d2h_bytes
    dev = sess.run_host(gts, h_prob, h_box, chunk=2, targets="device")
    assert sess.d2h_bytes < d2h_host // 10
    for k in ("labels", "label_weights", "bbox_targets", "bbox_inside_weights"):
        assert dev[k].is_cuda and torch.equal(dev[k].cpu(), host[k])
    for k in ("boxes", "scores", "img_ids", "keep", "n_keep"):
        assert not dev[k].is_cuda and torch.equal(dev[k], host[k])


def test_run_device_two_streams_matches_one_stream():
    """run_device forks the detection chain onto a second stream (pipeline.py): same bits as the serial order,
    also when the call is repeated back to back (the fork waits for every earlier use of the inputs); and the
    same again as one CUDA-graph replay (run_device_graph), including after the inputs were refilled in place."""
    pipeline = import_module("detectron-pytorch_b200.pipeline")
    B, ih, iw, top_n = 6, 256, 512, 300
    cfg, strides, scales, anc, gts, deltas, prob = make_case(B, ih, iw, 23)
    sess = pipeline.BoxOpsSession(strides, ih, iw, scales, cfg.anchor_ratios, cfg.anchor_base, B, 2, top_n=top_n,
                                  nms_thresh=0.5, bg=cfg.rpn_bg_threshold, fg=cfg.rpn_fg_threshold, max_gt_total=16 * B)
    sess.stage_gt(gts)
    args = (torch.from_numpy(np.concatenate(gts, 0)).cuda(), sess.h_off.cuda(), torch.from_numpy(prob).cuda(), torch.from_numpy(deltas).cuda())

    def same(x, y):
        torch.cuda.synchronize()
        n = int(x["n_keep"][-1])
        assert n == int(y["n_keep"][-1]) and n > 0
        for k in x:                                   # rows of `keep` past n_keep[-1] are not written
            assert torch.equal(x[k][:n] if k == "keep" else x[k], y[k][:n] if k == "keep" else y[k]), k

    serial = sess.run_device(*args, overlap=False)
    for _ in range(3):
        forked = sess.run_device(*args, overlap=True)
    same(serial, forked)
    same(serial, sess.run_device_graph(*args))
    assert sess.graph_launches >= 10
    args[2].copy_(args[2].flip(0)); args[3].copy_(args[3].flip(0))
    want = sess.run_device(*args, overlap=False)
    assert not torch.equal(want["scores"], serial["scores"])
    same(want, sess.run_device_graph(*args))


@pytest.mark.parametrize("box_rows", ["copy", "mapped"])
def test_run_host_compact_format_expands_to_reference(box_rows):
    """format="compact" (17 B / anchor over the link) + the host expander == format="reference" bit for bit, and the
    top-k gather reading rpn_box rows out of the pinned host buffer == the staged copy."""
    pipeline = import_module("detectron-pytorch_b200.pipeline")
    B, ih, iw, top_n = 6, 256, 512, 250
    cfg, strides, scales, anc, gts, deltas, prob = make_case(B, ih, iw, 21)
    sess = pipeline.BoxOpsSession(strides, ih, iw, scales, cfg.anchor_ratios, cfg.anchor_base, B, 2, top_n=top_n,
                                  nms_thresh=0.5, bg=cfg.rpn_bg_threshold, fg=cfg.rpn_fg_threshold, max_gt_total=16 * B)
    h_prob, h_box = torch.from_numpy(prob).pin_memory(), torch.from_numpy(deltas).pin_memory()
    ref = {k: v.clone() for k, v in sess.run_host(gts, h_prob, h_box, chunk=4, box_rows="copy").items() if torch.is_tensor(v)}
    h2d_ref, d2h_ref = sess.h2d_bytes, sess.d2h_bytes
    out = sess.run_host(gts, h_prob, h_box, chunk=4, format="compact", box_rows=box_rows)
    assert out["packed"].dtype == torch.uint8 and tuple(out["packed"].shape) == (B, sess.A)
    assert sess.d2h_bytes < 0.45 * d2h_ref
    if box_rows == "mapped":
        assert sess.h2d_bytes < 0.4 * h2d_ref
    labels, lw, bt, bw = sess.expand(out)
    for got, k in ((labels, "labels"), (lw, "label_weights"), (bt, "bbox_targets"), (bw, "bbox_inside_weights")):
        assert torch.equal(got, ref[k]), k
    for k in ("boxes", "scores", "img_ids", "n_keep"):
        assert torch.equal(out[k], ref[k]), k
    nk = int(out["n_keep"][B])
    assert torch.equal(out["keep"][:nk], ref["keep"][:nk])
    cnt = out["counts"].numpy()
    for b in range(B):
        lab = ref["labels"][b].numpy()
        assert cnt[b, 0] == (lab >= 1).sum() and cnt[b, 1] == (lab == 0).sum()


def test_run_host_reference_arrays_over_compact_wire():
    """format="reference", wire="compact": the caller gets the reference's four arrays; each chunk crosses the link in
    the compact format and is expanded on worker threads while later chunks are in flight.  Bit-identical to wire="same",
    on two consecutive calls (the second over the first's leftovers), with a third of the bytes device->host."""
    pipeline = import_module("detectron-pytorch_b200.pipeline")
    B, ih, iw, top_n = 7, 256, 512, 250
    cfg, strides, scales, anc, gts, deltas, prob = make_case(B, ih, iw, 31)
    sess = pipeline.BoxOpsSession(strides, ih, iw, scales, cfg.anchor_ratios, cfg.anchor_base, B, 2, top_n=top_n,
                                  nms_thresh=0.5, bg=cfg.rpn_bg_threshold, fg=cfg.rpn_fg_threshold, max_gt_total=16 * B)
    h_prob, h_box = torch.from_numpy(prob).pin_memory(), torch.from_numpy(deltas).pin_memory()
    ref = {k: v.clone() for k, v in sess.run_host(gts, h_prob, h_box, chunk=2).items() if torch.is_tensor(v)}
    d2h_ref = sess.d2h_bytes
    for h in (sess.h_labels, sess.h_lw, sess.h_bw, sess.h_bt):
        h.fill_(-123.0)                                        # stale contents must all be overwritten
    for rep in range(2):
        out = sess.run_host(gts, h_prob, h_box, chunk=2, wire="compact")
        assert out["format"] == "reference" and sess.d2h_bytes < 0.45 * d2h_ref
        for k in ("labels", "label_weights", "bbox_targets", "bbox_inside_weights", "boxes", "scores", "img_ids", "n_keep"):
            assert torch.equal(out[k], ref[k]), (rep, k)
        nk = int(out["n_keep"][B])
        assert torch.equal(out["keep"][:nk], ref["keep"][:nk])
