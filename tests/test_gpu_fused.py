"""GPU parity of the stage-one epilogue: the to_Dets* conversions against vectors recorded from
the reference's own methods, the device-side thresholding, and detbox.fused.stage_one (== get_final_results_stage1,
libs/nets/model.py:474-497) against the oracle's composition — class ids / probabilities / row order bit-exact,
boxes within 1e-5 — including the eager == CUDA-graph-replay check that shows the sequence has no host sync."""
from importlib import import_module

import numpy as np
import pytest
import torch

import detbox
from oracle import oracle as O
from oracle.make_golden import to_dets_cases, to_dets_inputs
from util import assert_boxes_close

pytestmark = pytest.mark.gpu


def cu(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def model():
    return import_module(detbox.libs.__name__ + ".nets.model")


@pytest.mark.parametrize("case", to_dets_cases(), ids=lambda c: c[0])
def test_to_dets_golden(golden, case):
    M = model()
    g = golden("to_dets")
    name, act, multi, C, thr, B, _ = case
    fn = {("softmax", False): M.to_Dets, ("sigmoid", False): M.to_Dets_sigmoid,
          ("softmax", True): M.to_Dets2, ("sigmoid", True): M.to_Dets2_sigmoid}[(act, multi)]
    b, p, i = g[name + "_boxes"], g[name + "_probs"], g[name + "_ids"]
    dets = fn(cu(b), cu(p), cu(i), thr, batch_size=B) if multi else fn(cu(b), cu(p), cu(i), batch_size=B)
    assert len(dets) == B and all(d.dtype == np.float64 for d in dets)
    assert np.array_equal(np.cumsum([0] + [d.shape[0] for d in dets]), g[name + "_offsets"])
    assert np.array_equal(np.concatenate(dets, 0), g[name + "_dets"])
    dets_np = fn(b, p, i, thr, batch_size=B) if multi else fn(b, p, i, batch_size=B)      # NumPy arguments, like upstream's callers
    assert all(np.array_equal(x, y) for x, y in zip(dets, dets_np))


def test_to_dets_keep_and_capacity():
    rng = np.random.RandomState(0)
    n, C, B = 500, 7, 3
    b = rng.uniform(0, 100, (n, 4)).astype(np.float32); p = rng.rand(n, C).astype(np.float32)
    ids = np.sort(rng.randint(0, B, n)).astype(np.int64)
    keep = np.sort(rng.choice(n, 200, replace=False)).astype(np.int64)
    buf = np.concatenate([keep, rng.randint(0, n, 50)])                        # rows beyond n_keep must be ignored
    dets, off = detbox.to_dets(cu(b), cu(p), cu(ids), B, "sigmoid", True, 0.6, keep=cu(buf), n_keep=cu(np.asarray([200], np.int32)))
    want = O.to_dets(b[keep], p[keep], ids[keep], B, "sigmoid", True, 0.6)
    got = detbox.fused.split(dets, off)
    assert all(np.array_equal(x, y) for x, y in zip(got, want))
    small, off2 = detbox.to_dets(cu(b), cu(p), cu(ids), B, "sigmoid", True, 0.6, cap=10)
    assert int(off2[-1]) == sum(x.shape[0] for x in O.to_dets(b, p, ids, B, "sigmoid", True, 0.6)) > 10
    with pytest.raises(RuntimeError):
        detbox.fused.split(small, off2)


@pytest.mark.parametrize("thr,none_pass", [(0.4, False), (2.0, True)])
def test_threshold_segments(thr, none_pass):
    rng = np.random.RandomState(1)
    s = rng.rand(3000).astype(np.float32); s[1234] = s.max()                  # duplicate maximum: lowest index wins
    ids = rng.randint(0, 5, 3000).astype(np.int64)
    seg32, seg64 = detbox.threshold_segments(cu(s), cu(ids), 5, thr)
    valid = s >= np.float32(thr)
    if none_pass:
        assert not valid.any()
        valid = np.zeros_like(valid); valid[int(np.argmax(s))] = True
    want = np.where(valid, ids, 5)
    assert np.array_equal(seg64.cpu().numpy(), want) and np.array_equal(seg32.cpu().numpy(), want.astype(np.int32))


@pytest.mark.parametrize("act,C,thr", [("softmax", 2, 0.3), ("softmax", 4, 0.2), ("sigmoid", 3, 0.5), ("softmax", 2, 5.0)])
def test_stage_one_matches_oracle(act, C, thr):
    cfg = O.citypersons_retinanet_cfg()
    strides, scales = cfg.strides[:3], cfg.anchor_scales[:3]
    ih, iw, B, max_dets = 192, 320, 3, 40
    rng = np.random.RandomState(C * 10 + int(thr * 10))
    anc = O.anchor_pyramid(strides, ih, iw, scales, cfg.anchor_ratios, cfg.anchor_base).astype(np.float32)
    A = anc.shape[0]
    deltas = rng.normal(0, 0.5, (B * A, 4)).astype(np.float32)
    if act == "softmax":
        fg = rng.permutation(np.linspace(0.0, 0.95, B * A)).astype(np.float32)            # unique objectness
        share = rng.dirichlet(np.ones(C - 1), B * A).astype(np.float32)
        prob = np.concatenate([(1 - fg)[:, None], fg[:, None] * share], 1).astype(np.float32)
    else:
        prob = rng.rand(B * A, C).astype(np.float32)
    want = O.final_results_stage1(deltas, prob, anc, B, thr, max_dets, 0.5, act)
    run = lambda: detbox.fused.stage_one(cu(deltas), cu(prob), cu(anc), B, thr, max_dets, 0.5, act)
    got = detbox.fused.split(*run())
    assert len(got) == B
    for g_, w_ in zip(got, want):
        assert g_.shape == w_.shape
        assert np.array_equal(g_[:, 4:], w_[:, 4:])
        assert_boxes_close(g_[:, :4], w_[:, :4])
    # the mirror's get_final_results_stage1 is the same call
    M = model()
    mcfg = import_module(detbox.libs.__name__ + ".configs.config")
    old = mcfg.rpn_box_encoding
    try:
        dets = M.get_final_results_stage1(cu(deltas), cu(prob), cu(anc), thr, max_dets, 0.5, activation=act, batch_size=B)
    finally:
        mcfg.rpn_box_encoding = old
    assert all(np.array_equal(x, y) for x, y in zip(dets, got))


def test_stage_one_is_graph_capturable():
    cfg = O.citypersons_retinanet_cfg()
    anc = cu(O.anchor_pyramid(cfg.strides[:3], 128, 256, cfg.anchor_scales[:3], cfg.anchor_ratios, cfg.anchor_base).astype(np.float32))
    A, B = anc.shape[0], 2
    g = torch.Generator(device="cuda"); g.manual_seed(0)
    deltas = torch.randn((B * A, 4), device="cuda", generator=g) * 0.5
    p1 = torch.rand((B * A,), device="cuda", generator=g)
    prob = torch.stack([1 - p1, p1], 1).contiguous()
    fn = lambda: detbox.fused.stage_one(deltas, prob, anc, B, 0.5, 30, 0.5)
    ref = fn()
    side = torch.cuda.Stream(); side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        fn(); fn()
    torch.cuda.current_stream().wait_stream(side); torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        out = fn()
    graph.replay(); torch.cuda.synchronize()
    n = int(ref[1][-1])
    assert torch.equal(out[1], ref[1]) and torch.equal(out[0][:n], ref[0][:n])
