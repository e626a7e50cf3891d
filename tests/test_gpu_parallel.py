"""The per-image sharded mode on real GPUs: the device-side detection packer against the host packer, and
two NCCL ranks — one GPU each, images dealt round-robin by detbox.parallel.shard_images — whose all-gathered detections
must equal the unsharded single-GPU result bit for bit (the path has no exchange step; the gather is the optional
epilogue).  The 2-rank test is skipped on a box with fewer than 2 GPUs."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import detbox
from oracle import oracle as O
from oracle.make_golden import synth_gt

pytestmark = pytest.mark.gpu
IH, IW, TOP_N, MAX_DETS = 256, 512, 300, 40


def make_inputs(n_images):
    cfg = O.citypersons_retinanet_cfg()
    strides, scales = cfg.strides[:3], cfg.anchor_scales[:3]
    A = O.anchor_pyramid(strides, IH, IW, scales, cfg.anchor_ratios, cfg.anchor_base).shape[0]
    per_image = []
    for i in range(n_images):                                  # every image from its own seed: independent of the sharding
        rng = np.random.RandomState(500 + i)
        gt = synth_gt(rng, 1 + i % 7, IH, IW, 0.2)
        deltas = rng.normal(0, 0.5, (A, 4)).astype(np.float32)
        p1 = rng.permutation(np.linspace(0.0, 1.0, A)).astype(np.float32)
        per_image.append((gt, deltas, np.stack([1 - p1, p1], 1).astype(np.float32)))
    return cfg, strides, scales, A, per_image


def run_shard(images, per_image, cfg, strides, scales):
    """The device-resident path on the given images -> packed detections (device) of those images."""
    B = len(images)
    sess = detbox.pipeline.BoxOpsSession(strides, IH, IW, scales, cfg.anchor_ratios, cfg.anchor_base, B, 2, top_n=TOP_N,
                                         nms_thresh=0.5, bg=cfg.rpn_bg_threshold, fg=cfg.rpn_fg_threshold, max_gt_total=16 * B)
    G = sess.stage_gt([per_image[i][0] for i in images])
    d_gt, d_off = sess.h_gt[:G].cuda(), sess.h_off.cuda()
    d_box = torch.from_numpy(np.concatenate([per_image[i][1] for i in images], 0)).cuda()
    d_prob = torch.from_numpy(np.concatenate([per_image[i][2] for i in images], 0)).cuda()
    out = sess.run_device(d_gt, d_off, d_prob, d_box)
    return out, detbox.parallel.pack_detections_device(out["boxes"], out["scores"], out["keep"], out["n_keep"], B, MAX_DETS)


def test_pack_detections_device_matches_host_packer():
    cfg, strides, scales, A, per_image = make_inputs(5)
    out, (buf, cnt) = run_shard(list(range(5)), per_image, cfg, strides, scales)
    torch.cuda.synchronize()
    nk = out["n_keep"].cpu().numpy()
    keep = out["keep"].cpu().numpy()[:nk[5]]
    boxes, scores = out["boxes"].cpu().numpy(), out["scores"].cpu().numpy()
    dets, o = [], 0
    for b in range(5):
        k = keep[o:o + nk[b]]; o += nk[b]
        dets.append(np.concatenate([boxes[k], scores[k, None], np.ones((k.size, 1), np.float32)], 1))
    want_buf, want_cnt = detbox.parallel.pack_detections(dets, MAX_DETS)
    assert np.array_equal(cnt.cpu().numpy(), want_cnt.numpy()) and (want_cnt.numpy() > 0).all()
    assert np.array_equal(buf.cpu().numpy(), want_buf.numpy())
    assert (nk[:5] > MAX_DETS).any()                           # truncation exercised


def _worker(rank, world, port, n_images, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        cfg, strides, scales, A, per_image = make_inputs(n_images)
        mine = detbox.parallel.shard_images(n_images, rank, world)
        _, (buf, cnt) = run_shard(mine, per_image, cfg, strides, scales)
        all_buf, all_cnt = detbox.parallel.gather_detections_device(buf, cnt, n_images)
        torch.cuda.synchronize()
        q.put((rank, mine, all_buf.cpu().numpy(), all_cnt.cpu().numpy()))
    finally:
        dist.destroy_process_group()


def test_two_rank_nccl_shard_and_gather_equals_single_gpu():
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (gpurun --gpus 2)")
    n_images = 8
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, n_images, q)) for r in range(2)]
    [p.start() for p in procs]
    res = sorted((q.get(timeout=300) for _ in procs), key=lambda t: t[0])
    [p.join(120) for p in procs]
    assert all(p.exitcode == 0 for p in procs)
    assert res[0][1] == [0, 2, 4, 6] and res[1][1] == [1, 3, 5, 7]
    cfg, strides, scales, A, per_image = make_inputs(n_images)
    _, (buf, cnt) = run_shard(list(range(n_images)), per_image, cfg, strides, scales)      # unsharded, this process, GPU 0
    for r in range(2):                                         # both ranks hold every image's detections, in image order
        assert np.array_equal(res[r][3], cnt.cpu().numpy())
        assert np.array_equal(res[r][2], buf.cpu().numpy())
