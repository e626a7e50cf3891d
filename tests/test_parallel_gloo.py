"""The N>1 path on CPU: two `gloo` ranks shard images round-robin and all-gather their
fixed-size detection buffers.  No kernel runs here — this covers the host-side sharding logic."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import detbox


def _dets(i):
    rng = np.random.RandomState(1000 + i)
    m = int(rng.randint(0, 7))
    d = rng.rand(m, 6).astype(np.float32)
    d[:, 5] = i                                   # tag rows with their image
    return d


def _worker(rank, world, port, n_images, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        mine = detbox.parallel.shard_images(n_images, rank, world)
        got = detbox.parallel.gather_detections([_dets(i) for i in mine], n_images, max_dets=5)
        ok = len(got) == n_images and all(np.array_equal(g, _dets(i)[:5]) for i, g in enumerate(got))
        q.put((rank, mine, bool(ok)))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("n_images", [8, 5, 1])
def test_two_rank_shard_and_gather(n_images):
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, n_images, q)) for r in range(2)]
    [p.start() for p in procs]
    res = sorted(q.get(timeout=120) for _ in procs)
    [p.join(60) for p in procs]
    assert all(p.exitcode == 0 for p in procs)
    assert res[0][1] == list(range(0, n_images, 2)) and res[1][1] == list(range(1, n_images, 2))
    assert res[0][2] and res[1][2]
    owned = sorted(res[0][1] + res[1][1])
    assert owned == list(range(n_images))         # a partition: every image exactly once


def test_shard_helpers_single_process():
    assert detbox.parallel.shard_images(10, 3, 4) == [3, 7]
    assert detbox.parallel.owner_of(7, 4) == 3
    buf, cnt = detbox.parallel.pack_detections([np.ones((3, 6)), np.zeros((0, 6))], 2)
    assert buf.shape == (2, 2, 6) and cnt.tolist() == [2, 0]
    out = detbox.parallel.gather_detections([np.ones((3, 6))], 1, max_dets=2)
    assert out[0].shape == (2, 6)
    with pytest.raises(ValueError):
        detbox.parallel.shard_images(4, 4, 4)
