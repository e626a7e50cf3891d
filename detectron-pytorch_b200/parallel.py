"""Per-image sharding across the GPUs of one box — host-side logic only.

The path has no exchange step: image i is owned by rank i % world (round-robin), every stage runs
device-local, and NO collective sits on the hot path.  The one optional collective is the epilogue
`gather_detections`: an all_gather of fixed-size [max_dets, 6] float32 buffers (+ counts, <= 2.4 KB
per image) so that every rank ends up with all detections.  It goes through torch.distributed
(NCCL over NVLink for CUDA tensors, gloo for the CPU tests); the reference's only parallelism is
torch.nn.DataParallel scatter/gather in one process (libs/nets/data_parallel.py:19-91)."""
from __future__ import annotations

import numpy as np
import torch
import torch.distributed as dist


def shard_images(n_images: int, rank: int, world: int):
    """Indices of the images rank `rank` owns: i with i % world == rank (ascending)."""
    if not 0 <= rank < world:
        raise ValueError("rank %d outside world %d" % (rank, world))
    return list(range(rank, n_images, world))


def owner_of(image_index: int, world: int) -> int:
    return image_index % world


def pack_detections(dets_list, max_dets: int, device="cpu"):
    """List of [m_i, 6] arrays (x1,y1,x2,y2,score,cls; best first) -> (buf [n, max_dets, 6] float32
    zero-padded, counts int32 [n]); rows beyond max_dets are dropped (libs/postprocessings/single_shot.py:59)."""
    n = len(dets_list)
    buf = np.zeros((n, max_dets, 6), np.float32)
    cnt = np.zeros((n,), np.int32)
    for i, d in enumerate(dets_list):
        d = np.asarray(d, np.float32).reshape(-1, 6)[:max_dets]
        buf[i, :d.shape[0]] = d
        cnt[i] = d.shape[0]
    return torch.from_numpy(buf).to(device), torch.from_numpy(cnt).to(device)


def unpack_detections(buf, counts):
    b, c = buf.cpu().numpy(), counts.cpu().numpy()
    return [b[i, :int(c[i])].copy() for i in range(b.shape[0])]


def gather_detections(local_dets, n_images: int, max_dets: int = 100, group=None, device="cpu"):
    """Every rank passes the detections of ITS images (in shard_images order) and receives the list of
    all n_images detection arrays in image order.  Ranks may own different image counts (n_images not
    a multiple of the world size): buffers are padded to the largest shard."""
    if not (dist.is_available() and dist.is_initialized()):
        if len(local_dets) != n_images:
            raise ValueError("single process: expected all %d images" % n_images)
        return [np.asarray(d, np.float32).reshape(-1, 6)[:max_dets] for d in local_dets]
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    mine = shard_images(n_images, rank, world)
    if len(local_dets) != len(mine):
        raise ValueError("rank %d owns %d images but got %d detection arrays" % (rank, len(mine), len(local_dets)))
    per = (n_images + world - 1) // world
    buf, cnt = pack_detections(list(local_dets) + [np.zeros((0, 6), np.float32)] * (per - len(mine)), max_dets, device)
    bufs = [torch.empty_like(buf) for _ in range(world)]
    cnts = [torch.empty_like(cnt) for _ in range(world)]
    dist.all_gather(bufs, buf, group=group)
    dist.all_gather(cnts, cnt, group=group)
    out = [None] * n_images
    for r in range(world):
        for j, d in enumerate(unpack_detections(bufs[r], cnts[r])[:len(shard_images(n_images, r, world))]):
            out[r + j * world] = d
    return out


# ---------------------------------------------------------------------------------- device-side epilogue --
def pack_detections_device(boxes, scores, keep, n_keep, n_local, max_dets=100, classes=None, out=None):
    """Device tensors from topk_decode / nms (n_local images, per-image segments) -> (buf [n_local, max_dets, 6] float32,
    counts int32 [n_local]) on the device with ONE kernel and no host synchronisation (detbox_pack_detections)."""
    from . import ops
    from ._lib import check, lib
    b = ops._dev(boxes).view(-1, 4)
    s = ops._dev(scores).view(-1)
    k = ops._dev(keep, torch.int64, "keep").view(-1)
    nk = ops._dev(n_keep, torch.int32, "n_keep").view(-1)
    c = ops._dev(classes).view(-1) if classes is not None else None
    if out is None:
        out = (torch.empty((n_local, max_dets, 6), dtype=torch.float32, device=b.device),
               torch.empty((n_local,), dtype=torch.int32, device=b.device))
    buf, cnt = out
    check(lib().detbox_pack_detections(b.data_ptr(), s.data_ptr(), ops._ptr(c), k.data_ptr(), nk.data_ptr(), int(n_local),
                                       int(max_dets), buf.data_ptr(), cnt.data_ptr(), ops._stream()), "pack_detections")
    ops._count(1 if n_local else 0)
    return buf, cnt


def gather_detections_device(buf, cnt, n_images: int, group=None, out=None):
    """The optional final all-gather on CUDA buffers (NCCL over NVLink): every rank passes the packed detections of ITS
    images (shard_images order, n_images a multiple of the world size) and receives (dets [n_images, max_dets, 6],
    counts [n_images]) in IMAGE order — views of the gathered buffers, no copy, no host synchronisation.
    Single process: returned as they are."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return buf, cnt
    world = dist.get_world_size(group)
    per = buf.shape[0]
    if per * world != n_images:
        raise ValueError("gather_detections_device needs n_images (%d) = world (%d) x images per rank (%d)" % (n_images, world, per))
    if out is None:
        out = (torch.empty((world,) + tuple(buf.shape), dtype=buf.dtype, device=buf.device),
               torch.empty((world, per), dtype=cnt.dtype, device=cnt.device))
    all_buf, all_cnt = out
    dist.all_gather_into_tensor(all_buf, buf, group=group)
    dist.all_gather_into_tensor(all_cnt, cnt, group=group)
    # image i = r + j * world lives at [r, j]: image order is the (j, r) traversal
    return all_buf.transpose(0, 1).reshape(n_images, buf.shape[1], buf.shape[2]), all_cnt.transpose(0, 1).reshape(n_images)
