"""Per-rank host<->device copy rates with N ranks running at once (VERDICT r1 next #1a): is the end-to-end step
limited by each GPU's own PCIe link or by the host's memory system that all ranks share?

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P profiles/pcie_probe_multi.py

Every rank copies what ONE e2e step of bench.py moves in the compact format (135 MB pinned host -> device, 287 MB
device -> pinned host), first one direction at a time, then both at once on two streams, all ranks between the same
barriers; rank 0 prints per-rank and aggregate GB/s.  Pinned buffers are allocated after the rank pinned itself to the
CPUs `nvidia-smi topo -m` lists for its GPU (one NUMA node on this VM: the placement cannot differ)."""
import json, os, time
import torch, torch.distributed as dist

rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
H2D, D2H = 135_000_000, 287_000_000
hin = torch.empty(H2D // 4, dtype=torch.float32, pin_memory=True); hout = torch.empty(D2H // 4, dtype=torch.float32, pin_memory=True)
hin.fill_(1.0); hout.zero_()
din, dout = torch.empty_like(hin, device="cuda"), torch.ones(D2H // 4, dtype=torch.float32, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()

def run(kind, iters=8):
    def once():
        if kind in ("h2d", "both"):
            with torch.cuda.stream(s1): din.copy_(hin, non_blocking=True)
        if kind in ("d2h", "both"):
            with torch.cuda.stream(s2): hout.copy_(dout, non_blocking=True)
    once(); torch.cuda.synchronize()
    if world > 1: dist.barrier()
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(iters): once()
    torch.cuda.synchronize(); dt = (time.perf_counter() - t0) / iters
    nbytes = (H2D if kind in ("h2d", "both") else 0) + (D2H if kind in ("d2h", "both") else 0)
    return nbytes / dt / 1e9, dt * 1e3

res = {}
for kind in ("h2d", "d2h", "both"):
    gbs, ms = run(kind)
    t = torch.tensor([gbs, ms], dtype=torch.float64, device="cuda")
    if world > 1:
        allt = [torch.empty_like(t) for _ in range(world)]; dist.all_gather(allt, t)
    else:
        allt = [t]
    if rank == 0:
        g = [float(x[0]) for x in allt]
        res[kind] = {"per_rank_gbs": [round(x, 1) for x in g], "aggregate_gbs": round(sum(g), 1), "ms_max": round(max(float(x[1]) for x in allt), 2)}
if rank == 0:
    print(json.dumps({"ranks": world, "bytes": {"h2d": H2D, "d2h": D2H}, **res}))
if world > 1:
    dist.destroy_process_group()
