#!/usr/bin/env python
"""bench.py — box-ops images/s @1024x2048 (BASELINE.json metric), one JSON line on stdout.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl detbox|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port P bench.py --gpus N --steps K --warmup W

A STEP is one pass of the box path over `--batch` (default 64) synthetic CityPersons-shaped
1024x2048 images PER GPU (BASELINE.json configs[4] sharded per image, weak scaling):
    anchor pyramid (261,888 anchors) -> anchor-target assignment -> objectness + per-image
    top-900 + decode -> per-image NMS (IoU 0.5).
`value`  : images/s with the inputs already resident in HBM (CUDA events, max over ranks).
`e2e`    : the same step through detbox.pipeline.BoxOpsSession.run_host — pinned HOST buffers in,
           HOST buffers out, the copies inside the timed region.
Extra keys report the two other parts of BASELINE.json's metric on configs[1] / configs[2]:
`nms_us` (1k/5k/20k/100k boxes), `roi_align` (GB/s fwd / bwd, 4096 RoIs x 256 ch x 7x7, P2-P5) and
`single_image` (batch-1 latency of the whole path, eager and CUDA-graph replay).
`roofline` is the dominant kernel of the step, timed with CUDA events around its launches
(detbox_trace_kernel); `cpu_baseline` is the reference's own CPU code (oracle/_ref when it was
built, else the oracle's C port) on the host cores, on a bounded sample of the same workload.
--impl reference times only that CPU path (all host cores) and prints the same line.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

# CP-R: exper/citypersons/retinanet/train_test_resnet50_softmax.yaml:19-46
IH, IW = 1024, 2048
STRIDES = [8, 16, 32, 64, 128]
SCALES = [[2, 2.520, 3.175], [4, 5.040, 6.349], [8, 10.079, 12.700], [16, 20.159, 25.398], [32, 40.317, 50.797]]
RATIOS = [1, 2.5]
BASE = 16
NUM_CLASSES = 2
BG, FG = 0.5, 0.6
TOP_N, NMS_THR = 900, 0.5
A_TOTAL = 261888
GT_PER_IMAGE = 12          # real CityPersons mean ~10-12 GT per image
METRIC, UNIT = "box_ops_images_per_s_1024x2048", "images/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--batch", type=int, default=64, help="images per GPU per step")
    ap.add_argument("--impl", default="detbox", choices=["detbox", "reference"])
    ap.add_argument("--no-extras", action="store_true", help="skip the NMS sweep / RoIAlign sub-benchmarks")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--breakdown", action="store_true", help="also print per-op device times to stderr")
    ap.add_argument("--cpu-roi-child", default=None, help=argparse.SUPPRESS)
    ap.add_argument("--trace", type=int, default=3, help="kernel id timed for the roofline (include/detbox.h)")
    return ap.parse_args()


# ------------------------------------------------------------------------- synthetic inputs ----
def synth_gt(rng, G):
    """pedestrian-shaped boxes, heights log-uniform [12,264] px, aspect
    h/w ~ U(2.0,2.8), centres uniform, 15% ignore regions (class -1)."""
    h = np.exp(rng.uniform(np.log(12), np.log(264), G))
    w = h / rng.uniform(2.0, 2.8, G)
    cx, cy = rng.uniform(0, IW, G), rng.uniform(0, IH, G)
    b = np.stack([cx - w / 2, cy - h / 2, cx + w / 2, cy + h / 2], 1)
    b[:, 0::2] = np.clip(b[:, 0::2], 0, IW - 1)
    b[:, 1::2] = np.clip(b[:, 1::2], 0, IH - 1)
    cls = np.ones(G)
    cls[rng.rand(G) < 0.15] = -1
    return np.concatenate([b, cls[:, None]], 1).astype(np.float32)


def synth_head(rng, n_rows, boost_rows=None):
    """Head outputs for n_rows anchors: deltas ~ N(0, 0.5^2); 2-class softmax probabilities from a
    background-biased foreground logit ~ N(-4.6, 1) (prior bias, libs/nets/head.py:76-79), rows in
    boost_rows ~ N(2, 1)."""
    deltas = rng.normal(0.0, 0.5, (n_rows, 4)).astype(np.float32)
    logit = rng.normal(-4.6, 1.0, n_rows)
    if boost_rows is not None and len(boost_rows):
        logit[boost_rows] = rng.normal(2.0, 1.0, len(boost_rows))
    p1 = (1.0 / (1.0 + np.exp(-logit))).astype(np.float32)
    prob = np.stack([1.0 - p1, p1], 1).astype(np.float32)
    return deltas, prob


def bench_config(B, world):
    """`config` of the JSON line — the SAME dict for both arms (the reference arm runs this workload, B images per step,
    on the host cores)."""
    return {"workload": "BASELINE configs[4] shard: %d CP-R images per GPU per step @1024x2048, 261888 anchors/img, "
                        "G=%d/img, anchor pyramid + anchor-target + per-image top-%d + decode + NMS(%.1f)"
                        % (B, GT_PER_IMAGE, TOP_N, NMS_THR),
            "images_per_gpu": B, "global_images": B * world, "parallelism": "per-image shard x%d, no collective" % world,
            "l2_policy": "inputs larger than L2 (%.0f MB of head outputs per step, never re-used inside a step)"
                         % (B * A_TOTAL * (NUM_CLASSES + 4) * 4 / 1e6)}


# --------------------------------------------------- the reference's CPU path for one image ----
_CPU = {}


def _cpu_setup():
    """Per-process: load the reference's own compiled CPU code (oracle/_ref) when it was built,
    else the oracle's C port.  Only the cpu_baseline / --impl reference legs come here."""
    if "kind" in _CPU:
        return _CPU
    from oracle import ref_loader
    if ref_loader.available():
        r = ref_loader.load()
        c = r.cfg
        c.rpn_bg_threshold, c.rpn_fg_threshold, c.rpn_clobber_positives = BG, FG, True
        c.rpn_sample_strategy, c.rpn_box_encoding = "simple", "fastrcnn"
        c.bbweights, c.ignored_area_intersection_fraction = [1., 1., 1., 1.], 0.3
        _CPU.update(kind="reference", r=r)
    else:
        from oracle import oracle as O
        cfg = O.citypersons_retinanet_cfg()
        _CPU.update(kind="port", O=O, cfg=cfg)
    return _CPU


def _cpu_inputs(seed):
    key = "inputs"                                    # one resident synthetic image per worker process
    if key not in _CPU:
        rng = np.random.RandomState(seed + os.getpid() % 1000)
        gt = synth_gt(rng, GT_PER_IMAGE)
        deltas, prob = synth_head(rng, A_TOTAL, rng.choice(A_TOTAL, 300, replace=False))
        _CPU[key] = (gt, deltas, prob)
    return _CPU[key]


def cpu_one_image(seed):
    """anchor_pyramid + anchor.encode + objectness/top-n/decode + cpu_nms for one image, the
    reference's per-image CPU composition.  Returns seconds (inputs untimed)."""
    c = _cpu_setup()
    gt, deltas, prob = _cpu_inputs(seed)
    t0 = time.perf_counter()
    if c["kind"] == "reference":
        r = c["r"]
        anchors = r.boxes_anchor.anchor_pyramid(STRIDES, IH, IW, SCALES, RATIOS, BASE, [[0.0, 0.0]]).astype(np.float32)
        r.layers_anchor.encode(gt, anchors)                                  # libs/layers/anchor.py:17
        obj = 1.0 - prob[:, 0]                                               # libs/nets/model.py:57
        order = np.argsort(-obj, kind="stable")[:TOP_N]                      # model.py:311-312
        boxes = r.bbox_transform.bbox_transform_inv(anchors[order], deltas[order])   # model.py:318
        dets = np.hstack([boxes, obj[order, None]]).astype(np.float32)
        r.cpu_nms.cpu_nms(dets, NMS_THR)                                     # libs/nms/cpu_nms.pyx:17
    else:
        O, cfg = c["O"], c["cfg"]
        anchors = O.anchor_pyramid(STRIDES, IH, IW, SCALES, RATIOS, BASE).astype(np.float32)
        O.anchor_encode(gt, anchors, cfg)
        boxes, probs, _, _, _ = O.decode_and_choose_top_n(deltas, prob, anchors, TOP_N, "softmax")
        O.nms(boxes, 1.0 - probs[:, 0], NMS_THR, O.NMS_GE_F64)
    return time.perf_counter() - t0


def _cpu_warm(seed):
    _cpu_setup()
    _cpu_inputs(seed)
    return _CPU["kind"]


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


class CpuPool:
    """`cores` worker processes (the reference parallelises this path over DataLoader workers,
    one image per worker call), each with its own resident synthetic image."""

    def __init__(self, cores):
        import multiprocessing as mp
        self.cores = cores
        _cpu_setup()          # in the PARENT, before the fork: oracle/_ref/*.so are mapped here too (the driver's
                              # native-library hook watches this process) and the workers inherit them
        self.pool = mp.get_context("fork").Pool(cores)
        self.kind = self.pool.map(_cpu_warm, range(1000, 1000 + cores), chunksize=1)[0]

    def step(self, images):
        """Process `images` images (round-robin over the workers); returns wall seconds."""
        seeds = [1000 + (i % self.cores) for i in range(images)]
        t0 = time.perf_counter()
        self.pool.map(cpu_one_image, seeds, chunksize=1)
        return time.perf_counter() - t0

    def close(self):
        self.pool.close()
        self.pool.join()


def cpu_baseline_leg(budget_s=12.0):
    """Bounded sample for the main arm's `cpu_baseline`: all host cores, a few rounds."""
    cores = host_cores()
    pool = CpuPool(cores)
    pool.step(cores)                                  # warm-up round
    imgs, wall = 0, 0.0
    while wall < budget_s and imgs < 16 * cores:
        wall += pool.step(cores)
        imgs += cores
    pool.close()
    return {"value": imgs / wall, "unit": UNIT, "cores": cores, "kind": pool.kind,
            "sample": "%d synthetic 1024x2048 images (G=%d, top_n=%d) over %d worker processes, %.1f s wall: "
                      "anchor_pyramid + anchor.encode + top-n/decode + cpu_nms per image" % (imgs, GT_PER_IMAGE, TOP_N, cores, wall)}


# ------------------------------------------- CPU baselines of the two other parts of the metric ----
def sweep_boxes(n):
    """SURVEY 8d config-2 boxes (seed 2): the NMS sweep's inputs, shared by the GPU leg and the CPU baseline."""
    rng = np.random.RandomState(2 + n)
    h = np.exp(rng.uniform(np.log(12), np.log(264), n)); w = h / 2.44
    cx, cy = rng.uniform(0, IW, n), rng.uniform(0, IH, n)
    b = np.stack([np.clip(cx - w / 2, 0, IW - 1), np.clip(cy - h / 2, 0, IH - 1),
                  np.clip(cx + w / 2, 0, IW - 1), np.clip(cy + h / 2, 0, IH - 1)], 1).astype(np.float32)
    return b, rng.permutation(np.linspace(0, 1, n)).astype(np.float32)


def cpu_nms_baseline(sizes=(1000, 5000, 20000, 100000), budget_s=25.0):
    """the reference's compiled Cython cpu_nms (libs/nms/cpu_nms.pyx:17-68, one thread: the loop is
    serial) on the sweep's boxes.  Bounded: a size is skipped once the budget is spent."""
    c = _cpu_setup()
    out, spent, last = {}, 0.0, None
    for n in sizes:
        if last is not None:                          # greedy NMS costs ~ n * kept ~ n^2 on this box distribution
            predict = last[1] * (n / last[0]) ** 2
            if spent + predict > budget_s:
                out[str(n)] = {"skipped": "bounded sample: predicted %.0f s from the n = %d run (cost ~ n^2), budget %.0f s" % (predict, last[0], budget_s),
                               "extrapolated_us": round(1e6 * predict, 0)}
                continue
        b, s = sweep_boxes(n)
        dets = np.ascontiguousarray(np.concatenate([b, s[:, None]], 1), np.float32)
        reps = 3 if n <= 5000 else 1
        ts = []
        for _ in range(reps):
            t0 = time.perf_counter()
            if c["kind"] == "reference":
                keep = c["r"].cpu_nms.cpu_nms(dets, NMS_THR)
            else:
                keep = c["O"].nms(b, s, NMS_THR, c["O"].NMS_GE_F64)
            ts.append(time.perf_counter() - t0)
        spent += sum(ts)
        last = (n, float(np.median(ts)))
        out[str(n)] = {"us": round(1e6 * float(np.median(ts)), 1), "kept": int(len(keep))}
    out["kind"], out["cores"] = c["kind"], 1
    out["what"] = "libs/nms/cpu_nms.pyx cpu_nms(dets, 0.5) ('>=' in float64), serial by construction"
    return out


def _cpu_roi_child(with_backward):
    """Child process of cpu_roi_align_baseline (OMP_NUM_THREADS is set in its environment, the only way to be sure which
    OpenMP runtime instance the reference's .so is bound to obeys it): prints one JSON line."""
    from oracle import oracle as O
    c = _cpu_setup()
    if c["kind"] != "reference":
        print(json.dumps({"unavailable": "oracle/_ref (compiled crop_and_resize.c) not built"})); return
    r = c["r"]
    _, _, P, strides, shapes, ref_areas, rois = roi_bench_inputs()
    levels = O.assign_levels(rois[:, 1:], np.asarray(ref_areas, np.float32))
    gen = np.random.default_rng(3)
    t_f = t_b = 0.0
    for l, (st, shp) in enumerate(zip(strides, shapes)):
        sel = levels == l
        if not sel.any():
            continue
        feat = gen.standard_normal(shp, dtype=np.float32)
        boxes, ind = O.tf_boxes(rois[sel], np.float32(1.0 / st), P, P, shp[2], shp[3])
        t0 = time.perf_counter()
        crops = r.crop_and_resize_forward(feat, boxes, ind, 0.0, P, P)
        t_f += time.perf_counter() - t0
        if with_backward:
            t0 = time.perf_counter()
            r.crop_and_resize_backward(crops, boxes, ind, shp)
            t_b += time.perf_counter() - t0
        del feat, crops
    print(json.dumps({"fwd_ms": round(1e3 * t_f, 1), "bwd_ms": round(1e3 * t_b, 1) if with_backward else None}))


def cpu_roi_align_baseline():
    """the reference's crop_and_resize.c (libs/layers/roi_align_tf/src/crop_and_resize.c:30 forward,
    `#pragma omp parallel for` over boxes; :207 backward, serial) on configs[2] — the same 4096 RoIs as the GPU leg,
    level by level as PyramidRoIAlign does (pyramid_roi_align.py:30-60).  Forward with OMP_NUM_THREADS = 1 and = all
    cores (one child process each), backward once."""
    cores = host_cores()
    res = {}
    for tag, nthr, bwd in (("omp1", 1, True), ("omp_all", cores, False)):
        env = dict(os.environ, OMP_NUM_THREADS=str(nthr))
        try:
            p = subprocess.run([sys.executable, os.path.abspath(__file__), "--cpu-roi-child", "1" if bwd else "0"], env=env,
                               capture_output=True, text=True, timeout=600)
            d = json.loads(p.stdout.strip().splitlines()[-1])
        except Exception as e:                          # keep the bench line; say why the row is missing
            return {"unavailable": "cpu RoIAlign child failed: %s" % str(e)[:120]}
        if "unavailable" in d:
            return d
        res["fwd_ms_" + tag] = d["fwd_ms"]
        if bwd:
            res["bwd_ms_serial"] = d["bwd_ms"]
    res.update(cores=cores, kind="reference",
               what="crop_and_resize.c forward (OpenMP over boxes; OMP_NUM_THREADS = 1 / = cores) and backward (serial) on the configs[2] "
                    "RoIs, level by level; times include the wrapper's zero fill of the output / dense gradient map")
    return res


# ------------------------------------------------------------------------------- GPU helpers ----
class ClockSampler:
    """nvidia-smi clocks / throttle reasons of one GPU, sampled every 200 ms while active."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.proc, self.thread = [], None, None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None
        self.mark0 = self.mark1 = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), [x.strip() for x in line.split(",")]))

    def begin(self):
        self.mark0 = time.time()

    def end(self):
        self.mark1 = time.time()

    def stop(self):
        if self.proc:
            self.proc.terminate()
        rows = [r for t, r in self.rows if self.mark0 is None or (self.mark0 - 0.25 <= t <= (self.mark1 or t) + 0.25)]
        rows = rows or [r for _, r in self.rows]
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in rows:
            try:
                sm.append(float(r[0])); mx.append(float(r[1]))
            except (ValueError, IndexError):
                continue
            for n, v in zip(names, r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": max(mx), "reasons": sorted(reasons), "samples": len(sm)}


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


def event_pairs(torch, n):
    return [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(n)]


def timed_loop(torch, dist, fn, steps, warmup, before_step=None):
    """W untimed + K timed calls of fn(); barrier + synchronize on both sides; device time by CUDA
    events on the launching stream; returns the max over ranks in milliseconds."""
    for i in range(warmup):
        if before_step:
            before_step(-1)
        fn()
    torch.cuda.synchronize()
    if dist is not None:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(steps):
        if before_step:
            before_step(i)
        fn()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    if dist is not None:
        t = torch.tensor([ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
        dist.barrier()
    return ms


# --------------------------------------------------------------- extras: NMS sweep, RoIAlign ----
def median_ms(torch, fn, iters=10, warm=3, flush=None):
    for _ in range(warm):
        fn()
    ts = []
    for _ in range(iters):
        if flush is not None:
            flush.add_(1.0)                          # > L2-sized write between timed iterations
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return float(np.median(ts))


def nms_sweep(torch, detbox):
    """BASELINE configs[1]: n in {1k,5k,20k,100k}, IoU 0.5, SURVEY §8d config-2 boxes (sweep_boxes).  Beside every
    device-resident number: the reference's OWN CUDA path (libs/nms/nms_kernel.cu compiled unmodified for sm_100a,
    oracle/_ref/libref_cuda_nms.so, called through its launcher _nms: host dets in, cudaMalloc + copies + kernel + host
    scan inside, gpu_nms.pyx:16-31) at 20k / 100k, and this library's drop-in for the same symbol — the same-box GPU
    baseline of SURVEY 8c."""
    import ctypes
    out = {}
    ref = None
    p = os.path.join(ROOT, "oracle", "_ref", "libref_cuda_nms.so")
    if os.path.exists(p):
        ref = ctypes.CDLL(p)._Z4_nmsPiS_PKfiifi
        ref.restype = None
        ref.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_float, ctypes.c_int]
    ours = detbox.lib()._nms
    for n in (1000, 5000, 20000, 100000):
        b, s = sweep_boxes(n)
        db, ds = torch.from_numpy(b).cuda(), torch.from_numpy(s).cuda()
        ms = median_ms(torch, lambda: detbox.nms(db, ds, NMS_THR, detbox.NMS_GT_F32), iters=7 if n > 20000 else 15)
        keep, nk = detbox.nms(db, ds, NMS_THR, detbox.NMS_GT_F32)
        out[str(n)] = {"us": round(ms * 1e3, 1), "kept": int(nk[-1].item())}
        if n >= 20000:
            order = np.argsort(-s, kind="stable")
            dets = np.ascontiguousarray(np.concatenate([b[order], s[order, None]], 1), np.float32)
            for name, fn in (("reference_gpu_nms_host_us", ref), ("detbox_nms_host_us", ours)):
                if fn is None:
                    continue
                kp, num = np.zeros(n, np.int32), np.zeros(1, np.int32)
                ts = []
                for _ in range(3):
                    torch.cuda.synchronize()
                    t0 = time.perf_counter()
                    fn(kp.ctypes.data, num.ctypes.data, dets.ctypes.data, n, 5, ctypes.c_float(NMS_THR), torch.cuda.current_device())
                    ts.append(time.perf_counter() - t0)
                out[str(n)][name] = round(1e6 * float(np.median(ts)), 1)
                out[str(n)][name.replace("_us", "_kept")] = int(num[0])
    return out


def measure_l2_peak(torch):
    """the L2 denominator — device-to-device copy over a resident set (2 x 24 MB, inside the 126 MB L2),
    read + write bytes over CUDA-event time, best of 5 batches of 20 copies (the way MEASURED_PEAKS.json's HBM figure
    was taken, on a set that never leaves L2)."""
    n = 6 * 1024 * 1024
    a = torch.randn(n, device="cuda"); b = torch.empty_like(a)
    for _ in range(5):
        b.copy_(a)
    best = 0.0
    for _ in range(5):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(20):
            b.copy_(a)
        e1.record(); torch.cuda.synchronize()
        best = max(best, 20 * 2 * n * 4 / (e0.elapsed_time(e1) * 1e-3) / 1e9)
    return round(best, 1)


def roi_bench_inputs():
    """BASELINE configs[2]: 512 RoIs/img x 256 ch x 7x7 over P2-P5 of a 1024x2048 canvas, batch 8 — shapes and RoIs,
    shared by the GPU leg and the CPU baseline."""
    B, C, P = 8, 256, 7
    strides = [4, 8, 16, 32]
    shapes = [(B, C, IH // s, IW // s) for s in strides]
    ref_areas = [(float(np.mean(sc)) * BASE) ** 2 for sc in SCALES[:4]]
    rng = np.random.RandomState(3)
    g = np.concatenate([synth_gt(rng, 512)[:, :4] for _ in range(B)], 0)
    g += rng.normal(0, 4.0, g.shape).astype(np.float32)
    g[:, 2:] = np.maximum(g[:, 2:], g[:, :2] + 4)
    rois = np.concatenate([np.repeat(np.arange(B), 512)[:, None].astype(np.float32), g], 1).astype(np.float32)
    return B, C, P, strides, shapes, ref_areas, rois


def roi_align_bench(torch, detbox, hbm_peak):
    """BASELINE configs[2]: 512 RoIs/img x 256 ch x 7x7 over P2-P5 of a 1024x2048 canvas, batch 8,
    fp32, bin-centre sampling (the variant MaskRCNN wires in), level assigned in-kernel."""
    B, C, P, strides, shapes, ref_areas, rois = roi_bench_inputs()
    R = rois.shape[0]
    d_rois = torch.from_numpy(rois).cuda()
    gen = torch.Generator(device="cuda").manual_seed(3)
    res = {}
    lv = detbox.assign_levels(d_rois[:, 1:5], ref_areas)
    # compulsory feature bytes = unique 32-byte sectors the taps touch: in NHWC a
    # position's C floats are contiguous (uniq positions x C x 4 B); in NCHW a sector holds 8
    # consecutive x of one (b,c,y) row (uniq (b,y,x//8) x C x 32 B)
    uniq, uniq_sec, uniq_g64 = 0, 0, 0
    for l, s in enumerate(strides):
        sel = (lv == l).nonzero().flatten()
        if sel.numel() == 0:
            continue
        r = d_rois[sel]; H, W = shapes[l][2], shapes[l][3]
        sc = 1.0 / s
        j = torch.arange(P, device="cuda", dtype=torch.float32)[None]
        sw = (r[:, 3:4] - r[:, 1:2]) * sc / P; sh = (r[:, 4:5] - r[:, 2:3]) * sc / P
        xs = r[:, 1:2] * sc + sw / 2 - 0.5 + j * sw; ys = r[:, 2:3] * sc + sh / 2 - 0.5 + j * sh
        xt = torch.cat([xs.floor(), xs.ceil()], 1).clamp(0, W - 1).long(); yt = torch.cat([ys.floor(), ys.ceil()], 1).clamp(0, H - 1).long()
        flat = (r[:, 0].long()[:, None, None] * H + yt[:, :, None]) * W + xt[:, None, :]
        m = torch.zeros(B * H * W, dtype=torch.bool, device="cuda"); m[flat.flatten()] = True
        uniq += int(m.sum().item())
        uniq_sec += int(m.view(B * H, W // 8, 8).any(dim=2).sum().item())
        uniq_g64 += int(m.view(B * H, W // 16, 16).any(dim=2).sum().item())      # DRAM moves 64-byte granules (2 sectors)
    out_bytes = R * C * P * P * 4
    feat_bytes_by_layout = {"nhwc": uniq * C * 4, "nchw": uniq_sec * C * 32}
    dense_bytes = sum(int(np.prod(s)) for s in shapes) * 4
    flush = torch.empty(64 * 1024 * 1024, dtype=torch.float32, device="cuda").zero_()      # 256 MB > 126 MB L2
    for layout in ("nchw", "nhwc"):
        feats = [torch.randn(s, generator=gen, device="cuda", dtype=torch.float32) for s in shapes]
        if layout == "nhwc":
            feats = [f.contiguous(memory_format=torch.channels_last) for f in feats]
        scales = [1.0 / s for s in strides]
        fwd = lambda: detbox.roi_align_forward(feats, d_rois, scales, P, P, detbox.ROI_BIN_CENTRE, ref_areas=ref_areas)
        y = fwd()
        gout = torch.randn(y.shape, generator=gen, device="cuda", dtype=torch.float32)
        if layout == "nhwc":
            gout = gout.contiguous(memory_format=torch.channels_last)
        grads = [torch.empty_like(f) for f in feats]
        bwd_acc = lambda: detbox.roi_align_backward(gout, shapes, d_rois, scales, detbox.ROI_BIN_CENTRE, ref_areas=ref_areas,
                                                    channels_last=(layout == "nhwc"), grads=grads)
        bwd_full = lambda: detbox.roi_align_backward(gout, shapes, d_rois, scales, detbox.ROI_BIN_CENTRE, ref_areas=ref_areas,
                                                     channels_last=(layout == "nhwc"))
        t_f = median_ms(torch, fwd, flush=flush)
        t_b = median_ms(torch, bwd_full, flush=flush)
        t_s = median_ms(torch, bwd_acc, flush=flush)
        feat_bytes = feat_bytes_by_layout[layout]
        fb, bb = out_bytes + feat_bytes + R * 20, out_bytes + dense_bytes + feat_bytes + R * 20
        g64 = out_bytes + uniq_g64 * C * 64 + R * 20 if layout == "nchw" else None
        res[layout] = {"fwd_ms": round(t_f, 4), "fwd_gbs": round(fb / t_f / 1e6, 1), "fwd_frac_hbm": round(fb / t_f / 1e6 / hbm_peak, 3),
                       "bwd_ms": round(t_b, 4), "bwd_gbs": round(bb / t_b / 1e6, 1), "bwd_frac_hbm": round(bb / t_b / 1e6 / hbm_peak, 3),
                       "bwd_scatter_only_ms": round(t_s, 4)}
        if layout == "nchw":
            # bwd_ms is the tile-centric dense backward (every gradient map written once: no memset, no atomics, deterministic);
            # the RoI-centric path it replaced as the default — memset of the maps + separable scatter — is timed beside it
            t_bs = median_ms(torch, lambda: detbox.roi_align_backward(gout, shapes, d_rois, scales, detbox.ROI_BIN_CENTRE, ref_areas=ref_areas,
                                                                      deterministic=False), flush=flush)
            res[layout]["bwd_memset_scatter_ms"] = round(t_bs, 4)
            res[layout]["bwd_memset_scatter_frac_hbm"] = round(bb / t_bs / 1e6 / hbm_peak, 3)
            res[layout]["bwd_kernel"] = "roi_bwd_nchw_tile_kernel (detbox_roi_align_backward_tiled): 1.41 GB written + 0.21 GB read per launch (ncu) against 2.9 GB for memset + scatter"
        if g64:
            res[layout]["fwd_frac_hbm_64B_granules"] = round(g64 / t_f / 1e6 / hbm_peak, 3)
            res[layout]["note"] = ("fwd_frac_hbm counts the unique 32-byte sectors the taps touch; DRAM fetches 64-byte granules, and a NCHW window row is "
                                   "48-64 bytes at a 2 KB stride: fwd_frac_hbm_64B_granules counts those.  ncu (profiles/r02*_ncu_roi.txt): the staged forward "
                                   "moves 1.13 GB read + 0.19 GB written in 232 us = 5.7 TB/s, 0.87 of the measured peak")
        del feats, grads, gout, y
    # L2-level view of the forward gather: 4 taps x 4 B per output element (SURVEY 8d upper bound) + the output write
    l2_peak = measure_l2_peak(torch)
    gather = out_bytes * 4 + out_bytes
    res["l2"] = {"peak_gbs": l2_peak, "how": "copy over a 48 MB resident set (read + write), best of 5 x 20",
                 "fwd_gather_bytes": gather,
                 "nchw_fwd_gather_gbs": round(gather / res["nchw"]["fwd_ms"] / 1e6, 1),
                 "nhwc_fwd_gather_gbs": round(gather / res["nhwc"]["fwd_ms"] / 1e6, 1),
                 "note": "fwd_gather_bytes is the 4-tap upper bound: taps shared by neighbouring bins are L1 hits, so the L2-level "
                         "rate is below these figures (ncu lts__t_bytes in profiles/r01_ncu_roi.txt is the measured one)"}
    res["bytes"] = {"out": out_bytes, "features_touched": feat_bytes_by_layout, "features_touched_nchw_64B_granules": uniq_g64 * C * 64,
                    "dense_grad_zero_fill": dense_bytes}
    res["workload"] = "B=8, R=4096, C=256, 7x7, P2-P5 @1024x2048, bin-centre, level assigned in-kernel, fp32; L2 flushed between iterations"
    return res


def coco_post_bench(torch, detbox, hbm_peak):
    """BASELINE configs[3]: COCO-shape 800x1333 RetinaNet post-processing, A = 132,450 anchors/img, batch 16, 80 sigmoid
    classes: objectness -> top-k -> decode -> per-image NMS (0.5), inputs resident in HBM.  Two budgets: 1000 per image
    and the reference's global 900*B (libs/nets/model.py:311-312).  Algorithmic bytes ~ 4*C*A per image (the class
    scores are read once; SURVEY 8d: 678 MB per batch)."""
    B, A, C = 16, 132450, 80
    gen = torch.Generator(device="cuda").manual_seed(5)
    prob = torch.sigmoid(torch.randn((B * A, C), generator=gen, device="cuda") - 4.6)
    deltas = torch.randn((B * A, 4), generator=gen, device="cuda") * 0.5
    xy = torch.rand((A, 2), generator=gen, device="cuda") * torch.tensor([1333.0, 800.0], device="cuda")
    wh = torch.exp(torch.rand((A, 2), generator=gen, device="cuda") * 3.0 + 2.5)
    anc = torch.cat([xy - wh / 2, xy + wh / 2], 1).contiguous()
    out = {"workload": "B=16, A=132450, C=80 sigmoid, fp32; inputs %.0f MB > L2" % ((prob.numel() + deltas.numel()) * 4 / 1e6)}
    for name, top_n, per_image in (("per_image_1000", 1000, True), ("global_14400", 900 * B, False)):
        def step():
            sel = detbox.topk_decode(deltas, prob, anc, top_n, "sigmoid", "fastrcnn", per_image, want=("img_ids", "scores"))
            return detbox.nms(sel["boxes"], sel["scores"], NMS_THR, detbox.NMS_GT_F32, seg_ids=sel["img_ids"], n_seg=B,
                              presorted=per_image)
        t_topk = median_ms(torch, lambda: detbox.topk_decode(deltas, prob, anc, top_n, "sigmoid", "fastrcnn", per_image,
                                                             want=("img_ids", "scores")))
        t = median_ms(torch, step)
        # the score read on its own: events around objectness_kernel (C = 80 instantiation) inside one top-k call
        ea, eb = event_pairs(torch, 1)[0]
        ea.record(); eb.record()
        detbox.lib().detbox_trace_kernel(5, ea.cuda_event, eb.cuda_event)
        detbox.topk_decode(deltas, prob, anc, top_n, "sigmoid", "fastrcnn", per_image, want=("img_ids", "scores"))
        torch.cuda.synchronize()
        detbox.lib().detbox_trace_kernel(0, None, None)
        t_obj = ea.elapsed_time(eb)
        out["objectness_kernel"] = {"ms": round(t_obj, 4), "gbs": round(prob.numel() * 4 / t_obj / 1e6, 1),
                                    "frac_hbm": round(prob.numel() * 4 / t_obj / 1e6 / hbm_peak, 3),
                                    "note": "the 678 MB score read itself; score_read_frac_hbm below divides the same bytes by the whole top-k + decode time "
                                            "(select, partition, candidate select, sort, gather: latency-bound tail, profiles/r02_coco_path.txt)"}
        out[name] = {"ms": round(t, 4), "topk_decode_ms": round(t_topk, 4), "images_per_s": round(B / (t / 1e3), 1),
                     "score_read_gbs": round(prob.numel() * 4 / t_topk / 1e6, 1),
                     "score_read_frac_hbm": round(prob.numel() * 4 / t_topk / 1e6 / hbm_peak, 3)}
    return out


# ------------------------------------------------- BASELINE configs[4] as written: strong scaling ----
def strong_64_leg(torch, dist, detbox, rank, world, steps, warmup, n_images=64, max_dets=100):
    """ONE batch of 64 CP-R images dealt round-robin over the GPUs (image i -> GPU i mod N, detbox.parallel.shard_images;
    SURVEY 8d config 5), every rank running the device-resident step on ITS 64/N images (one CUDA-graph replay), then
    the optional epilogue inside the timed region: the kept detections packed into fixed [64/N, 100, 6] buffers
    (detbox_pack_detections) and all-gathered over NCCL (detbox.parallel.gather_detections_device), so every rank
    ends the step with all 64 images' detections.  Images are seeded by their GLOBAL index: the batch is the same
    for every N.  Time = CUDA events, barrier on both sides, max over ranks."""
    P = detbox.parallel
    mine = P.shard_images(n_images, rank, world)
    Bl = len(mine)
    sess = detbox.pipeline.BoxOpsSession(STRIDES, IH, IW, SCALES, RATIOS, BASE, Bl, NUM_CLASSES, top_n=TOP_N, nms_thresh=NMS_THR,
                                         bg=BG, fg=FG, max_gt_total=GT_PER_IMAGE * Bl)
    A = sess.A
    gts = [synth_gt(np.random.RandomState(100 + i), GT_PER_IMAGE) for i in mine]
    G = sess.stage_gt(gts)
    d_gt, d_off = sess.h_gt[:G].cuda(), sess.h_off.cuda()
    lab = detbox.anchor_target(sess.anchors, d_gt, d_off, sess.tcfg)[0].cpu().numpy()
    d_prob = torch.empty((Bl * A, NUM_CLASSES), dtype=torch.float32, device="cuda")
    d_box = torch.empty((Bl * A, 4), dtype=torch.float32, device="cuda")
    for j, i in enumerate(mine):
        deltas, prob = synth_head(np.random.RandomState(7000 + i), A, np.nonzero(lab[j] != 0)[0])
        d_box[j * A:(j + 1) * A].copy_(torch.from_numpy(deltas)); d_prob[j * A:(j + 1) * A].copy_(torch.from_numpy(prob))
    del lab
    buf = torch.empty((Bl, max_dets, 6), dtype=torch.float32, device="cuda")
    cnt = torch.empty((Bl,), dtype=torch.int32, device="cuda")
    gathered = (torch.empty((world, Bl, max_dets, 6), dtype=torch.float32, device="cuda"),
                torch.empty((world, Bl), dtype=torch.int32, device="cuda")) if world > 1 else None
    state = {}

    def step():
        out = sess.run_device_graph(d_gt, d_off, d_prob, d_box)
        P.pack_detections_device(out["boxes"], out["scores"], out["keep"], out["n_keep"], Bl, max_dets, out=(buf, cnt))
        state["all"] = P.gather_detections_device(buf, cnt, n_images, out=gathered)

    ms = timed_loop(torch, dist, step, steps, warmup)
    ms_nogather = timed_loop(torch, dist, lambda: sess.run_device_graph(d_gt, d_off, d_prob, d_box), steps, warmup)
    all_buf, all_cnt = state["all"]
    torch.cuda.synchronize()
    res = {"images": n_images, "images_per_gpu": Bl, "ms_per_step": round(ms / steps, 4),
           "images_per_s": round(n_images * steps / (ms / 1e3), 1),
           "ms_per_step_without_gather": round(ms_nogather / steps, 4),
           "gather": ("NCCL all_gather_into_tensor of [%d,%d,6] f32 + counts per rank, inside the timed region" % (Bl, max_dets)) if world > 1
                     else "none (1 GPU): pack kernel only",
           "gather_bytes_per_rank": int(buf.numel() * 4 + cnt.numel() * 4),
           "detections_total": int(all_cnt.sum().item()), "checksum": round(float(all_buf.double().sum().item()), 3),
           "note": "strong scaling: total work fixed at 64 images; detections_total / checksum must be the same for every N"}
    del sess, d_prob, d_box
    return res


def g_sweep_leg(torch, dist, detbox, L, steps, warmup, hbm_peak, d_prob, d_box, B):
    """the step at G in {1, 12, 32, 99} synthetic GT per image and on one batch of 64 REAL CityPersons
    validation GT sets (tests/golden/real_gt64.npz: G = 0 ... 46, mean 12.2, 475 ignore regions).  Per point: the
    device-resident step (graph replay) and the dominant kernel (events around its launches in an eager one-stream run)."""
    out = {}
    real = np.load(os.path.join(ROOT, "tests", "golden", "real_gt64.npz"))
    roff = real["offsets"]
    cases = [("G=%d" % g, [synth_gt(np.random.RandomState(100 + i), g) for i in range(B)]) for g in (1, 12, 32, 99)]
    if B <= 64:
        cases.append(("real_anno_val_64", [real["gt"][roff[i]:roff[i + 1]] for i in range(B)]))
    sess = detbox.pipeline.BoxOpsSession(STRIDES, IH, IW, SCALES, RATIOS, BASE, B, NUM_CLASSES, top_n=TOP_N, nms_thresh=NMS_THR,
                                         bg=BG, fg=FG, max_gt_total=100 * B)
    A = sess.A
    for name, gts in cases:
        Gt = sess.stage_gt(gts)
        d_gt, d_off = sess.h_gt[:Gt].cuda(), sess.h_off.cuda()
        ms = timed_loop(torch, dist, lambda: sess.run_device_graph(d_gt, d_off, d_prob, d_box), steps, warmup)
        pairs = event_pairs(torch, steps)
        for a, b in pairs:
            a.record(); b.record()

        def arm(i):
            if i >= 0:
                L.detbox_trace_kernel(3, pairs[i][0].cuda_event, pairs[i][1].cuda_event)
            else:
                L.detbox_trace_kernel(0, None, None)
        timed_loop(torch, dist, lambda: sess.run_device(d_gt, d_off, d_prob, d_box, overlap=False), steps, warmup, before_step=arm)
        L.detbox_trace_kernel(0, None, None)
        k_ms = float(np.mean([a.elapsed_time(b) for a, b in pairs]))
        tb = trace_bytes(3, B, A, Gt, NUM_CLASSES)
        out[name] = {"total_gt": int(Gt), "ms_per_step": round(ms / steps, 4), "images_per_s": round(B * steps / (ms / 1e3), 1),
                     "anchor_target_kernel_ms": round(k_ms, 4), "roofline_frac": round(tb / k_ms / 1e6 / hbm_peak, 4)}
    del sess
    return out


# -------------------------------------------------------------------------------------- main ----
def single_image_latency(torch, detbox, gt1, h_prob, h_box, A):
    """one CP-R image per call (the launch-latency regime) — the whole device-resident path
    eagerly launched, and the same launch sequence captured once in a CUDA graph and replayed."""
    sess1 = detbox.pipeline.BoxOpsSession(STRIDES, IH, IW, SCALES, RATIOS, BASE, 1, NUM_CLASSES, top_n=TOP_N, nms_thresh=NMS_THR,
                                          bg=BG, fg=FG, max_gt_total=GT_PER_IMAGE)
    G = sess1.stage_gt([gt1])
    d_gt, d_off = sess1.h_gt[:G].cuda(), sess1.h_off.cuda()
    d_prob, d_box = h_prob[:A].cuda(), h_box[:A].cuda()
    fn = lambda: sess1.run_device(d_gt, d_off, d_prob, d_box)
    n0 = detbox.launches(); ref = fn(); per_call = detbox.launches() - n0
    out = {"launches_per_image": int(per_call), "eager_us": round(median_ms(torch, fn, iters=40, warm=5) * 1e3, 1)}
    try:
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(3):
                fn()
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            got = fn()
        graph.replay(); torch.cuda.synchronize()
        same = all(torch.equal(got[k], ref[k]) for k in ("labels", "bbox_targets", "n_keep")) and \
            torch.equal(got["keep"][:int(got["n_keep"][-1])], ref["keep"][:int(ref["n_keep"][-1])])
        out["graph_us"] = round(median_ms(torch, graph.replay, iters=40, warm=5) * 1e3, 1)
        out["graph_matches_eager"] = bool(same)
        out["images_per_s_graph"] = round(1e6 / out["graph_us"], 1)
    except Exception as e:                             # keep the bench line; say why there is no graph number
        out["graph_error"] = str(e).splitlines()[0][:160]
    return out


def trace_bytes(kid, B, A, G_total, C):
    """ALGORITHMIC bytes of ONE launch of the traced kernel (kernels table)."""
    return {2: 16 * A * B + 20 * G_total,                       # colmax: anchors (re-read per image) + GT
            3: 16 * A + B * 36 * A + 20 * G_total,              # assign: anchors once, labels+targets+inside per image
            4: B * 8 * A,                                       # weights: labels read + label_weights written
            5: B * A * (4 * C + 4)}.get(kid)                    # objectness: probabilities read + keys written


def run_reference(args):
    """--impl reference: the reference's own CPU implementation of the step on all host cores.  A step is the SAME
    batch the GPU arm processes per GPU (--batch images, default 64), one image per worker call, round-robin over
    one worker process per core (the reference parallelises this path over DataLoader workers)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = host_cores()
    pool = CpuPool(cores)
    per_step = args.batch
    for _ in range(args.warmup):
        pool.step(per_step)
    wall = sum(pool.step(per_step) for _ in range(args.steps))
    pool.close()
    v = per_step * args.steps / wall
    line = {"impl": "reference", "metric": METRIC, "value": round(v, 3), "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(1e3 * wall / args.steps, 3),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64 IoU / f32 boxes",
            "data": "synthetic",
            "config": bench_config(per_step, max(1, args.gpus)),
            "reference_path": "oracle/_ref (the reference's own Cython / NumPy compiled in place): anchor_pyramid + layers.anchor.encode + "
                              "objectness argsort + bbox_transform_inv on the top-%d rows only (conservative: single_shot.compute_detection "
                              "decodes all 261888 rows) + cpu_nms; %d images per step over %d worker processes" % (TOP_N, per_step, cores),
            "cpu_baseline": {"value": round(v, 3), "unit": UNIT, "cores": cores, "kind": pool.kind,
                             "sample": "%d images per step x %d steps, one image per worker call, %d worker processes" % (per_step, args.steps, cores)},
            "e2e": {"value": round(v, 3), "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def main():
    args = parse_args()
    if args.cpu_roi_child is not None:
        return _cpu_roi_child(args.cpu_roi_child == "1")
    if args.impl == "reference":
        return run_reference(args)

    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    cpu_base = cpu_nms = cpu_roi = None
    if rank == 0 and world == 1 and not args.no_cpu:
        cpu_base = cpu_baseline_leg()                 # before CUDA is initialised (fork pool)
        if not args.no_extras:                        # the two other parts of the metric, on the same host cores
            cpu_nms = cpu_nms_baseline()
            cpu_roi = cpu_roi_align_baseline()        # (the cpu_baseline leg is the one place bench.py may use oracle/)

    import torch
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: detbox has no CPU path")
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    import detbox
    pipeline = detbox.pipeline
    L = detbox.lib()
    hbm_peak, peak_src = measured_peaks()

    B, A, C = args.batch, A_TOTAL, NUM_CLASSES
    sess = pipeline.BoxOpsSession(STRIDES, IH, IW, SCALES, RATIOS, BASE, B, C, top_n=TOP_N, nms_thresh=NMS_THR,
                                  bg=BG, fg=FG, max_gt_total=GT_PER_IMAGE * B)
    assert sess.A == A
    # synthetic inputs: images (rank*B .. rank*B+B-1), seeds 100+
    gts = [synth_gt(np.random.RandomState(100 + rank * B + i), GT_PER_IMAGE) for i in range(B)]
    G_total = sess.stage_gt(gts)
    d_gt, d_off = sess.h_gt[:G_total].cuda(), sess.h_off.cuda()
    lab = detbox.anchor_target(sess.anchors, d_gt, d_off, sess.tcfg)[0]
    boost = (lab.flatten() != 0).nonzero().flatten().cpu().numpy()        # fg + ignored anchors score high
    rng = np.random.RandomState(7 + rank)
    deltas, prob = synth_head(rng, B * A, boost)
    h_prob, h_box = torch.from_numpy(prob).pin_memory(), torch.from_numpy(deltas).pin_memory()
    d_prob, d_box = h_prob.cuda(), h_box.cuda()
    del lab, deltas, prob
    torch.cuda.synchronize()

    sampler = ClockSampler(local) if rank == 0 else None
    # ---- device-resident leg -------------------------------------------------------------------
    pairs = event_pairs(torch, args.steps)
    for a, b in pairs:
        a.record(); b.record()                       # materialise the cudaEvent_t handles

    def arm(i):
        if i >= 0 and args.trace:
            L.detbox_trace_kernel(args.trace, pairs[i][0].cuda_event, pairs[i][1].cuda_event)
        else:
            L.detbox_trace_kernel(0, None, None)

    # value: the step as the session ships it — ONE CUDA-graph replay per step, the anchor-target chain and the
    # detection chain on two streams inside the graph (pipeline.run_device_graph).  The same step launched eagerly
    # (two streams / one stream) is timed beside it; the one-stream eager run is also where the dominant kernel is
    # timed for the roofline (CUDA events around each of its launches, nothing else running beside it).
    step_dev = lambda: sess.run_device_graph(d_gt, d_off, d_prob, d_box)
    step_eager = lambda: sess.run_device(d_gt, d_off, d_prob, d_box)
    step_serial = lambda: sess.run_device(d_gt, d_off, d_prob, d_box, overlap=False)
    L.detbox_trace_kernel(0, None, None)
    if sampler:
        sampler.begin()
    ms_dev = timed_loop(torch, dist, step_dev, args.steps, args.warmup)
    launches = int(sess.graph_launches) * args.steps
    ms_eager = timed_loop(torch, dist, step_eager, args.steps, args.warmup)
    ms_serial = timed_loop(torch, dist, step_serial, args.steps, args.warmup, before_step=arm)
    L.detbox_trace_kernel(0, None, None)
    k_ms = float(np.mean([a.elapsed_time(b) for a, b in pairs])) if args.trace else None
    # ---- end-to-end leg: pinned host buffers in, host buffers out --------------------------------
    # e2e: the session's compact host format (packed label byte + float32 targets + per-image counts: 17 B / anchor
    # instead of the reference's 40, same information, detbox.expand_targets rebuilds the four arrays bit for bit) and the
    # regression deltas gathered from the pinned host buffer (only the top-900 rows per image cross the link).
    def e2e_leg(fn, note=None):
        ms = timed_loop(torch, dist, fn, args.steps, args.warmup)
        d = {"value": round(B * world * args.steps / (ms / 1e3), 1), "unit": UNIT, "ms_per_step": round(ms / args.steps, 4),
             "h2d_bytes_per_step": int(sess.h2d_bytes), "d2h_bytes_per_step": int(sess.d2h_bytes)}
        if note:
            d["note"] = note
        return d
    e2e = e2e_leg(lambda: sess.run_host(gts, h_prob, h_box, format="compact"))
    e2e["format"] = "compact: uint8 (label+1 | inside-weight flag<<7) [B,A] + bbox_targets f32 [B,A,4] + counts i32 [B,2]; rpn_box rows gathered from pinned host memory"
    if sampler:
        sampler.end()
    clocks = sampler.stop() if sampler else None
    e2e_more = {}
    if not args.no_extras:
        e2e_more["e2e_reference_format"] = e2e_leg(lambda: sess.run_host(gts, h_prob, h_box, format="reference", box_rows="copy"),
                                                   "round-1 path: the reference's four float32 arrays (40 B / anchor) copied back, all of rpn_box staged on the device")
        e2e_more["e2e_reference_format_mapped_rows"] = e2e_leg(lambda: sess.run_host(gts, h_prob, h_box, format="reference"),
                                                               "the four float32 arrays copied back; rpn_box rows gathered from pinned host memory")
        ncore = host_cores()
        e2e_more["e2e_reference_arrays"] = e2e_leg(lambda: sess.run_host(gts, h_prob, h_box, format="reference", wire="compact"),
                                                   "the reference's four float32 arrays delivered on the HOST: every chunk crosses the link in the compact format and "
                                                   "detbox_expand_targets_host materialises them on worker threads while later chunks are in flight (inside the timed region)")
        e2e_more["e2e_compact_expanded_serial"] = e2e_leg(lambda: sess.expand(sess.run_host(gts, h_prob, h_box, format="compact"), n_threads=ncore),
                                                          "the same result with the expansion run AFTER the transfers (%d threads)" % ncore)
        # the same call with the anchor targets left in HBM for a GPU-side loss: host head
        # outputs in, detections out.  Reported beside e2e, not instead of it.
        e2e_more["e2e_targets_resident"] = e2e_leg(lambda: sess.run_host(gts, h_prob, h_box, targets="device"),
                                                   "anchor targets stay on the device; detections (boxes, scores, keep) are read back")

    if args.breakdown and rank == 0:
        anc = sess.anchors
        ops_t = {"anchor_pyramid": median_ms(torch, lambda: detbox.anchor_pyramid(STRIDES, IH, IW, SCALES, RATIOS, BASE)),
                 "anchor_target": median_ms(torch, lambda: detbox.anchor_target(anc, d_gt, d_off, sess.tcfg)),
                 "topk_decode": median_ms(torch, lambda: detbox.topk_decode(d_box, d_prob, anc, TOP_N, "softmax", "fastrcnn", True, want=("img_ids", "scores")))}
        sel = detbox.topk_decode(d_box, d_prob, anc, TOP_N, "softmax", "fastrcnn", True, want=("img_ids", "scores"))
        seg = sel["img_ids"].int()
        ops_t["nms_segmented"] = median_ms(torch, lambda: detbox.nms(sel["boxes"], sel["scores"], NMS_THR, detbox.NMS_GT_F32, seg_ids=seg, n_seg=B))
        for kid, name in ((2, "k_colmax"), (3, "k_assign"), (4, "k_weights"), (5, "k_objectness"), (6, "k_nms_mask"), (7, "k_nms_scan")):
            ea, eb = event_pairs(torch, 1)[0]
            ea.record(); eb.record()
            L.detbox_trace_kernel(kid, ea.cuda_event, eb.cuda_event)
            step_serial(); torch.cuda.synchronize()
            L.detbox_trace_kernel(0, None, None)
            ops_t[name] = ea.elapsed_time(eb)
        print("breakdown (ms, batch %d): %s" % (B, json.dumps({k: round(v, 4) for k, v in ops_t.items()})), file=sys.stderr)

    extras = {}
    if not args.no_extras:                            # every rank: the leg ends in a collective
        strong = strong_64_leg(torch, dist, detbox, rank, world, args.steps, args.warmup)
        if rank == 0:
            extras["strong_64"] = strong
    if rank == 0 and world == 1 and not args.no_extras:
        extras["g_sweep"] = g_sweep_leg(torch, dist, detbox, L, args.steps, args.warmup, hbm_peak, d_prob, d_box, B)
        extras["nms_us"] = nms_sweep(torch, detbox)
        if cpu_nms:
            extras["nms_us"]["cpu_baseline"] = cpu_nms
        extras["roi_align"] = roi_align_bench(torch, detbox, hbm_peak)
        if cpu_roi:
            extras["roi_align"]["cpu_baseline"] = cpu_roi
        extras["single_image"] = single_image_latency(torch, detbox, gts[0], h_prob, h_box, A)
        extras["coco_r_post"] = coco_post_bench(torch, detbox, hbm_peak)

    if rank == 0:
        total_images = B * world
        roof = None
        tb = trace_bytes(args.trace, B, A, G_total, C)
        if k_ms and tb:
            ach = tb / k_ms / 1e6
            traffic = None
            try:                                     # per-launch DRAM bytes of the same kernel from the committed ncu capture
                traffic = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json"))).get(str(args.trace), {}).get("bytes")
                if B != 64:
                    traffic = None                   # the capture is of the default batch
            except Exception:
                pass
            roof = {"bound": "hbm", "kernel_id": args.trace, "achieved": round(ach, 1), "peak": hbm_peak, "unit": "GB/s",
                    "frac": round(ach / hbm_peak, 4), "traffic": traffic, "kernel_ms": round(k_ms, 4),
                    "algorithmic_bytes_per_launch": tb, "peak_source": peak_src,
                    "timed_in": "%d eager one-stream steps after the main timed region (CUDA events around every launch of the kernel; "
                                "inside the graph the two chains share the SMs and events cannot bracket a node)" % args.steps}
        line = {"metric": METRIC, "value": round(total_images * args.steps / (ms_dev / 1e3), 1), "unit": UNIT,
                "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(ms_dev / args.steps, 4),
                "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64 IoU / f32 boxes",
                "data": "synthetic",
                "config": bench_config(B, world),
                "launch": "one CUDA-graph replay per step; anchor-target chain and detection chain on two streams inside the graph",
                "e2e": e2e,
                "gpu_launches": int(launches), "clocks": clocks,
                "step_modes_ms": {"graph_two_streams": round(ms_dev / args.steps, 4), "eager_two_streams": round(ms_eager / args.steps, 4),
                                  "eager_one_stream": round(ms_serial / args.steps, 4)}}
        if roof:
            line["roofline"] = roof
        line.update(e2e_more)
        if cpu_base:
            line["cpu_baseline"] = cpu_base
        line.update(extras)
        print(json.dumps(line), flush=True)
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
