"""Batched box path with HOST buffers in and out — the call a user of the reference makes.

The reference runs this path per image on NumPy arrays:
    anchors  = anchor_pyramid(...)                                  libs/boxes/anchor.py:30
    targets  = compute_rpn_targets_in_batch(gt_boxes_list, anchors) libs/layers/data_layer.py:228
    dets     = get_final_results_stage1(outputs, anchors, ...)      libs/nets/model.py:474
             = objectness -> top_n -> decode -> per-image NMS       model.py:308-355
`BoxOpsSession.run_host` is that composite for a whole batch: pinned host buffers are copied to
the device, every stage runs as hand-written CUDA kernels of libdetbox.so on one stream with no
intermediate host synchronisation, and the results are copied back into pinned host buffers.
`run_device` is the same sequence on tensors already resident in HBM.  There is no CPU path.
"""
from __future__ import annotations

import os

import numpy as np
import torch

from . import ops, ops_post
from ._lib import DetboxError


class BoxOpsSession:
    """Pre-allocated staging for a fixed (batch, canvas, anchor config).

    top_n is taken PER IMAGE (the sharded mode of SURVEY §8e, identical to the reference's
    behaviour at its own eval batch size of 1, test/test.py:71).
    """

    def __init__(self, strides, ih, iw, anchor_scales, anchor_ratios, anchor_base, batch, num_classes,
                 top_n=900, nms_thresh=0.5, nms_cmp=ops.NMS_GT_F32, activation="softmax", bg=0.5, fg=0.6,
                 max_gt_total=None, anchor_shift=((0.0, 0.0),), device="cuda", gt0_tables=False):
        if not torch.cuda.is_available():
            raise DetboxError("BoxOpsSession needs a CUDA device: detbox has no CPU path")
        self.dev = torch.device(device)
        self.anchor_args = (list(strides), int(ih), int(iw), anchor_scales, anchor_ratios, anchor_base, anchor_shift)
        self.B, self.C = int(batch), int(num_classes)
        self.top_n, self.nms_thresh, self.nms_cmp, self.activation = int(top_n), float(nms_thresh), nms_cmp, activation
        self.tcfg = ops.make_target_cfg(bg, fg)
        self.anchors = ops.anchor_pyramid(*self.anchor_args, device=self.dev)
        # grid geometry of those rows: lets the anchor-target pass take GT-0 targets from separable tables (the
        # library verifies separability on the device on every call, so this is a hint, not a promise)
        # gt0_tables=True hands that geometry to the anchor-target pass.  Bit-identical results; measured on B200 at
        # CP-R, batch 64 it is SLOWER (anchor_target_kernel 0.165 vs 0.153 ms: the two table loads and the index
        # arithmetic cost more issue slots than the packed encode they replace), hence off by default.
        self.pyr = ops.pyramid_desc(*self.anchor_args, device=self.dev) if gt0_tables else None
        self.A = A = self.anchors.shape[0]
        B = self.B
        self.K = B * min(self.top_n, A)
        G = int(max_gt_total or 128 * B)
        pin = lambda shape, dt: torch.empty(shape, dtype=dt, pin_memory=True)
        dv = lambda shape, dt: torch.empty(shape, dtype=dt, device=self.dev)
        # inputs: host staging + device copies
        self.h_gt, self.d_gt = pin((G, 5), torch.float32), dv((G, 5), torch.float32)
        self.h_off, self.d_off = pin((B + 1,), torch.int32), dv((B + 1,), torch.int32)
        self.d_prob, self.d_box = dv((B * A, self.C), torch.float32), dv((B * A, 4), torch.float32)
        # outputs on the host (the device tensors are allocated by the ops)
        self.h_labels, self.h_lw = pin((B, A), torch.float32), pin((B, A), torch.float32)
        self.h_bt, self.h_bw = pin((B, A, 4), torch.float32), pin((B, A, 4), torch.float32)
        self.h_boxes, self.h_scores = pin((self.K, 4), torch.float32), pin((self.K,), torch.float32)
        self.h_img = pin((self.K,), torch.int64)
        self.h_keep, self.h_nkeep = pin((self.K,), torch.int64), pin((B + 1,), torch.int32)
        self.h_packed = self.h_counts = self.d_packed = None     # compact host format: allocated on first use
        self._keep_out = pin((self.K,), torch.int64)
        self._h_nk = None
        self._streams = [torch.cuda.Stream(self.dev) for _ in range(3)]     # copy-in, kernels, copy-out
        # run_device's second stream (detection chain).  DETBOX_SIDE_PRIORITY: 0 = same priority, -1 = above the
        # anchor-target chain's stream.
        self._side_stream = torch.cuda.Stream(self.dev, priority=int(os.environ.get("DETBOX_SIDE_PRIORITY", "-1")))
        self.overlap = os.environ.get("DETBOX_OVERLAP", "1") != "0"
        self.h2d_bytes = 0
        self.d2h_bytes = 0

    # ------------------------------------------------------------------ device-resident path --
    def run_device(self, gt, gt_offsets, rpn_prob, rpn_box, regenerate_anchors=True, overlap=None):
        """gt [sumG,5], gt_offsets int32 [B+1], rpn_prob [B*A,C], rpn_box [B*A,4] — CUDA tensors.

        The two halves of the path share nothing but the anchors: the anchor-target chain (tile boxes ->
        column max -> assign -> label weights; long, bandwidth-heavy kernels) and the detection chain
        (objectness -> top-k -> decode -> NMS; a dozen short kernels, several with one block per image).
        `overlap` (default: self.overlap) forks the detection chain onto a second stream after the anchors
        exist and joins it before returning, so its latency-bound kernels run in the gaps of the other chain.
        Results are identical either way; fork / join are stream waits, so the whole call stays capturable
        in one CUDA graph."""
        overlap = self.overlap if overlap is None else overlap
        if regenerate_anchors:      # the reference rebuilds the pyramid for every batch (data_layer.py:138)
            self.anchors = ops.anchor_pyramid(*self.anchor_args, device=self.dev)

        def detections():
            sel = ops_post.topk_decode(rpn_box, rpn_prob, self.anchors, self.top_n, self.activation, "fastrcnn",
                                       per_image=True, want=("img_ids", "scores"))
            keep, n_keep = ops_post.nms(sel["boxes"], sel["scores"], self.nms_thresh, self.nms_cmp,
                                        seg_ids=sel["img_ids"], n_seg=self.B, presorted=True)
            return sel, keep, n_keep

        if overlap:
            cur = torch.cuda.current_stream(self.dev)
            side = self._side_stream
            side.wait_stream(cur)                                  # fork: after the anchors and every earlier use of the inputs
            with torch.cuda.stream(side):
                sel, keep, n_keep = detections()
            labels, lw, bt, bw = ops.anchor_target(self.anchors, gt, gt_offsets, self.tcfg, pyramid=self.pyr)
            cur.wait_stream(side)                                  # join
        else:
            labels, lw, bt, bw = ops.anchor_target(self.anchors, gt, gt_offsets, self.tcfg, pyramid=self.pyr)
            sel, keep, n_keep = detections()
        return {"labels": labels, "label_weights": lw, "bbox_targets": bt, "bbox_inside_weights": bw,
                "boxes": sel["boxes"], "scores": sel["scores"], "img_ids": sel["img_ids"],
                "keep": keep, "n_keep": n_keep}

    def run_device_graph(self, gt, gt_offsets, rpn_prob, rpn_box):
        """run_device captured ONCE in a CUDA graph (both streams, fork and join included) and replayed: one
        launch call per step instead of ~17, so the host never paces the device.  The graph is bound to the
        argument tensors (addresses and shapes): refill them in place between calls; a different tensor
        re-captures.  Returns the SAME output tensors on every call — each replay overwrites them."""
        key = tuple((t.data_ptr(), tuple(t.shape)) for t in (gt, gt_offsets, rpn_prob, rpn_box))
        if getattr(self, "_graph_key", None) != key:
            cur = torch.cuda.current_stream(self.dev)
            warm = torch.cuda.Stream(self.dev)
            warm.wait_stream(cur)
            with torch.cuda.stream(warm):                          # workspaces and allocator pools settle before capture
                for _ in range(2):
                    self.run_device(gt, gt_offsets, rpn_prob, rpn_box)
            cur.wait_stream(warm)
            torch.cuda.synchronize(self.dev)
            graph = torch.cuda.CUDAGraph()
            n0 = ops.launches()
            with torch.cuda.graph(graph):
                out = self.run_device(gt, gt_offsets, rpn_prob, rpn_box)
            self.graph_launches = ops.launches() - n0              # kernels per replay
            self._graph, self._graph_out, self._graph_key = graph, out, key
        self._graph.replay()
        return self._graph_out

    # ----------------------------------------------------------------------- host-buffer path --
    def stage_gt(self, gt_boxes_list):
        """Pack a list of [G_i,5] arrays into the pinned GT buffer; returns total G."""
        if len(gt_boxes_list) != self.B:
            raise DetboxError("expected %d GT arrays, got %d" % (self.B, len(gt_boxes_list)))
        off = 0
        self.h_off[0] = 0
        hg = self.h_gt.numpy()
        for i, g in enumerate(gt_boxes_list):
            g = np.asarray(g, np.float32).reshape(-1, 5)
            if off + g.shape[0] > hg.shape[0]:
                raise DetboxError("more GT boxes than max_gt_total=%d" % hg.shape[0])
            hg[off:off + g.shape[0]] = g
            off += g.shape[0]
            self.h_off[i + 1] = off
        return off

    def run_host(self, gt_boxes_list, rpn_prob_host, rpn_box_host, chunk=None, targets="host", format="reference",
                 box_rows="auto", wire="same", expand_threads=None):
        """gt_boxes_list: B NumPy arrays [G_i,5]; rpn_prob_host [B*A,C] / rpn_box_host [B*A,4]:
        float32 host tensors (pinned for asynchronous copies).  Returns pinned host tensors; the
        call returns after all streams have drained, so they are ready to read.

        The batch is cut into chunks of `chunk` images that flow through three streams — copy-in,
        kernels, copy-out — so the PCIe link runs in both directions while the kernels of the
        neighbouring chunk execute (images are independent: no result changes).  chunk=None picks the size measured
        best for the mode on B200 (profiles/chunk_probe_r02.py): 4 when the expansion is overlapped (more of it hides), 16
        for the compact format and for targets="device", 8 for the four float32 arrays over the link.

        targets="device": the four anchor-target tensors stay in HBM ([B,A] / [B,A,4] CUDA tensors in the result)
        for a loss that runs on the GPU — the training integration of INTEGRATION.md §2, where the reference
        itself copies its CPU-computed targets TO the device — and only the detections travel back.

        format="reference": labels / label_weights / bbox_targets / bbox_inside_weights as float32 arrays, the
        reference's return types (40 bytes per anchor over the link).  format="compact": 'packed' uint8 [B,A]
        ((label + 1) | flag << 7), 'bbox_targets' [B,A,4] and 'counts' int32 [B,2] — 17 bytes per anchor, the
        same information; `expand(result)` rebuilds the reference's arrays on the host, bit for bit.

        wire="compact" (with format="reference", targets="host"): the caller still receives the reference's four float32
        arrays, but every chunk crosses the link in the compact format and is expanded by detbox_expand_targets_host on
        worker threads as soon as its device->host copy has landed — the expansion (400 MB of host writes per 64 CP-R
        images) hides behind the transfers of the following chunks instead of running after them.  Bit-identical to
        wire="same".

        box_rows: "copy" stages all of rpn_box on the device (16 bytes x A per image); "mapped" lets the top-k
        gather read the top_n selected rows straight out of the pinned host buffer (model.py:314-318 indexes
        rpn_box[order] the same way); "auto" = mapped when rpn_box_host is pinned."""
        if targets not in ("host", "device"):
            raise DetboxError("targets must be 'host' or 'device'")
        if format not in ("reference", "compact"):
            raise DetboxError("format must be 'reference' or 'compact'")
        if box_rows == "auto":
            box_rows = "mapped" if rpn_box_host.is_pinned() else "copy"
        if box_rows not in ("copy", "mapped"):
            raise DetboxError("box_rows must be 'auto', 'copy' or 'mapped'")
        if wire not in ("same", "compact"):
            raise DetboxError("wire must be 'same' or 'compact'")
        on_dev = targets == "device"
        expand_here = wire == "compact" and format == "reference" and not on_dev
        compact = format == "compact" or expand_here
        if chunk is None:
            chunk = 4 if expand_here else (16 if (compact or on_dev) else 8)
        chunk = max(1, min(int(chunk), self.B))
        B, A, C = self.B, self.A, self.C
        dv = lambda shape, dt=torch.float32: torch.empty(shape, dtype=dt, device=self.dev)
        if getattr(self, "d_bt", None) is None:                   # device-side result buffers, reused by every call
            self.d_bt = dv((B, A, 4))
        if compact and self.d_packed is None:
            self.d_packed, self.d_counts = dv((B, A), torch.uint8), dv((B, 2), torch.int32)
            self.h_packed = torch.empty((B, A), dtype=torch.uint8, pin_memory=True)
            self.h_counts = torch.empty((B, 2), dtype=torch.int32, pin_memory=True)
        if not compact and getattr(self, "d_labels", None) is None:
            self.d_labels, self.d_lw, self.d_bw = dv((B, A)), dv((B, A)), dv((B, A, 4))
        futures = []
        if expand_here:
            if getattr(self, "_pool", None) is None:
                from concurrent.futures import ThreadPoolExecutor
                self._pool = ThreadPoolExecutor(max_workers=2)       # two chunks' expansions in flight; each call threads itself
            bbw = [float(self.tcfg.bbweights[i]) for i in range(4)]
            nthr = int(expand_threads or max(1, min(16, len(os.sched_getaffinity(0))) // 2))

            def expand_chunk(ev, sl):
                ev.synchronize()                                 # this chunk's device->host copies have landed (releases the GIL)
                ops.expand_targets(self.h_packed[sl], self.h_counts[sl], bbw, nthr, out=(self.h_labels[sl], self.h_lw[sl], self.h_bw[sl]))
        G = self.stage_gt(gt_boxes_list)
        s_in, s_k, s_out = self._streams
        cur = torch.cuda.current_stream(self.dev)
        for s in self._streams:
            s.wait_stream(cur)
        prob_h, box_h = rpn_prob_host.view(B * A, C), rpn_box_host.view(B * A, 4)
        kper = min(self.top_n, A)
        with torch.cuda.stream(s_in):
            self.d_gt[:G].copy_(self.h_gt[:G], non_blocking=True)
            self.d_off.copy_(self.h_off, non_blocking=True)
        self.anchors = ops.anchor_pyramid(*self.anchor_args, device=self.dev)     # data_layer.py:138, once per batch
        s_k.wait_stream(cur)
        hold, d2h, h2d = [], 0, G * 20 + self.h_off.numel() * 4
        tl = getattr(self, "timeline", None)        # optional instrumentation (profiles/timeline_probe.py)
        if tl is not None:
            import time as _time
            _t0 = _time.perf_counter()
            e = torch.cuda.Event(enable_timing=True); e.record(cur); tl.append((("start", e), None))
        def mark(name, stream):
            if tl is not None:
                e = torch.cuda.Event(enable_timing=True); e.record(stream); tl.append((name, e))
        # chunk schedule: a short first chunk starts the (bottleneck) copy-out stream early
        starts, c0 = [], 0
        while c0 < B:
            n = min(max(1, chunk // 4) if c0 == 0 and chunk < B else chunk, B - c0)
            starts.append((c0, n)); c0 += n
        if self._h_nk is None or self._h_nk.shape != (len(starts), chunk + 1):
            self._h_nk = torch.zeros((len(starts), chunk + 1), dtype=torch.int32, pin_memory=True)
        h_nk = self._h_nk
        for ci, (c0, n) in enumerate(starts):
            r0, r1 = c0 * A, (c0 + n) * A
            ev_in = torch.cuda.Event()
            with torch.cuda.stream(s_in):
                self.d_prob[r0:r1].copy_(prob_h[r0:r1], non_blocking=True)
                h2d += (r1 - r0) * C * 4
                if box_rows == "copy":
                    self.d_box[r0:r1].copy_(box_h[r0:r1], non_blocking=True)
                    h2d += (r1 - r0) * 16
                else:
                    h2d += n * kper * 16                           # rows the gather reads over the link
                ev_in.record(s_in)
                mark("h2d%d" % c0, s_in)
            ev_k = torch.cuda.Event()
            with torch.cuda.stream(s_k):
                s_k.wait_event(ev_in)
                sl = slice(c0, c0 + n)
                if compact:
                    tg = ops.anchor_target(self.anchors, self.d_gt[:G], self.d_off[c0:c0 + n + 1], self.tcfg, pyramid=self.pyr,
                                           format="compact", out=(self.d_packed[sl], self.d_bt[sl], self.d_counts[sl]))
                else:
                    tg = ops.anchor_target(self.anchors, self.d_gt[:G], self.d_off[c0:c0 + n + 1], self.tcfg, pyramid=self.pyr,
                                           out=(self.d_labels[sl], self.d_lw[sl], self.d_bt[sl], self.d_bw[sl]))
                sel = ops_post.topk_decode(self.d_box[r0:r1] if box_rows == "copy" else box_h[r0:r1], self.d_prob[r0:r1],
                                           self.anchors, self.top_n, self.activation, "fastrcnn", per_image=True,
                                           want=("img_ids", "scores"))
                keep, n_keep = ops_post.nms(sel["boxes"], sel["scores"], self.nms_thresh, self.nms_cmp,
                                            seg_ids=sel["img_ids"], n_seg=n, presorted=True)
                ev_k.record(s_k)
                mark("krn%d" % c0, s_k)
            k0, k1 = c0 * kper, (c0 + n) * kper
            pairs = [(self.h_boxes[k0:k1], sel["boxes"]), (self.h_scores[k0:k1], sel["scores"]),
                     (self.h_img[k0:k1], sel["img_ids"]), (self.h_keep[k0:k1], keep), (h_nk[ci, :n + 1], n_keep)]
            if not on_dev:
                if compact:
                    pairs = [(self.h_packed[sl], tg[0]), (self.h_bt[sl], tg[1]), (self.h_counts[sl], tg[2])] + pairs
                else:
                    pairs = [(self.h_labels[sl], tg[0]), (self.h_lw[sl], tg[1]), (self.h_bt[sl], tg[2]), (self.h_bw[sl], tg[3])] + pairs
            with torch.cuda.stream(s_out):
                s_out.wait_event(ev_k)
                for h, d in pairs:
                    h.copy_(d, non_blocking=True)
                    d2h += h.numel() * h.element_size()
                mark("d2h%d" % c0, s_out)
                if expand_here:
                    ev_out = torch.cuda.Event()
                    ev_out.record(s_out)
                    futures.append(self._pool.submit(expand_chunk, ev_out, sl))
            hold.append((sel, keep, n_keep))                          # keep device buffers alive until the drain
        if tl is not None:
            self.issue_ms = 1e3 * (_time.perf_counter() - _t0)
        for s in self._streams:
            s.synchronize()
        for f in futures:
            f.result()                                           # the last chunks' expansions; re-raises a worker's error
        # chunk-local indices -> batch-global (host side, a few thousand integers)
        nk_all = np.zeros(B + 1, np.int32)
        keep_np, img_np, nk_np = self.h_keep.numpy(), self.h_img.numpy(), h_nk.numpy()
        out_keep, pos = self._keep_out.numpy(), 0
        for ci, (c0, n) in enumerate(starts):
            k0 = c0 * kper
            tot = int(nk_np[ci, n])
            out_keep[pos:pos + tot] = keep_np[k0:k0 + tot] + k0
            pos += tot
            nk_all[c0:c0 + n] = nk_np[ci, :n]
            img_np[k0:k0 + n * kper] += c0
        nk_all[B] = pos
        self.h_nkeep.numpy()[:] = nk_all
        self.h2d_bytes, self.d2h_bytes = h2d, d2h
        res = {"boxes": self.h_boxes, "scores": self.h_scores, "img_ids": self.h_img, "keep": self._keep_out,
               "n_keep": self.h_nkeep, "format": format}
        if expand_here:
            res.update(labels=self.h_labels, label_weights=self.h_lw, bbox_targets=self.h_bt, bbox_inside_weights=self.h_bw,
                       counts=self.h_counts)
        elif compact:
            tg = (self.d_packed, self.d_bt, self.d_counts) if on_dev else (self.h_packed, self.h_bt, self.h_counts)
            res.update(packed=tg[0], bbox_targets=tg[1], counts=tg[2])
        else:
            tg = (self.d_labels, self.d_lw, self.d_bt, self.d_bw) if on_dev else (self.h_labels, self.h_lw, self.h_bt, self.h_bw)
            res.update(labels=tg[0], label_weights=tg[1], bbox_targets=tg[2], bbox_inside_weights=tg[3])
        return res

    def expand(self, host_out, n_threads=None):
        """The reference's four arrays from a format="compact" result of run_host (host side): labels [B,A],
        label_weights [B,A], bbox_targets [B,A,4] (passed through), bbox_inside_weights [B,A,4]."""
        if host_out.get("format") != "compact":
            return host_out["labels"], host_out["label_weights"], host_out["bbox_targets"], host_out["bbox_inside_weights"]
        bbw = [float(self.tcfg.bbweights[i]) for i in range(4)]
        labels, lw, bw = ops.expand_targets(host_out["packed"], host_out["counts"], bbw, n_threads,
                                            out=(self.h_labels, self.h_lw, self.h_bw))
        return labels, lw, host_out["bbox_targets"], bw

    def detections(self, host_out):
        """Per-image [m,5] = (x1,y1,x2,y2,score) arrays from run_host()'s result (host-side views)."""
        nk = host_out["n_keep"].numpy()
        keep = host_out["keep"].numpy()[:int(nk[self.B])]
        boxes, scores = host_out["boxes"].numpy(), host_out["scores"].numpy()
        dets, o = [], 0
        for b in range(self.B):
            k = keep[o:o + int(nk[b])]
            dets.append(np.concatenate([boxes[k], scores[k, None]], 1))
            o += int(nk[b])
        return dets
