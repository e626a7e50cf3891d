"""Scratch probe: run_host step time against the chunk size of the three-stream pipeline."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, bench, detbox
B, A = 64, bench.A_TOTAL
sess = detbox.pipeline.BoxOpsSession(bench.STRIDES, bench.IH, bench.IW, bench.SCALES, bench.RATIOS, bench.BASE, B, 2,
                                     top_n=900, max_gt_total=12 * B)
gts = [bench.synth_gt(np.random.RandomState(i), 12) for i in range(B)]
rng = np.random.RandomState(0)
deltas, prob = bench.synth_head(rng, B * A, rng.choice(B * A, 20000, replace=False))
hp, hb = torch.from_numpy(prob).pin_memory(), torch.from_numpy(deltas).pin_memory()
for targets in ("host", "device"):
    for chunk in (2, 4, 8, 16, 32, 64):
        for _ in range(2):
            sess.run_host(gts, hp, hb, chunk=chunk, targets=targets)
        torch.cuda.synchronize(); t0 = time.perf_counter()
        for _ in range(5):
            sess.run_host(gts, hp, hb, chunk=chunk, targets=targets)
        torch.cuda.synchronize()
        print(targets, chunk, round(1e3 * (time.perf_counter() - t0) / 5, 3), "ms/step")
