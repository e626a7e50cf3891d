"""Scratch probe: per-chunk timeline of BoxOpsSession.run_host (copy-in / kernels / copy-out streams)."""
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
for _ in range(3):
    sess.run_host(gts, hp, hb)
sess.timeline = []
t0 = time.perf_counter()
sess.run_host(gts, hp, hb)
wall = 1e3 * (time.perf_counter() - t0)
ev0 = sess.timeline[0][0]
print("wall ms", wall, "cpu issue done at ms", sess.issue_ms)
for name, ev in sess.timeline[1:]:
    print("%-10s %.3f" % (name, ev0[1].elapsed_time(ev)))
