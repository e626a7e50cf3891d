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
kw = dict(format=sys.argv[1]) if len(sys.argv) > 1 else {}
if len(sys.argv) > 2:
    kw["chunk"] = int(sys.argv[2])
for _ in range(3):
    sess.run_host(gts, hp, hb, **kw)
sess.timeline = []
t0 = time.perf_counter()
sess.run_host(gts, hp, hb, **kw)
wall = 1e3 * (time.perf_counter() - t0)
ev0 = sess.timeline[0][0]
print(kw, "wall ms %.3f" % wall, "cpu issue done at ms %.3f" % sess.issue_ms)
print(" ".join("%s@%.2f" % (name, ev0[1].elapsed_time(ev)) for name, ev in sess.timeline[1:]))
