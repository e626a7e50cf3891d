"""run_host chunk size vs step time for the round-2 formats (round 1 tuned chunk=8 for 402 MB in / 672 MB out)."""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, bench, detbox
B, A = 64, bench.A_TOTAL
sess = detbox.pipeline.BoxOpsSession(bench.STRIDES, bench.IH, bench.IW, bench.SCALES, bench.RATIOS, bench.BASE, B, 2, top_n=900, max_gt_total=12 * B)
gts = [bench.synth_gt(np.random.RandomState(100 + i), 12) for i in range(B)]
rng = np.random.RandomState(0)
deltas, prob = bench.synth_head(rng, B * A, rng.choice(B * A, 20000, replace=False))
hp, hb = torch.from_numpy(prob).pin_memory(), torch.from_numpy(deltas).pin_memory()
def t(fn, n=7):
    fn(); fn(); torch.cuda.synchronize(); ts = []
    for _ in range(n):
        t0 = time.perf_counter(); fn(); torch.cuda.synchronize(); ts.append(time.perf_counter() - t0)
    return round(1e3 * float(np.median(ts)), 2)
out = {}
for chunk in (2, 4, 8, 16, 32):
    out[chunk] = {"compact": t(lambda: sess.run_host(gts, hp, hb, chunk=chunk, format="compact")),
                  "reference_arrays": t(lambda: sess.run_host(gts, hp, hb, chunk=chunk, wire="compact")),
                  "targets_resident": t(lambda: sess.run_host(gts, hp, hb, chunk=chunk, targets="device"))}
print(json.dumps(out))
