"""Scratch probe: PCIe floor for the e2e step (402 MB in, 672 MB out) and run_host chunk sweep."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, bench, detbox

def t(fn, n=5):
    fn(); torch.cuda.synchronize(); ts = []
    for _ in range(n):
        t0 = time.perf_counter(); fn(); torch.cuda.synchronize(); ts.append(time.perf_counter() - t0)
    return 1e3 * float(np.median(ts))

hin = torch.empty(402 * 1000 * 1000 // 4, dtype=torch.float32, pin_memory=True)
hout = torch.empty(672 * 1000 * 1000 // 4, dtype=torch.float32, pin_memory=True)
din, dout = torch.empty_like(hin, device="cuda"), torch.empty_like(hout, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
print("H2D 402MB ms", t(lambda: din.copy_(hin, non_blocking=True)))
print("D2H 672MB ms", t(lambda: hout.copy_(dout, non_blocking=True)))
def both():
    with torch.cuda.stream(s1): din.copy_(hin, non_blocking=True)
    with torch.cuda.stream(s2): hout.copy_(dout, non_blocking=True)
print("both concurrently ms", t(both))

B, A = 64, bench.A_TOTAL
sess = detbox.pipeline.BoxOpsSession(bench.STRIDES, bench.IH, bench.IW, bench.SCALES, bench.RATIOS, bench.BASE, B, 2,
                                     top_n=900, max_gt_total=12 * B)
gts = [bench.synth_gt(np.random.RandomState(i), 12) for i in range(B)]
rng = np.random.RandomState(0)
deltas, prob = bench.synth_head(rng, B * A, rng.choice(B * A, 20000, replace=False))
hp, hb = torch.from_numpy(prob).pin_memory(), torch.from_numpy(deltas).pin_memory()
for chunk in (2, 4, 8, 16, 32, 64):
    print("run_host chunk", chunk, "ms", t(lambda: sess.run_host(gts, hp, hb, chunk=chunk)))
