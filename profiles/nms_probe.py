"""Scratch probe: one NMS call per size (after warm-up) for ncu launch lists."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, bench, detbox
n = int(sys.argv[1]) if len(sys.argv) > 1 else 20000
rng = np.random.RandomState(2)
h = np.exp(rng.uniform(np.log(12), np.log(264), n)); w = h / 2.44
cx, cy = rng.uniform(0, 2048, n), rng.uniform(0, 1024, n)
b = np.stack([np.clip(cx - w / 2, 0, 2047), np.clip(cy - h / 2, 0, 1023), np.clip(cx + w / 2, 0, 2047), np.clip(cy + h / 2, 0, 1023)], 1).astype(np.float32)
s = rng.permutation(np.linspace(0, 1, n)).astype(np.float32)
db, ds = torch.from_numpy(b).cuda(), torch.from_numpy(s).cuda()
for _ in range(3):
    keep, nk = detbox.nms(db, ds, 0.5, detbox.NMS_GT_F32)
torch.cuda.synchronize()
print(n, int(nk[-1]))
