"""Where does the one-block segment sort beat the radix route?  top-k + decode time for K winners per segment with
DETBOX_SEG_SORT_LIMIT (read once per process) given on the command line: python profiles/topk_sort_probe.py"""
import json, os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CHILD = r'''
import sys, json; sys.path.insert(0, %r)
import torch, bench, detbox
gen = torch.Generator(device="cuda").manual_seed(5)
out = {}
for name, B, A, C, K, per in (("per_image_3000", 8, 132450, 2, 3000, True), ("per_image_6000", 8, 132450, 2, 6000, True),
                              ("global_8000", 16, 132450, 2, 8000, False), ("global_14400", 16, 132450, 2, 14400, False)):
    p1 = torch.rand((B * A, 1), generator=gen, device="cuda"); prob = torch.cat([1 - p1, p1], 1).contiguous()
    deltas = torch.randn((B * A, 4), generator=gen, device="cuda") * 0.5
    anc = torch.rand((A, 4), generator=gen, device="cuda") * 500; anc[:, 2:] += anc[:, :2]
    f = lambda: detbox.topk_decode(deltas, prob, anc, K, "softmax", "fastrcnn", per, want=("img_ids", "scores"))
    out[name] = round(bench.median_ms(torch, f, iters=15, warm=4) * 1e3, 1)
print(json.dumps(out))
''' % ROOT
for lim in (2048, 4096, 8192, 16384):
    r = subprocess.run([sys.executable, "-c", CHILD], env=dict(os.environ, DETBOX_SEG_SORT_LIMIT=str(lim)), capture_output=True, text=True)
    print("limit %5d us:" % lim, r.stdout.strip().splitlines()[-1] if r.stdout.strip() else r.stderr[-300:])
