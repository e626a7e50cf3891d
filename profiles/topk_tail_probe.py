"""A/B of the fused top-k tail (candidate select + segment sort + gather/decode in one launch) against the three separate
kernels (DETBOX_TOPK_UNFUSED_TAIL=1): microseconds per call inside a CUDA graph holding 16 calls.  Prints [launches, 3 timings]."""
import os, sys, torch, json
sys.path.insert(0, '/root/repo')
import bench, detbox
from detbox import ops_post
dev = torch.device('cuda:0')
torch.manual_seed(0)
REP = 16
def t(fn, it=100):
    for _ in range(20): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(REP): fn()          # many calls per replay: one replay of a 5-kernel graph is bound by the Python call
    for _ in range(5): g.replay()
    e0.record()
    for _ in range(it): g.replay()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / it / REP * 1e3
res = {}
for name, B, A, C, K, per, act in [("step_64x16368", 64, 16368, 2, 900, True, "softmax"),
                                   ("single_16368", 1, 16368, 2, 900, True, "softmax"),
                                   ("coco_132450x81", 1, 132450, 81, 1000, False, "sigmoid")]:
    prob = torch.rand(B, A, C, device=dev); box = torch.randn(B, A, 4, device=dev) * 0.1
    anc = torch.rand(A, 4, device=dev) * 500; anc[:, 2:] += anc[:, :2] + 8
    fn = lambda: ops_post.topk_decode(box.view(-1, 4), prob.view(-1, C), anc, top_n=K, per_image=per, activation=act)
    n0 = detbox.launches(); fn(); nl = detbox.launches() - n0
    res[name] = [nl] + [round(t(fn), 2) for _ in range(3)]
print(json.dumps({"unfused" if os.environ.get("DETBOX_TOPK_UNFUSED_TAIL") else "fused": res}))
