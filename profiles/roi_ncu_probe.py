"""RoIAlign launches for ncu (BASELINE configs[2] inputs of bench.roi_bench_inputs): exactly two launches of every
flavour in a fixed order, so a capture of all `roi_` kernels can be read by KERNEL NAME (template arguments included)
instead of by skip count:  fwd NCHW x2, bwd NCHW x2, fwd channels-last x2, bwd channels-last x2.  The second launch of
each pair is the warm one.  A 256 MB write between launches evicts L2 (as bench.py does between timed iterations)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, bench, detbox

B, C, P, strides, shapes, ref_areas, rois = bench.roi_bench_inputs()
d_rois = torch.from_numpy(rois).cuda()
gen = torch.Generator(device="cuda").manual_seed(3)
flush = torch.zeros(64 * 1024 * 1024, dtype=torch.float32, device="cuda")
scales = [1.0 / s for s in strides]
for layout in ("nchw", "nhwc"):
    feats = [torch.randn(s, generator=gen, device="cuda") for s in shapes]
    if layout == "nhwc":
        feats = [f.contiguous(memory_format=torch.channels_last) for f in feats]
    y = None
    for _ in range(2):
        flush.add_(1.0)
        y = detbox.roi_align_forward(feats, d_rois, scales, P, P, detbox.ROI_BIN_CENTRE, ref_areas=ref_areas)
    g = torch.randn(y.shape, generator=gen, device="cuda")
    if layout == "nhwc":
        g = g.contiguous(memory_format=torch.channels_last)
    grads = [torch.zeros_like(f) for f in feats]
    for _ in range(2):
        flush.add_(1.0)
        detbox.roi_align_backward(g, shapes, d_rois, scales, detbox.ROI_BIN_CENTRE, ref_areas=ref_areas,
                                  channels_last=(layout == "nhwc"), grads=grads)
    torch.cuda.synchronize()
    del feats, grads, g, y
print("roi_ncu_probe done")
