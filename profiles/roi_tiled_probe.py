"""Two launches of the tile-centric dense NCHW backward on BASELINE configs[2] inputs (for ncu: -k regex:roi_bwd_nchw_tile)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, bench, detbox

B, C, P, strides, shapes, ref_areas, rois = bench.roi_bench_inputs()
d_rois = torch.from_numpy(rois).cuda()
gen = torch.Generator(device="cuda").manual_seed(3)
flush = torch.zeros(64 * 1024 * 1024, dtype=torch.float32, device="cuda")
scales = [1.0 / s for s in strides]
g = torch.randn((rois.shape[0], C, P, P), generator=gen, device="cuda")
for _ in range(2):
    flush.add_(1.0)
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    gi = detbox.roi_align_backward(g, shapes, d_rois, scales, detbox.ROI_BIN_CENTRE, ref_areas=ref_areas, deterministic=True)
    ev1.record(); torch.cuda.synchronize()
    print("backward %.3f ms" % ev0.elapsed_time(ev1))
    del gi
print("roi_tiled_probe done")
