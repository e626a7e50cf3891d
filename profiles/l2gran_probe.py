"""Probe: does cudaLimitMaxL2FetchGranularity (32 / 64 / 128 B) change the gather-bound RoIAlign kernels?"""
import ctypes, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, bench, detbox
torch.cuda.init(); torch.zeros(1).cuda()
rt = None
for name in ("libcudart.so.12", "libcudart.so"):
    try:
        rt = ctypes.CDLL(name); break
    except OSError:
        pass
if rt is None:
    import glob
    c = glob.glob(os.path.join(os.path.dirname(torch.__file__), "..", "nvidia", "cuda_runtime", "lib", "libcudart.so*"))
    rt = ctypes.CDLL(c[0])
LIMIT = 0x05   # cudaLimitMaxL2FetchGranularity
for g in (None, 32, 64, 128):
    if g is not None:
        rc = rt.cudaDeviceSetLimit(LIMIT, ctypes.c_size_t(g))
    v = ctypes.c_size_t(0); rt.cudaDeviceGetLimit(ctypes.byref(v), LIMIT)
    r = bench.roi_align_bench(torch, detbox, 6549.1)
    print(g, v.value, json.dumps({k: {kk: r[k][kk] for kk in ("fwd_ms", "bwd_ms", "bwd_scatter_only_ms")} for k in ("nchw", "nhwc")}))
