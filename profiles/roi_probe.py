"""Scratch probe: RoIAlign config-3 forward/backward timings (used for ncu captures of the NHWC kernels)."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, bench, detbox
print(json.dumps(bench.roi_align_bench(torch, detbox, 6549.1)))
