"""Scratch probe: BASELINE configs[3] post-processing extra of bench.py on its own."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, bench, detbox
print(json.dumps(bench.coco_post_bench(torch, detbox, 6429.6)))
