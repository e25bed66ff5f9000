"""Median step latency of the graph-replayed step over a range of batch sizes (one GPU): python tools/latency_sweep.py [B ...]
Environment switches (DDP_BWD_DENSE, DDP_SPLIT_HEAD, ...) are read once per process, so A/B = two runs."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
import pytorch_distributed_examples_b200 as pkg
from pytorch_distributed_examples_b200 import _native
nat = _native.get()
sizes = [int(a) for a in sys.argv[1:]] or [64, 256, 512, 1024, 2048, 4096]
print(json.dumps({b: round(bench.step_latency_stats(pkg, nat, b, 0.9, steps=400)["us_median"], 1) for b in sizes}))
