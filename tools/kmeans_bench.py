"""Device time of PointClusters.cluster at the C2 shape: 3M positions, 1024 clusters (reference trainer/trainer.py:327-370
clusters every evaluation with 100 iterations)."""
import os
import sys

import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "splat-trainer_b200"))
from gsb200 import ops  # noqa: E402

gen = torch.Generator(device="cuda").manual_seed(0)
x = torch.randn((3_000_000, 3), device="cuda", generator=gen) * 5
init = x[torch.randperm(x.shape[0], device="cuda", generator=gen)[:1024]].clone()
ops.kmeans(x, iters=2, centroids=init)
for iters in (10, 100):
  e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
  e0.record()
  ops.kmeans(x, iters=iters, centroids=init)
  e1.record()
  torch.cuda.synchronize()
  ms = e0.elapsed_time(e1)
  pairs = x.shape[0] * 1024 * iters
  print(f"kmeans 3M x 1024, {iters} iterations: {ms:.1f} ms ({ms / iters:.3f} ms/iter, {pairs / ms / 1e9:.2f} T point-centroid pairs/s)")
