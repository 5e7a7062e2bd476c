"""Device time of the colour MLP kernels at the C2 shape (3M splats, ~1.67M in view): CUDA events, L2-exceeding inputs."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "splat-trainer_b200"))
from gsb200 import ops  # noqa: E402


def main():
  n, p_view = 3_000_000, 0.558
  gen = torch.Generator(device="cuda").manual_seed(0)
  G = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden", "color_mlp.npz"))
  model = ops.color_model({k: torch.from_numpy(G[k]).cuda() for k in ops.COLOR_WEIGHTS})
  feature = torch.randn((n, 16), device="cuda", generator=gen)
  position = torch.randn((n, 3), device="cuda", generator=gen) * 4
  glo = torch.randn((32,), device="cuda", generator=gen)
  iv = (torch.rand((n,), device="cuda", generator=gen) < p_view).float()
  rgb = torch.zeros((n, 3), device="cuda")
  flush = torch.empty((256 << 20,), device="cuda", dtype=torch.uint8)
  times = []
  for it in range(8):
    flush.zero_()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    ops.color_mlp_fwd(model, feature, position, [0.0, 0.0, -5.0], glo, in_view=iv, rgb=rgb)
    e1.record()
    torch.cuda.synchronize()
    times.append(e0.elapsed_time(e1))
  v = int(iv.sum().item())
  t = float(np.median(times[2:]))
  flop = v * 2 * (48 * 64 + 32 * 4 + 36 * 96 + 48 * 64 + 32 * 4)
  print(f"color fwd: {v} rows in view, {t * 1e3:.1f} us, {flop / t / 1e9:.1f} TFLOP/s (useful), "
        f"{(n * 4 + v * (64 + 12 + 12)) / t / 1e6:.0f} GB/s algorithmic")


if __name__ == "__main__":
  main()
