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


def bwd():
  n = 3_000_000
  gen = torch.Generator(device="cuda").manual_seed(0)
  G = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden", "color_mlp.npz"))
  model = ops.color_model({k: torch.from_numpy(G[k]).cuda() for k in ops.COLOR_WEIGHTS})
  grads_t = {k: torch.zeros(G[k].shape, device="cuda") for k in ops.COLOR_WEIGHTS}
  grads = ops.color_model(grads_t)
  feature = torch.randn((n, 16), device="cuda", generator=gen)
  position = torch.randn((n, 3), device="cuda", generator=gen) * 4
  glo = torch.randn((32,), device="cuda", generator=gen)
  d_feature, d_glo = torch.zeros((n, 16), device="cuda"), torch.zeros((32,), device="cuda")
  acc = torch.randn((n, 12), device="cuda", generator=gen) * 1e-4
  flush = torch.empty((256 << 20,), device="cuda", dtype=torch.uint8)
  for p_vis in (0.0166, 0.25):
    w = (torch.rand((n,), device="cuda", generator=gen) < p_vis).float()
    times = []
    for it in range(6):
      flush.zero_()
      e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
      e0.record()
      ops.color_mlp_bwd(model, grads, feature, position, [0.0, 0.0, -5.0], glo, w, acc[:, 6:9], d_feature, d_glo)
      e1.record()
      torch.cuda.synchronize()
      times.append(e0.elapsed_time(e1))
    print(f"color bwd: {int(w.sum().item())} blended rows, {float(np.median(times[2:])) * 1e3:.1f} us")


if __name__ == "__main__":
  if "fwd" not in sys.argv[1:]:
    bwd()
  main()
