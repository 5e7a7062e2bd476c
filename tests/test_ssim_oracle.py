"""CPU: the oracle's multi-scale SSIM loss (oracle/gs_oracle.c section 8) against an independent torch float64
restatement with autograd.  fused-ssim itself is absent from /root/reference and this image (parity unpinned at that
boundary); what is pinned is the reference's own composition, trainer/trainer.py:450-462: valid padding, 2x2 average
pooling between levels, loss = mean over levels of (1 - ssim)."""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

from oracle import gs_oracle as go


def torch_ssim(pred, ref):
  """fused_ssim(pred, ref, padding='valid') as published: 11x11 Gaussian window sigma 1.5, zero-padded conv, map cropped by 5."""
  g = torch.exp(-(torch.arange(11, dtype=torch.float64) - 5) ** 2 / (2 * 1.5 ** 2))
  g = g / g.sum()
  C = pred.shape[1]
  win = (g[:, None] * g[None, :]).expand(C, 1, 11, 11)
  blur = lambda x: F.conv2d(x, win, padding=5, groups=C)
  mu1, mu2 = blur(pred), blur(ref)
  s1, s2, s12 = blur(pred * pred) - mu1 * mu1, blur(ref * ref) - mu2 * mu2, blur(pred * ref) - mu1 * mu2
  C1, C2 = 0.01 ** 2, 0.03 ** 2
  m = ((2 * mu1 * mu2 + C1) * (2 * s12 + C2)) / ((mu1 * mu1 + mu2 * mu2 + C1) * (s1 + s2 + C2))
  return m[:, :, 5:-5, 5:-5].mean()


def torch_ssim_loss(pred, ref, levels):
  """reference trainer/trainer.py:450-462 (compute_ssim_loss) on (H,W,C) images."""
  ref = ref.unsqueeze(0).permute(0, 3, 1, 2)
  pred = pred.unsqueeze(0).permute(0, 3, 1, 2)
  ssim = torch_ssim(pred, ref)
  loss = 1.0 - ssim
  for _ in range(1, levels):
    pred = F.avg_pool2d(pred, kernel_size=2, stride=2)
    ref = F.avg_pool2d(ref, kernel_size=2, stride=2)
    loss = loss + (1.0 - torch_ssim(pred, ref))
  return loss / levels, ssim.item()


@pytest.mark.parametrize("size,levels,C", [((61, 83), 3, 3), ((48, 48), 1, 1), ((97, 64), 2, 4)])
def test_ssim_loss_and_gradient_match_torch(size, levels, C):
  H, W = size
  rng = np.random.default_rng(H + levels)
  base = rng.uniform(0, 1, (H, W, C))
  pred = np.clip(base + rng.normal(0, 0.1, base.shape), 0, 1)
  target = np.clip(base + rng.normal(0, 0.05, base.shape), 0, 1)
  tp = torch.from_numpy(pred).requires_grad_()
  loss, s0 = torch_ssim_loss(tp, torch.from_numpy(target), levels)
  (0.7 * loss).backward()
  o64 = go.Oracle("f64")
  l64, d64, s64 = o64.ssim_loss(pred, target, levels=levels, weight=0.7)
  assert abs(l64 - 0.7 * loss.item()) < 1e-10 and abs(s64 - s0) < 1e-10
  np.testing.assert_allclose(d64, tp.grad.numpy(), rtol=1e-7, atol=1e-12)
  o32 = go.Oracle("f32")
  l32, d32, s32 = o32.ssim_loss(pred.astype(np.float32), target.astype(np.float32), levels=levels, weight=0.7)
  assert abs(l32 - 0.7 * loss.item()) < 2e-5
  scale = np.abs(d64).max()
  assert np.abs(d32 - d64).max() / scale < 2e-3


def test_ssim_identical_images_is_one():
  rng = np.random.default_rng(0)
  img = rng.uniform(0, 1, (40, 50, 3)).astype(np.float32)
  loss, d, s0 = go.Oracle("f32").ssim_loss(img, img, levels=2, weight=1.0)
  assert abs(s0 - 1.0) < 1e-5 and abs(loss) < 1e-5 and np.abs(d).max() < 1e-3
