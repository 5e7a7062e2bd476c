"""GPU parity of the multi-scale SSIM loss (include/gsb200.h section 8) against the CPU oracle (which tests/
test_ssim_oracle.py checks against an independent torch float64 restatement with autograd)."""
import numpy as np
import pytest
import torch

from gsb200 import ops
from gpu_util import cu, npy, oracle

pytestmark = pytest.mark.gpu


def _images(H, W, C, seed, overshoot=0.0):
  rng = np.random.default_rng(seed)
  base = rng.uniform(0, 1, (H, W, C))
  pred = base + rng.normal(0, 0.1, base.shape)
  if overshoot == 0.0:
    pred = np.clip(pred, 0, 1)
  target = np.clip(base + rng.normal(0, 0.05, base.shape), 0, 1)
  return pred.astype(np.float32), target.astype(np.float32)


@pytest.mark.parametrize("size,levels,C", [((61, 83), 3, 3), ((48, 48), 1, 1), ((97, 64), 2, 4), ((240, 320), 3, 3),
                                           ((33, 21), 2, 3)])
def test_ssim_loss_and_gradient_match_oracle(size, levels, C):
  H, W = size
  pred, target = _images(H, W, C, H + levels)
  base = torch.full((H, W, C), 0.125, device="cuda")
  d = base.clone()
  sums = ops.ssim_loss(cu(pred), cu(target), levels, 0.7, d_pred=d)
  loss_ref, d_ref, s0_ref = oracle().ssim_loss(pred, target, levels=levels, weight=0.7)
  counts = ops.ssim_counts(H, W, C, levels)
  ssim = [float(s) / c if c else 0.0 for s, c in zip(npy(sums), counts)]
  loss = 0.7 * np.mean([1.0 - v for v in ssim])
  assert abs(ssim[0] - s0_ref) < 2e-5 and abs(loss - loss_ref) < 2e-5
  got = npy(d - base)  # the gradient is ADDED
  scale = np.abs(d_ref).max()
  assert np.abs(got - d_ref).max() / scale < 2e-3, np.abs(got - d_ref).max() / scale
  assert np.quantile(np.abs(got - d_ref) / scale, 0.999) < 5e-4


def test_ssim_uint8_target_and_clamp_mask():
  """level 0 clamps the raw render on load and passes no gradient where it left [0,1]; uint8 targets are /255."""
  H, W, C = 70, 90, 3
  pred, target = _images(H, W, C, 3, overshoot=1.0)
  assert (pred < 0).any() and (pred > 1).any()
  tu8 = np.round(target * 255).astype(np.uint8)
  d = torch.zeros((H, W, C), device="cuda")
  sums = ops.ssim_loss(cu(pred), cu(tu8), 3, 1.0, d_pred=d)
  loss_ref, d_ref, _ = oracle().ssim_loss(np.clip(pred, 0, 1), tu8.astype(np.float32) / 255.0, levels=3, weight=1.0)
  d_ref = np.where((pred >= 0) & (pred <= 1), d_ref, 0)
  counts = ops.ssim_counts(H, W, C, 3)
  loss = np.mean([1.0 - float(s) / c for s, c in zip(npy(sums), counts)])
  assert abs(loss - loss_ref) < 2e-5
  got = npy(d)
  assert np.all(got[(pred < 0) | (pred > 1)] == 0)
  assert np.abs(got - d_ref).max() / np.abs(d_ref).max() < 2e-3


def test_ssim_identical_images():
  img = _images(64, 64, 3, 9)[1]
  d = torch.zeros((64, 64, 3), device="cuda")
  sums = ops.ssim_loss(cu(img), cu(img), 2, 1.0, d_pred=d)
  counts = ops.ssim_counts(64, 64, 3, 2)
  assert all(abs(float(s) / c - 1.0) < 1e-5 for s, c in zip(npy(sums), counts))
  assert d.abs().max().item() < 1e-4


@pytest.mark.parametrize("size,levels,C,u8", [((70, 90), 3, 3, True), ((97, 64), 2, 4, False), ((48, 48), 1, 1, False),
                                              ((240, 323), 3, 3, True)])
def test_photometric_loss_is_l1_mse_then_ssim(size, levels, C, u8):
  """gs_photometric_loss (the fused call of the training path, reference trainer/trainer.py:465-498) against the two
  separate entry points and against the oracle's gso_loss_l1_mse + gso_ssim_loss on an over-shooting render."""
  H, W = size
  pred, target = _images(H, W, C, H + 7, overshoot=1.0)
  tu8 = np.round(target * 255).astype(np.uint8)
  tgt = cu(tu8) if u8 else cu(target)
  tref = tu8.astype(np.float32) / np.float32(255.0) if u8 else target
  ws = ops.ssim_workspace(H, W, C, levels)
  d = torch.full((H, W, C), 7.0, device="cuda")  # WRITTEN, whatever was there
  ls = torch.full((2,), 3.0, device="cuda"); ss = torch.full((levels,), 3.0, device="cuda")
  ops.photometric_loss(cu(pred), tgt, 0.3, 10.0, 0.7, levels, d, ls, ss, ws)
  d2 = torch.zeros((H, W, C), device="cuda")
  _, ls2 = ops.loss_l1_mse(cu(pred), tgt, 0.3, 10.0, d_pred=d2)
  ss2 = ops.ssim_loss(cu(pred), tgt, levels, 0.7, d_pred=d2)
  assert np.allclose(npy(ls), npy(ls2), rtol=2e-6) and np.allclose(npy(ss), npy(ss2), rtol=2e-6)
  assert np.abs(npy(d) - npy(d2)).max() <= 1e-7 * max(1.0, float(d2.abs().max()))
  o = oracle()
  d_lm, sums_ref = o.loss_l1_mse(pred, None if u8 else target, 0.3, 10.0, target_u8=tu8 if u8 else None)
  _, d_ss, _ = o.ssim_loss(np.clip(pred, 0, 1), tref, levels=levels, weight=0.7)
  d_ref = d_lm + np.where((pred >= 0) & (pred <= 1), d_ss, 0)
  assert np.allclose(npy(ls), sums_ref, rtol=1e-4)
  assert np.abs(npy(d) - d_ref).max() / np.abs(d_ref).max() < 2e-3
