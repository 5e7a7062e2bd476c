"""GPU parity of the per-step controller / scene passes (include/gsb200.h section 7) against the golden vectors
produced by the reference's own code (tests/golden/mcmc.npz, reg_loss.npz) and against the CPU oracle."""
import numpy as np
import pytest
import torch

from gsb200 import lib as L
from gsb200 import ops
from gsb200 import synthetic as syn
from gsb200.scene import SceneConfig, SplatScene
from gpu_util import cu, npy, oracle
from oracle import gs_oracle as go
from test_golden_oracle import gold

pytestmark = pytest.mark.gpu


def test_mcmc_noise_matches_reference_and_oracle():
  g = gold("mcmc.npz")
  pos = cu(g["position"])
  ops.mcmc_noise(cu(g["alpha_logit"][:, 0]), cu(g["log_scaling"]), cu(g["rotation"]), cu(g["points_in_view"]),
                 int(g["min_views"]), float(g["opacity_threshold"]), float(g["noise_level"]), cu(g["randn"]), pos)
  out = npy(pos)
  moved = g["points_in_view"] > int(g["min_views"])
  assert np.array_equal(out[~moved], g["position"][~moved])
  np.testing.assert_allclose(out, g["out_position"], rtol=2e-5, atol=2e-6)
  ref = g["position"].copy()
  oracle().mcmc_noise(g["alpha_logit"][:, 0], g["log_scaling"], g["rotation"], g["points_in_view"], int(g["min_views"]),
                      float(g["opacity_threshold"]), float(g["noise_level"]), g["randn"], ref)
  np.testing.assert_allclose(out, ref, rtol=2e-5, atol=2e-6)


def test_mcmc_prune_respawn_masks_match_reference():
  g = gold("mcmc.npz")
  n = g["position"].shape[0]
  sh = np.zeros((n, 3, 1), np.float32)
  scene = SplatScene(dict(position=g["position"], log_scaling=g["log_scaling"], rotation=g["rotation"],
                          alpha_logit=g["alpha_logit"][:, 0], sh=sh), SceneConfig(sh_degree=0))
  scene.stats["max_scale_px"].copy_(cu(g["max_scale_px"]))
  scene.stats["split_score"].copy_(cu(g["split_score"]))
  split_mask, prune_mask = scene.mcmc_find_prune_split(float(g["opacity_threshold"]), 200.0, float(g["min_split_px"]))
  assert np.array_equal(npy(prune_mask), g["prune_mask"])
  assert np.array_equal(npy(split_mask), g["split_mask"])
  assert int(split_mask.sum()) == int(prune_mask.sum())


def test_reg_loss_matches_reference_and_oracle():
  g = gold("reg_loss.npz")
  m = g["opacity"].shape[0]
  w = (float(g["w_scale"]), float(g["w_opacity"]), float(g["w_aspect"]))
  g2d = np.zeros((m, 6), np.float32); g2d[:, 5] = g["opacity"]
  # strided views like the fused path: vis is a plane, d_opacity is column 5 of the raster accumulator
  acc = torch.zeros((m, 12), device="cuda")
  acc[:, 5] = 0.25
  d_depth = torch.full((m,), 7.0, device="cuda")
  d_ls = torch.full((m, 3), 0.5, device="cuda")
  args = (w, cu(g2d), cu(g["depths"][:, 0]), cu(g["log_scaling"]), cu(g["weight"]))
  sums = ops.reg_loss(*args, compute_sums=True, d_opacity=acc[:, 5], d_depth=d_depth)
  ops.reg_loss(*args, sums=sums, compute_sums=False, d_log_scaling=d_ls)
  s = npy(sums)
  assert s[3] == m
  np.testing.assert_allclose(s[0] / m, g["reg_scale"], rtol=2e-5)
  np.testing.assert_allclose(s[1] / m, g["reg_opacity"], rtol=2e-5)
  np.testing.assert_allclose(s[2] / m, g["reg_aspect"], rtol=2e-5)
  np.testing.assert_allclose(npy(acc[:, 5]) - 0.25, g["d_opacity"], rtol=2e-4, atol=6e-8)   # ADDED (onto 0.25: fp32 ulp)
  np.testing.assert_allclose(npy(d_depth), g["d_depths"][:, 0], rtol=2e-4, atol=1e-9)        # WRITTEN
  np.testing.assert_allclose(npy(d_ls) - 0.5, g["d_log_scaling"], rtol=2e-4, atol=1.2e-7)    # ADDED (onto 0.5: fp32 ulp)
  assert not acc[:, :5].any() and not acc[:, 6:].any()
  # oracle, with a third of the rows not blended
  vis = g["weight"].copy(); vis[::3] = 0
  od = [np.zeros(m, np.float32), np.zeros(m, np.float32), np.zeros((m, 3), np.float32)]
  osums = oracle().reg_loss(w, g2d, g["depths"][:, 0], g["log_scaling"], vis, *od)
  gd = [torch.zeros((m,), device="cuda"), torch.full((m,), -3.0, device="cuda"), torch.zeros((m, 3), device="cuda")]
  gs = ops.reg_loss(w, cu(g2d), cu(g["depths"][:, 0]), cu(g["log_scaling"]), cu(vis), d_opacity=gd[0], d_depth=gd[1],
                    d_log_scaling=gd[2])
  np.testing.assert_allclose(npy(gs), osums, rtol=2e-5)
  np.testing.assert_allclose(npy(gd[0]), od[0], rtol=2e-4, atol=1e-9)
  np.testing.assert_allclose(npy(gd[1])[vis > 0], od[1][vis > 0], rtol=2e-4, atol=1e-9)
  assert np.all(npy(gd[1])[vis == 0] == -3.0), "d_depth is only written for blended rows"
  np.testing.assert_allclose(npy(gd[2]), od[2], rtol=2e-4, atol=1e-9)


def _scene(n=20000, W=320, H=240, **cfg):
  g = syn.garden_scene(n, 2, seed=7, focal=0.5 * W / np.tan(np.radians(25.0)))
  c = syn.ring_cameras(8, W, H)[1]
  cam = L.make_camera(c["T"], c["fx"], c["fy"], c["cx"], c["cy"], W, H)
  ocam = go.make_camera(c["T"], c["fx"], c["fy"], c["cx"], c["cy"], W, H)
  return g, cam, ocam, SplatScene(g, SceneConfig(sh_degree=2, **cfg))


def test_training_image_with_reg_matches_oracle_chain():
  """forward_backward with the reference's default loss (MSE 10 + 3-level SSIM 1.0 + reg_loss, reference
  trainer/trainer.py:465-498, config/trainer/default.yaml:52-55) == the oracle chain with gso_ssim_loss after the MSE and
  gso_reg_loss between the raster and the projection backward."""
  W, H = 320, 240
  wts = (0.01, 1.0, 1e-4)
  g, cam, ocam, scene = _scene(W=W, H=H, reg_weights=wts)
  rng = np.random.default_rng(0)
  target = rng.uniform(0, 1, (H, W, 3)).astype(np.float32)
  scene.forward_backward(cam, cu(target), first=True)
  orc = oracle(); cfg = go.make_config()
  K = 9
  sh = np.ascontiguousarray(g["sh"][:, :, :K])
  g2d, depth, rgb = orc.project_fwd(ocam, cfg, g["position"], g["log_scaling"], g["rotation"], g["alpha_logit"], sh, 2)
  o2p, ranges, total, _ = orc.tile_map(cfg, g2d, depth, W, H)
  image, fT, nc, _, _ = orc.raster_fwd(cfg, g2d, rgb, depth, o2p, ranges, W, H)
  dimg, _ = orc.loss_l1_mse(image, target, 0.0, 10.0)
  # + multi-scale SSIM with the reference's default weight / levels (through the clamp)
  ssim_ref, dssim, _ = orc.ssim_loss(np.clip(image, 0, 1), target, levels=3, weight=1.0)
  dimg = dimg + np.where((image >= 0) & (image <= 1), dssim, 0).astype(np.float32)
  acc = orc.raster_bwd(cfg, g2d, rgb, o2p, ranges, W, H, fT, nc, dimg)
  vis = np.ascontiguousarray(acc[:, 11])
  d_op = np.ascontiguousarray(acc[:, 5]); d_depth = np.zeros_like(vis); d_ls = np.zeros((vis.shape[0], 3), np.float32)
  sums = orc.reg_loss(wts, g2d, depth, g["log_scaling"], vis, d_op, d_depth, d_ls)
  acc[:, 5] = d_op
  grads = orc.project_bwd(ocam, cfg, g["position"], g["log_scaling"], g["rotation"], g["alpha_logit"], sh, 2, g2d,
                          acc[:, 0:6], acc[:, 6:9], d_depth, row_vis=vis)
  grads["log_scaling"] = grads["log_scaling"] + d_ls
  torch.cuda.synchronize()
  np.testing.assert_allclose(npy(scene._work["reg_sums"]), sums, rtol=1e-3)
  assert sums[3] > 500
  reg_ref = (wts[0] * sums[0] + wts[1] * sums[1] + wts[2] * sums[2]) / sums[3]
  np.testing.assert_allclose(float(scene.reg_value().item()), reg_ref, rtol=1e-3)
  np.testing.assert_allclose(float(scene.ssim_value()[0].item()), ssim_ref, rtol=1e-4)
  for name in ("position", "log_scaling", "rotation", "alpha_logit", "sh"):
    a, r = npy(scene.g[name]).reshape(vis.shape[0], -1), grads[name].reshape(vis.shape[0], -1)
    scale = np.abs(r).max(axis=0) + 1e-12
    e = np.abs(a - r) / scale
    assert np.quantile(e, 0.999) < 1e-4, (name, np.quantile(e, [0.5, 0.99, 0.999, 1.0]))
  # and the regularisation really contributes: switching it off changes the log_scaling gradient
  _, _, _, plain = _scene(W=W, H=H, reg_weights=(0.0, 0.0, 0.0))
  plain.forward_backward(cam, cu(target), first=True)
  assert (plain.g["log_scaling"] - scene.g["log_scaling"]).abs().max().item() > 1e-6


def test_query_visibility_equals_forward_visibility():
  g, cam, ocam, scene = _scene()
  idx, vis = scene.query_visibility(cam)
  full = scene.render(cam, compute_visibility=True)["visibility"]
  assert idx.dtype == torch.int64 and torch.all(idx[1:] > idx[:-1])
  # visibility is summed with fp32 atomics across warps: equal up to summation order
  assert int((full > 0).sum()) == idx.shape[0] > 500 and torch.equal((full > 0).nonzero().squeeze(1), idx)
  torch.testing.assert_close(full[idx], vis, rtol=1e-4, atol=1e-6)
