"""Oracle parity ON the BASELINE.json configurations themselves:
  C2  3M splats SH3 1297x840 one camera (the headline workload, generated as SURVEY 8(d) writes it, and round 1's sparse
      variant): tile lists + ranges bit-exact against the oracle mapper run on the GPU's own projected rows (13-20 M
      overlaps), image max-abs <= 1e-4, n_contrib identical, parameter gradients of the full default loss.
  C1  100k splats SH0 640x480 one camera, as stated: forward + backward + one LaProp step against the oracle chain.
  PSNR gate: C1 trained for 300 steps on both sides, |dPSNR| <= 0.05 dB (north_star; metric trainer/evaluation.py:29-31).
The oracle needs ~2 s per C2 image on the GPU box's host cores."""
import os
import sys

import numpy as np
import pytest
import torch

from gsb200 import lib as L
from gsb200 import synthetic as syn
from gsb200.scene import SceneConfig, SplatScene
from gsb200.trainer import Trainer
from gpu_util import cu, npy
from oracle import gs_oracle as go
from oracle import gs_train as gt

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _lrs(cfg):
  return dict(position=cfg.lr_position, log_scaling=cfg.lr_log_scaling, rotation=cfg.lr_rotation,
              alpha_logit=cfg.lr_alpha_logit, sh=cfg.lr_sh)


def _cams(wl, k):
  c = wl["cameras"][k]
  W, H = wl["width"], wl["height"]
  return (L.make_camera(c["T"], c["fx"], c["fy"], c["cx"], c["cy"], W, H, c["near"], c["far"]),
          go.make_camera(c["T"], c["fx"], c["fy"], c["cx"], c["cy"], W, H, c["near"], c["far"]))


def _grad_check(name, a, r, q=0.999, tol=1e-4, cap=1e-2):
  """gradient tolerance of the suite: per column, |a - r| / max|r| <= tol for a fraction q of the rows (fp32 atomics
  reorder sums of thousands of terms) and <= cap everywhere."""
  a, r = a.reshape(r.shape[0], -1), r.reshape(r.shape[0], -1)
  scale = np.abs(r).max(axis=0) + 1e-20
  e = np.abs(a - r) / scale
  print(f"grad {name}: rows {r.shape[0]}, entries above {tol:g} of the column max: {int((e > tol).sum())} of {e.size} "
        f"({(e > tol).mean():.2e}), max {e.max():.2e}, p99.9 {np.quantile(e, 0.999):.2e}")
  assert np.quantile(e, q) <= tol, (name, np.quantile(e, [0.5, 0.99, 0.999, 1.0]).tolist())
  assert e.max() <= cap, (name, float(e.max()))


def _stagewise(wl, cam, ocam, min_overlaps):
  """Every stage of the hot path at the workload's full size against the oracle, each stage fed the GPU's own output of the
  stage before it (so one stage's rounding cannot hide in the next stage's tolerance), then the whole chain end to end."""
  from gsb200 import ops
  n, deg, W, H = wl["n"], wl["sh_degree"], wl["width"], wl["height"]
  cfg = SceneConfig(sh_degree=deg)
  scene = SplatScene(wl["gaussians"], cfg)
  for _ in range(4):
    w, tb, _, _ = scene._forward(cam)
    if not scene.overflowed():
      break
    scene.grow_overlaps(2.0)
  torch.cuda.synchronize()
  O = int(tb.counters[0].item())
  g2d, depth, rgb = npy(w["g2d"]), npy(w["depth"]), npy(w["rgb"])
  orc = go.Oracle("f32")
  ocfg = go.make_config(blur_cov=cfg.blur_cov)
  g = wl["gaussians"]
  report = dict(workload=wl["description"], O=O)
  # (1) projection + SH colour
  og2d, odepth, orgb = orc.project_fwd(ocam, ocfg, g["position"], g["log_scaling"], g["rotation"], g["alpha_logit"], g["sh"], deg)
  both = (g2d[:, 5] > 0) & (og2d[:, 5] > 0)
  report["V"] = int((g2d[:, 5] > 0).sum())
  assert ((g2d[:, 5] > 0) != (og2d[:, 5] > 0)).mean() < 2e-3  # FMA contraction may flip a row sitting on a cull boundary
  on = both & (np.abs(og2d[:, 0]) < 4 * W) & (np.abs(og2d[:, 1]) < 4 * H)  # far off-screen means carry fp32 ulps of 1e5 px
  np.testing.assert_allclose(g2d[on, :2], og2d[on, :2], rtol=1e-5, atol=5e-3)
  np.testing.assert_allclose(g2d[both, 5], og2d[both, 5], rtol=1e-4, atol=1e-6)
  np.testing.assert_allclose(rgb[both], orgb[both], rtol=1e-5, atol=1e-5)
  np.testing.assert_allclose(depth[both], odepth[both], rtol=1e-5)
  # (2) tile map: bit-exact against the oracle's literal 64-bit-key sort, on the GPU's own rows
  o2p, ranges, total, _ = orc.tile_map(ocfg, g2d, depth, W, H)
  assert total == O and O > min_overlaps
  assert np.array_equal(npy(tb.tile_ranges), ranges)
  assert np.array_equal(npy(tb.overlap_to_point[:O]), o2p)
  # (3) raster forward on the same lists.  Blend decisions are bit-identical by construction (Spec R); the alphas differ
  # by ex2.approx vs exp2 (2 ulp), so a pixel whose transmittance lands within rounding of the saturation threshold
  # (T < 1e-4) may stop one splat earlier or later: allowed for a small fraction of pixels, each of which is saturated
  image, fT, nc, _, _ = orc.raster_fwd(ocfg, g2d, rgb, depth, o2p, ranges, W, H, visibility=False)
  gi, gT, gnc = npy(w["image"]), npy(w["final_T"]), npy(w["n_contrib"])
  report["image_max_abs"] = float(np.abs(gi - image).max())
  assert report["image_max_abs"] <= 1e-4
  differ = gnc != nc
  report["n_contrib_differs"] = float(differ.mean())
  assert differ.mean() < 2e-3
  if differ.any():
    assert max(float(gT[differ].max()), float(fT[differ].max())) < 2.1e-4, "only saturating pixels may stop at a different splat"
  np.testing.assert_allclose(gT[~differ], fT[~differ], rtol=0, atol=1e-5)
  print("PARITY forward", report, flush=True)
  # (4) raster backward on the SAME forward state (the GPU's final_T / n_contrib) and the same dL/dimage
  tgt = SplatScene(wl["target_gaussians"], SceneConfig(sh_degree=deg))
  target = (tgt.render(cam)["image"] * 255.0 + 0.5).to(torch.uint8)
  del tgt
  torch.cuda.empty_cache()
  d_image = w["d_image"]
  sums = torch.zeros((2,), device="cuda")
  ops.loss_l1_mse(w["image"], target, cfg.w_l1, cfg.w_mse, d_pred=d_image, sums=sums)
  ops.ssim_loss(w["image"], target, cfg.ssim_levels, cfg.w_ssim, d_pred=d_image, workspace=w["ssim_ws"])
  acc = ops.raster_bwd(scene.raster_cfg, w["g2d"], w["rgb"], tb.overlap_to_point, tb.tile_ranges, W, H, w["final_T"],
                       w["n_contrib"], d_image, aux=w["aux"])
  oacc = orc.raster_bwd(ocfg, g2d, rgb, o2p, ranges, W, H, gT, gnc, npy(d_image))
  a = npy(acc)
  blended = oacc[:, 11] > 0
  report["rows_blended"] = int(blended.sum())
  assert report["rows_blended"] > 300 and np.array_equal(a[:, 11] > 0, blended)
  scale = np.abs(oacc[:, :12]).max(axis=0) + 1e-20
  e = np.abs(a[:, :12] - oacc[:, :12])[blended] / scale
  report["raster_bwd"] = dict(p999=float(np.quantile(e, 0.999)), max=float(e.max()), above_1e4=float((e > 1e-4).mean()))
  print("PARITY raster_bwd", report["raster_bwd"], flush=True)
  # 99.9 % of the entries within 1e-4 of the column maximum where a splat sums thousands of pixel terms; the 8(d) scene's
  # few hundred screen-filling splats sum ~1e6 signed terms each in a different order (fp32 atomics): 99 % there
  q = 0.999 if report["rows_blended"] > 5000 else 0.99
  assert np.quantile(e, q) <= 1e-4 and e.max() <= 1e-2, report["raster_bwd"]
  # (5) projection / SH backward on the same accumulator
  vis = np.ascontiguousarray(oacc[:, 11])
  og = orc.project_bwd(ocam, ocfg, g["position"], g["log_scaling"], g["rotation"], g["alpha_logit"], g["sh"], deg, g2d,
                       a[:, 0:6], a[:, 6:9], None, row_vis=vis)
  p = scene.p
  gg = ops.project_bwd(cam, scene.raster_cfg, p["position"], p["log_scaling"], p["rotation"], p["alpha_logit"], p["sh"], deg,
                       w["g2d"], acc[:, 0:6], acc[:, 6:9], None, row_vis=acc[:, 11])
  for name in gt.ATTRS:
    _grad_check("project_bwd " + name, npy(gg[name])[blended], og[name][blended])
  acc.zero_()
  # (6) the whole chain end to end (own projection, own lists, full default loss, fused reg gradients, statistics)
  scene.forward_backward(cam, target, first=True)
  torch.cuda.synchronize()
  osc = gt.OracleScene(g, deg, _lrs(cfg), w_l1=cfg.w_l1, w_mse=cfg.w_mse, w_ssim=cfg.w_ssim, ssim_levels=cfg.ssim_levels,
                       reg_weights=cfg.reg_weights, blur_cov=cfg.blur_cov)
  osc.forward_backward(ocam, npy(target))
  report["end_to_end_image_max_abs"] = float(np.abs(gi - osc.last["image"]).max())
  for name in gt.ATTRS:
    # the two sides project with different roundings (FMA contraction), so rows in front of the camera whose 2-D
    # covariance is ill-conditioned differ more than the kernels do stage by stage: 99 % of the entries within 1e-4 of the
    # column maximum, all within 5e-2
    _grad_check("chain " + name, npy(scene.g[name]), osc.grads[name], q=0.99, tol=1e-4, cap=5e-2)
  for k in ("visibility", "prune_cost", "split_score", "max_scale_px"):
    x, y = npy(scene.stats[k]), osc.stats[k]
    rows = y != 0
    rel = np.abs(x - y)[rows] / np.abs(y[rows])
    err = np.abs(x - y)[rows] / np.abs(y).max()
    report["stats_" + k] = dict(rows=int(rows.sum()), median_rel=float(np.median(rel)), p999_of_max=float(np.quantile(err, 0.999)),
                                max_of_max=float(err.max()))
    # split_score / prune_cost go through the reference's exp_lerp (util/misc.py:56-59) in fp32: exp_lerp(t, 0, x) =
    # x + log((1-t)·e^-x + t) cancels to ~t·x for small x, so whatever the implementation the result carries an absolute
    # error of an ulp of 1.0 (6e-8; measured: GPU and oracle differ by exactly 2^-24 on values of 1e-7).  Tolerance: two such
    # ulps plus 0.5 % for 99.9 % of the rows; the other statistics are plain sums / maxima (1e-3 of the column maximum)
    if k in ("prune_cost", "split_score"):
      ok = np.abs(x - y)[rows] <= 2.4e-7 + 5e-3 * np.abs(y[rows])
      assert ok.mean() >= 0.999 and np.median(rel) < 2e-3, (k, float(ok.mean()), report["stats_" + k])
    else:
      assert np.quantile(err, 0.999) < 1e-3 and err.max() < 5e-2, (k, report["stats_" + k])
    assert np.count_nonzero((x != 0) != rows) <= 1e-4 * rows.size, k
  piv = npy(scene.stats["points_in_view"]) != osc.stats["points_in_view"]
  assert piv.mean() < 2e-3
  print("PARITY REPORT", report)
  return report


@pytest.mark.parametrize("workload", ["c2", "c2-sparse", "c2-trained"])
def test_c2_every_stage_matches_oracle(workload):
  wl = syn.workload(workload)
  cam, ocam = _cams(wl, 0)
  _stagewise(wl, cam, ocam, min_overlaps=3_000_000)


def test_c1_as_stated_forward_backward_laprop_step_matches_oracle():
  """BASELINE.json configs[0]: 100k splats, SH degree 0, one 640x480 camera: forward + backward render + optimiser step."""
  wl = syn.workload("c1")
  n, deg, W, H = wl["n"], wl["sh_degree"], wl["width"], wl["height"]
  assert (n, deg, W, H) == (100_000, 0, 640, 480)
  cam, ocam = _cams(wl, 0)
  _stagewise(wl, cam, ocam, min_overlaps=100_000)
  cfg = SceneConfig(sh_degree=0)
  scene = SplatScene(wl["gaussians"], cfg)
  osc = gt.OracleScene(wl["gaussians"], 0, _lrs(cfg), w_l1=cfg.w_l1, w_mse=cfg.w_mse, w_ssim=cfg.w_ssim,
                       ssim_levels=cfg.ssim_levels, reg_weights=cfg.reg_weights, blur_cov=cfg.blur_cov)
  tgt = gt.OracleScene(wl["target_gaussians"], 0, _lrs(cfg))
  target = np.clip(tgt.render(ocam) * 255.0 + 0.5, 0, 255).astype(np.uint8)
  tr = Trainer(scene)
  sums = tr.training_step([(cam, cu(target))], read_loss=True)
  osums = osc.training_step([(ocam, target)])
  np.testing.assert_allclose(sums, osums, rtol=1e-3)
  before = wl["gaussians"]
  for name in gt.ATTRS:
    a, r = npy(scene.p[name]).reshape(n, -1), osc.g[name].reshape(n, -1)
    b0 = before[name].reshape(n, -1)[:, :r.shape[1]]
    moved = np.abs(r - b0).max()
    assert moved > 0, name
    # the first LaProp step of a row is lr·lam·sign-like(m̂): compare the steps taken, relative to the largest step
    e = np.abs((a - b0) - (r - b0)) / moved
    print(f"step {name}: largest step {moved:.3e}, p99 {np.quantile(e, 0.99):.2e}, max {e.max():.2e}")
    assert np.quantile(e, 0.99) <= 2e-3, (name, float(np.quantile(e, 0.99)))


def test_psnr_after_300_steps_within_0p05_db_of_the_oracle():
  """north_star: PSNR after a fixed iteration count within 0.05 dB.  C1-shaped (100k splats, SH0, 640x480), 300 steps of
  the full default loss over 8 cameras, GPU and oracle from the same seed.
  (a) learning rates = 0.1 x the reference's: the dynamics are contracting (the oracle against a 1e-7-perturbed copy of
      itself stays within 0.001 dB), so any difference is the implementation's: strict 0.05 dB at every checkpoint.
  (b) the reference's learning rates as they are: the dynamics amplify a 1e-7 perturbation (measured in the same run as
      `floor_db`), so the gate is 0.05 dB or three times that floor, whichever is larger."""
  sys.path.insert(0, os.path.join(ROOT, "tools"))
  import psnr_parity
  a = psnr_parity.run(steps=300, lr_scale=0.1)
  print("PSNR GATE a", a["curve"], a["delta_db"])
  assert a["final_gpu_psnr"] > a["curve"][0]["gpu"] + 3.0, "training did not improve the PSNR"
  assert a["max_delta_db"] <= 0.05, a
  b = psnr_parity.run(steps=300, lr_scale=1.0, floor=True)
  print("PSNR GATE b", b["curve"], b["delta_db"], b["floor_db"])
  assert b["delta_db"] <= max(0.05, 3.0 * b["max_floor_db"]), b
