"""The `taichi_splatting`-named operator shim (gsb200/compat), driven the way the reference's MLPScene drives the real
package (scene/mlp_scene.py:372-381 query_visibility, :411-427 render, :214-239 step; controller/point_state.py:34-50)
and checked against the CPU oracle on the same seeded inputs.  Tolerance: 1e-4 of the image range / per-tensor max."""
import os
import sys

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "splat-trainer_b200", "gsb200", "compat"))

import taichi_splatting as ts  # noqa: E402
from taichi_splatting.optim import ParameterClass, VisibilityAwareLaProp, VisibilityOptimizer  # noqa: E402
from taichi_splatting.perspective import CameraParams  # noqa: E402
from taichi_splatting.renderer import project_to_image, render_projected  # noqa: E402

from oracle import gs_oracle as go  # noqa: E402
from gpu_util import cu, npy, oracle  # noqa: E402
from util import look_at, random_gaussians  # noqa: E402

W, H = 160, 112


def scene(n=5000, seed=31):
  g = random_gaussians(n, seed=seed, sh_degree=0, scale=0.08)
  rng = np.random.default_rng(seed + 1)
  g["feature"] = rng.uniform(0, 1, (n, 3)).astype(np.float32)
  T = look_at((3.0, 0.4, 0.8))
  fx = 0.5 * W / np.tan(np.radians(25.0))
  ocam = go.make_camera(T, fx, fx, W / 2, H / 2, W, H, 0.1, 100.0)
  cam = CameraParams(T_camera_world=torch.tensor(T, dtype=torch.float32, device="cuda"),
                     projection=torch.tensor([fx, fx, W / 2, H / 2], dtype=torch.float32, device="cuda"),
                     image_size=(W, H), near_plane=0.1, far_plane=100.0)
  return g, ocam, cam


def rel_err(a, b):
  return np.abs(a - b).max() / max(np.abs(b).max(), 1e-12)


def test_render_forward_backward_like_mlp_scene():
  g, ocam, cam = scene()
  orc = oracle()
  ocfg = go.make_config()
  names = ("position", "log_scaling", "rotation", "alpha_logit", "feature")
  t = {k: cu(g[k]).requires_grad_(True) for k in names}
  gaussians = ts.Gaussians3D(position=t["position"], log_scaling=t["log_scaling"], rotation=t["rotation"],
                             alpha_logit=t["alpha_logit"].unsqueeze(1), feature=t["feature"])
  config = ts.RasterConfig(compute_visibility=True, compute_point_heuristic=True, antialias=False, blur_cov=0.3)
  g2d, depth, idx = project_to_image(gaussians, cam, config)
  assert depth.shape == (idx.shape[0], 1) and g2d.shape == (idx.shape[0], 6) and idx.dtype == torch.int64
  colors = gaussians.feature[idx]
  r = render_projected(idx, g2d, colors, depth, cam, config, render_median_depth=True)
  assert isinstance(r, ts.Rendering) and r.image.shape == (H, W, 3)
  rng = np.random.default_rng(5)
  wimg = rng.normal(0, 1, (H, W, 3)).astype(np.float32)
  wop = rng.normal(0, 0.1, g["position"].shape[0]).astype(np.float32)
  wdep = rng.normal(0, 0.1, g["position"].shape[0]).astype(np.float32)
  # image loss + a reg_loss-style term through points.opacity and points.depths (scene/mlp_scene.py:269-279)
  loss = (r.image * cu(wimg)).sum() + (r.points.opacity * cu(wop)[idx]).sum() + (r.points.depths[:, 0] * cu(wdep)[idx]).sum()
  loss.backward()

  # ---- oracle on the N-row layout
  og2d, odepth, _ = orc.project_fwd(ocam, ocfg, g["position"], g["log_scaling"], g["rotation"], g["alpha_logit"])
  o2p, ranges, total, _ = orc.tile_map(ocfg, og2d, odepth, W, H)
  oimg, fT, nc, ovis, omed = orc.raster_fwd(ocfg, og2d, g["feature"], odepth, o2p, ranges, W, H, visibility=True, median=True)
  inview = np.nonzero(og2d[:, 5] > 0)[0]
  assert np.array_equal(npy(idx), inview)
  assert np.abs(npy(r.image) - oimg).max() < 1e-4
  assert np.abs(npy(r.image_weight) - (1 - fT)).max() < 1e-4
  assert rel_err(npy(r.points.visibility), ovis[inview]) < 1e-4
  assert np.abs(npy(r.median_depth_image) - omed).max() < 1e-3
  acc = orc.raster_bwd(ocfg, og2d, g["feature"], o2p, ranges, W, H, fT, nc, wimg)
  dg = acc[:, 0:6].copy()
  dg[:, 5] += wop * (og2d[:, 5] > 0)
  grads = orc.project_bwd(ocam, ocfg, g["position"], g["log_scaling"], g["rotation"], g["alpha_logit"], None, 0, og2d,
                          dg, None, wdep * (og2d[:, 5] > 0))
  for k in ("position", "log_scaling", "rotation", "alpha_logit"):
    assert rel_err(npy(t[k].grad).reshape(grads[k].shape), grads[k]) < 2e-4, k
  assert rel_err(npy(t["feature"].grad), acc[:, 6:9]) < 1e-4
  # heuristics are filled by the backward into the tensors handed out by the forward (trainer/trainer.py:511-514)
  assert rel_err(npy(r.points.prune_cost), acc[inview, 9]) < 1e-4
  assert rel_err(npy(r.points.split_score), acc[inview, 10]) < 1e-4
  vis = r.points.visible
  assert vis.idx.shape[0] == r.points.num_visible == int((ovis[inview] > 0).sum())
  assert r.points.screen_scale.shape == (idx.shape[0], 2) and bool((r.points.screen_scale[:, 0] >= r.points.screen_scale[:, 1]).all())
  d = r.detach()
  assert not d.image.requires_grad


def test_query_visibility_single_channel():
  """MLPScene.query_visibility: forward-only render of a 1-channel zero feature (scene/mlp_scene.py:372-381)."""
  g, ocam, cam = scene(3000, seed=7)
  t = {k: cu(g[k]) for k in ("position", "log_scaling", "rotation", "alpha_logit")}
  gaussians = ts.Gaussians3D(t["position"], t["log_scaling"], t["rotation"], t["alpha_logit"].unsqueeze(1), cu(g["feature"]))
  config = ts.RasterConfig(compute_visibility=True)
  with torch.no_grad():
    g2d, depth, idx = project_to_image(gaussians, cam, config)
    r = render_projected(idx, g2d, torch.zeros((idx.shape[0], 1), device="cuda"), depth, cam, config)
  orc, ocfg = oracle(), go.make_config()
  og2d, odepth, _ = orc.project_fwd(ocam, ocfg, g["position"], g["log_scaling"], g["rotation"], g["alpha_logit"])
  o2p, ranges, _, _ = orc.tile_map(ocfg, og2d, odepth, W, H)
  _, _, _, ovis, _ = orc.raster_fwd(ocfg, og2d, np.zeros((3000, 1), np.float32), odepth, o2p, ranges, W, H, visibility=True)
  visible = r.points.visible
  assert np.array_equal(npy(visible.idx), np.nonzero(ovis > 0)[0])
  assert rel_err(npy(visible.visibility), ovis[ovis > 0]) < 1e-4


GROUPS = dict(position=dict(lr=0.3, type="local_vector"), log_scaling=dict(lr=0.08, type="scalar"),
              rotation=dict(lr=0.01, type="vector"), alpha_logit=dict(lr=0.1, type="scalar"),
              feature=dict(lr=0.5, type="vector"))
KIND = dict(scalar=0, vector=1, local_vector=2)
OPTS = dict(betas=(0.8, 0.95), vis_beta=0.999, bias_correction=True, vis_smooth=0.01, grad_clip=2.0)


def test_parameter_class_step_matches_oracle():
  """ParameterClass(point_tensors, parameter_groups, optimizer=VisibilityAwareLaProp, ...).step(visibility, indexes,
  basis) as called at scene/mlp_scene.py:79,226-228 — two steps, against the oracle's per-group optimiser."""
  n = 3000
  g = random_gaussians(n, seed=3, sh_degree=0)
  rng = np.random.default_rng(9)
  tensors = dict(position=g["position"], log_scaling=g["log_scaling"], rotation=g["rotation"],
                 alpha_logit=g["alpha_logit"][:, None], feature=rng.normal(0, 1, (n, 16)).astype(np.float32),
                 visible=np.zeros(n, np.float32))
  pc = ParameterClass({k: cu(v) for k, v in tensors.items()}, GROUPS, optimizer=VisibilityAwareLaProp, **OPTS)
  assert isinstance(pc.optimizer, VisibilityOptimizer) and set(pc.keys()) == set(tensors)
  assert pc.position.requires_grad and not pc.visible.requires_grad
  orc = oracle()
  ocfg = go.make_optim_config(mode=1, beta1=0.8, beta2=0.95, vis_beta=0.999, vis_smooth=0.01, grad_clip=2.0, bias_correction=1)
  ref = {k: v.copy() for k, v in tensors.items()}
  m1 = {k: np.zeros_like(ref[k]) for k in GROUPS}
  m2 = {k: (np.zeros_like(ref[k]) if GROUPS[k]["type"] == "scalar" else np.zeros(n, np.float32)) for k in GROUPS}
  running, step = np.zeros(n, np.float32), np.zeros(n, np.float32)
  for it in range(2):
    vis = (rng.uniform(0, 1, n) * (rng.uniform(0, 1, n) < 0.4)).astype(np.float32)
    grads = {k: rng.normal(0, 1, ref[k].shape).astype(np.float32) * (vis > 0).reshape((n,) + (1,) * (ref[k].ndim - 1)) for k in GROUPS}
    for k in GROUPS:
      getattr(pc, k).grad = cu(grads[k])
    idx = torch.nonzero(cu(vis)).squeeze(1)
    pc.step(visibility=cu(vis)[idx], indexes=idx, basis=None)
    pc.zero_grad()
    orc.optim_vis_update(ocfg, vis, running, step)
    bls, bq = ref["log_scaling"].copy(), ref["rotation"].copy()
    for k in ("position", "log_scaling", "rotation", "alpha_logit", "feature"):
      orc.optim_step(ocfg, KIND[GROUPS[k]["type"]], GROUPS[k]["lr"], ref[k], grads[k], m1[k], m2[k], vis, running, step,
                     bls, bq, 0, 0.0, 0.0, 0)
  for k in GROUPS:
    assert rel_err(npy(getattr(pc, k)), ref[k]) < 1e-5, k
    assert rel_err(npy(pc.tensor_state[k]["exp_avg"]), m1[k]) < 1e-4, k
  # row selection keeps parameters AND optimiser state; appended rows start with zero state (mlp_scene.py:306-310)
  keep = cu((rng.uniform(0, 1, n) < 0.7))
  kept = pc[keep]
  k0 = int(keep.sum().item())
  assert kept.position.shape[0] == k0
  assert torch.equal(kept.tensor_state["position"]["exp_avg"], pc.tensor_state["position"]["exp_avg"][keep])
  grown = kept.append_tensors({k: v.data[:5] for k, v in pc.tensors.items()})
  assert grown.position.shape[0] == k0 + 5 and float(grown.tensor_state["feature"]["exp_avg"][k0:].abs().sum()) == 0.0
  # checkpoint round trip with plain tensors (trainer/checkpoint.py:38 loads with weights_only=True)
  sd = grown.state_dict()
  back = ParameterClass.from_state_dict(sd, optimizer=VisibilityAwareLaProp, **OPTS)
  assert torch.equal(back.rotation.data, grown.rotation.data)
  assert torch.equal(back.tensor_state["rotation"]["exp_avg_sq"], grown.tensor_state["rotation"]["exp_avg_sq"])
  assert back.update_groups(position=dict(lr=0.1))["position"] == 0.1


def test_evaluate_sh_at_and_queue():
  n = 2000
  g = random_gaussians(n, seed=11, sh_degree=3)
  sh, pos = cu(g["sh"]).requires_grad_(True), cu(g["position"])
  idx = torch.arange(0, n, 3, device="cuda")
  cam_pos = torch.tensor([3.0, 0.4, 0.8], device="cuda")
  ts.TaichiQueue.init(arch=None, threaded=True)
  rgb = ts.TaichiQueue.run_sync(ts.evaluate_sh_at, sh, pos, idx, cam_pos)
  rgb.sum().backward()
  orc = oracle()
  ref = orc.sh_eval_fwd(npy(idx), g["sh"], 3, g["position"], np.array([3.0, 0.4, 0.8], np.float32))
  assert np.abs(npy(rgb) - ref).max() < 1e-5
  assert float(sh.grad[1].abs().sum()) == 0.0 and float(sh.grad[0].abs().sum()) > 0.0
