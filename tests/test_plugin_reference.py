"""The reference's UNCHANGED controller modules driving the gsb200 scene plugin.

Runs in the build container only (needs /root/reference; skipped on the GPU box): the reference's
controller/target_controller.py, controller/point_state.py, config/__init__.py and gaussians/split.py are imported from
where they lie, over the operator shim (gsb200/compat/taichi_splatting) and the third-party stand-ins
(gsb200/compat/standins), and `TargetController.step` / `add_rendering` are run against `gsb200.plugin.SHScene` on a CPU
device.  The one library kernel on that path (gs_split_positions) is replaced by the CPU oracle's restatement for the
occasion — test infrastructure; the product has no CPU fallback."""
import os
import sys

import numpy as np
import pytest
import torch

REF = "/root/reference"
pytestmark = pytest.mark.skipif(not os.path.isdir(os.path.join(REF, "splat_trainer")), reason="reference tree absent")


@pytest.fixture(scope="module")
def ref():
  from gsb200 import compat
  compat.enable(REF)
  import torch._dynamo
  torch._dynamo.config.disable = True  # the reference wraps small helpers in torch.compile; this image's host compiler
                                       # cannot build inductor's C++ (no libgomp.spec), and eager is the same arithmetic
  import splat_trainer.scene.scene  # the reference's own import order (its entry points import the scene package first)
  import splat_trainer.controller.target_controller as tc
  import splat_trainer.controller.point_state as ps
  import splat_trainer.config as cfg
  import splat_trainer.gaussians.split as split
  from splat_trainer.logger.logger import NullLogger
  from gsb200 import plugin
  assert plugin.register_with_reference()
  return dict(tc=tc, ps=ps, cfg=cfg, split=split, NullLogger=NullLogger, plugin=plugin)


def _oracle_split(monkeypatch, plugin):
  from oracle import gs_oracle as go
  orc = go.Oracle("f32")

  def split_positions(split_idx, position, log_scaling, rotation, uniform, sep, out_position, out_log_scaling, out_offset=0):
    p, ls = orc.split_positions(split_idx.numpy(), position.numpy(), log_scaling.numpy(), rotation.numpy(), uniform.numpy(), sep)
    s2 = p.shape[0]
    out_position[out_offset:out_offset + s2] = torch.from_numpy(p)
    out_log_scaling[out_offset:out_offset + s2] = torch.from_numpy(ls)

  monkeypatch.setattr(plugin.ops, "split_positions", split_positions)


def _scene(ref, n=4000, seed=0):
  from taichi_splatting import Gaussians3D
  g = torch.Generator().manual_seed(seed)
  gaussians = Gaussians3D(position=torch.randn((n, 3), generator=g), log_scaling=torch.randn((n, 3), generator=g) * 0.3 - 2,
                          rotation=torch.nn.functional.normalize(torch.randn((n, 4), generator=g), dim=1),
                          alpha_logit=torch.randn((n, 1), generator=g), feature=torch.rand((n, 3), generator=g))
  config = ref["plugin"].SHSceneConfig(sh_degree=2)
  return config, config.from_color_gaussians(gaussians, None, torch.device("cpu"), ref["NullLogger"]())


def test_reference_modules_import_over_the_shims(ref):
  import splat_trainer.trainer.trainer as trainer  # the whole training loop module: every import resolves
  import splat_trainer.scene.mlp_scene as mlp_scene
  import taichi_splatting
  assert "gsb200" in taichi_splatting.__file__
  assert trainer.Trainer.training_step and mlp_scene.MLPScene.render
  from splat_trainer.scene.scene import GaussianScene, GaussianSceneConfig
  config, scene = _scene(ref, 50)
  assert isinstance(scene, GaussianScene) and isinstance(config, GaussianSceneConfig)
  # every abstract member of the reference's scene interface exists on the plugin
  for name in GaussianScene.__abstractmethods__:
    assert hasattr(scene, name), name
  for name in GaussianSceneConfig.__abstractmethods__:
    assert hasattr(config, name), name


def test_unchanged_target_controller_densifies_the_plugin_scene(ref, monkeypatch):
  tc, cfg = ref["tc"], ref["cfg"]
  _oracle_split(monkeypatch, ref["plugin"])
  n, target, total = 4000, 6000, 1000
  config, scene = _scene(ref, n)
  tconf = tc.TargetConfig(prune_rate=0.025, target_count_t=0.8, min_views=5, max_scale_px=200.0,
                          densify_prune_interval=cfg.Piecewise(50, [(0.1, 100), (0.5, 200)]))
  progress = cfg.Progress(step=0, total_steps=total)
  controller = tconf.make_controller(scene, target, progress, ref["NullLogger"]())  # beartype-checked: scene is a GaussianScene
  assert controller.next_densify == 50 and controller.points.batch_size[0] == n
  # statistics of 8 renderings through the reference's own PointState.add_rendering, fed the shim's RenderedPoints
  from taichi_splatting import RasterConfig, RenderedPoints, Rendering
  g = torch.Generator().manual_seed(1)
  for image in range(8):
    v = 3000
    idx = torch.randperm(n, generator=g)[:v].sort().values
    g2d = torch.zeros((v, 6))
    g2d[:, 2] = 1.0 / (torch.rand((v,), generator=g) * 30 + 0.5) ** 2
    g2d[:, 4] = 1.0 / (torch.rand((v,), generator=g) * 30 + 0.5) ** 2
    g2d[:, 5] = torch.rand((v,), generator=g)
    vis = torch.rand((v,), generator=g) * (torch.rand((v,), generator=g) > 0.2)
    pts = RenderedPoints(idx=idx, gaussians2d=g2d, depths=torch.rand((v, 1), generator=g) + 1,
                         prune_cost=torch.rand((v,), generator=g) * vis, split_score=torch.rand((v,), generator=g) * vis,
                         visibility=vis)
    rendering = Rendering(image=torch.zeros((4, 4, 3)), image_weight=torch.zeros((4, 4)), points=pts, camera=None,
                          config=RasterConfig())
    controller.add_rendering(image, rendering, progress)
    scene.add_rendering(image, rendering)
  assert int((controller.points.points_in_view > 0).sum()) > 0.9 * n
  assert float(scene.points.visible.sum()) > 0
  first_rows = {k: getattr(scene.points, k).data.clone() for k in scene.points.keys()}
  # not yet due
  controller.step(cfg.Progress(step=49, total_steps=total))
  assert scene.num_points == n
  # due: the reference computes the masks (two argsorts) and calls scene.split_and_prune(keep_mask, split_idx)
  stats = controller.points
  progress = cfg.Progress(step=50, total_steps=total)
  want = controller.target_points(progress)
  split_mask, prune_mask = controller.find_split_prune_indexes(progress.t, want)
  controller.points = stats  # find_split_prune_indexes edits split_score in place; step() recomputes the same masks
  controller.step(progress)
  assert scene.num_points == want == n - int(prune_mask.sum()) + int(split_mask.sum())  # K3: the count invariant
  assert controller.points.batch_size[0] == want and not controller.points.prune_cost.any()  # fresh PointState
  assert controller.next_densify == 100
  # kept rows carry their parameters over in order; children obey K4 (gaussians/split.py:102-112)
  keep = ~(split_mask | prune_mask)
  kept = int(keep.sum())
  for k, v in first_rows.items():
    assert torch.equal(getattr(scene.points, k).data[:kept], v[keep]), k
  sidx = split_mask.nonzero().squeeze(1)
  pos, ls = scene.points.position.data[kept:], scene.points.log_scaling.data[kept:]
  parent_pos, parent_ls = first_rows["position"][sidx], first_rows["log_scaling"][sidx]
  mid = 0.5 * (pos[0::2] + pos[1::2])
  assert torch.allclose(mid, parent_pos, atol=1e-5)
  changed = (ls[0::2] - parent_ls).abs() > 1e-6
  assert (changed.sum(1) == 1).all() and torch.allclose((ls[0::2] - parent_ls)[changed], torch.full((int(changed.sum()),), float(np.log(1 / np.sqrt(2)))), atol=1e-5)
  assert torch.equal(ls[0::2], ls[1::2])
  basis = ref["split"].point_basis(parent_ls, first_rows["rotation"][sidx])  # the reference's own R·S basis
  axis = changed.float().argmax(1)
  offset = 0.7 * basis[torch.arange(sidx.shape[0]), :, axis]
  assert torch.allclose((pos[0::2] - parent_pos).abs(), offset.abs(), atol=1e-4)
  for k in ("rotation", "alpha_logit", "feature"):
    assert torch.equal(getattr(scene.points, k).data[kept:][0::2], first_rows[k][sidx]), k
  # a second round through the schedule still works on the resized scene (optimiser state was resliced too)
  assert scene.points.tensor_state == {} or all(v["exp_avg"].shape[0] == want for v in scene.points.tensor_state.values())


def test_checkpoint_has_the_reference_layout_and_round_trips(ref):
  config, scene = _scene(ref, 300)
  sd = scene.state_dict()
  assert set(sd) == {"points"} and {"tensors", "parameter_groups", "state"} <= set(sd["points"])
  assert set(sd["points"]["tensors"]) == {"position", "log_scaling", "rotation", "alpha_logit", "feature", "visible"}
  clone = scene.clone()  # reference viewer: scene.clone() (viewer/splatview.py:65) = from_state_dict(state_dict())
  for k in scene.points.keys():
    assert torch.equal(getattr(clone.points, k).data, getattr(scene.points, k).data)
  lrs = clone.update_learning_rate(0.5)
  assert lrs["position"] == 0.3 and lrs["log_scaling"] == 0.08 and lrs["rotation"] == 0.01 and lrs["alpha_logit"] == 0.1
  # the trainer-level checkpoint dict of the reference (trainer/trainer.py:188-194) saves and loads with weights_only
  tc, cfg = ref["tc"], ref["cfg"]
  controller = tc.TargetConfig().make_controller(scene, 400, cfg.Progress(step=0, total_steps=1000), ref["NullLogger"]())
  ckpt = dict(step=123, scene=scene.state_dict(), controller=controller.state_dict())
  import io
  buf = io.BytesIO()
  torch.save(ckpt, buf)
  buf.seek(0)
  back = torch.load(buf, weights_only=True)
  assert back["step"] == 123 and torch.equal(back["scene"]["points"]["tensors"]["position"], scene.points.position.data)
  assert set(back["controller"]["points"]) >= {"prune_cost", "split_score", "max_scale_px", "points_in_view", "visibility"}
  # reg_loss schedule (config/scene/mlp.yaml:16-19 log_decay): weights at t = 0 and t = 1
  w = config.reg_weight
  assert abs(w["scale"](0.0) - 0.01) < 1e-12 and abs(w["scale"](1.0) - 1e-4) < 1e-12
  assert abs(w["opacity"](0.0) - 1.0) < 1e-12 and abs(w["opacity"](1.0) - 100.0) < 1e-9
