"""The MLP colour scene (SceneConfig.color_mlp: the reference's default `scene: mlp`) end to end on the GPU: render uses
the colour MLP for the in-view rows, training moves point features, GLO rows and MLP weights, and fits a teacher."""
import numpy as np
import pytest
import torch

from gsb200 import lib as L
from gsb200 import ops
from gsb200 import synthetic as syn
from gsb200.scene import SceneConfig, SplatScene, camera_position
from gsb200.trainer import Trainer
from gpu_util import npy
from oracle import gs_oracle as go

pytestmark = pytest.mark.gpu
W, H, N = 320, 240, 20000


def _gaussians(seed, feature_scale=1.0):
  g = syn.garden_scene(N, 0, seed=3, focal=0.5 * W / np.tan(np.radians(25.0)))
  rng = np.random.default_rng(seed)
  g = dict(g)
  del g["sh"]
  g["feature"] = (rng.normal(0, 1, (N, 16)) * feature_scale).astype(np.float32)
  return g


def _cams(k):
  return [L.make_camera(c["T"], c["fx"], c["fy"], c["cx"], c["cy"], W, H) for c in syn.ring_cameras(8, W, H)[:k]]


def _scene(seed, **kw):
  return SplatScene(_gaussians(seed), SceneConfig(color_mlp=True, num_images=4, **kw))


def test_render_uses_the_colour_mlp_for_in_view_rows():
  scene = _scene(1)
  with torch.no_grad():
    for k, v in scene.color.w.items():  # leave the near-linear range of the fresh initialisation
      v.mul_(2.0)
    scene.color.glo[1].normal_(0, 0.5, generator=torch.Generator(device="cuda").manual_seed(2))
  cam = _cams(2)[1]
  out = scene.render(cam, image_idx=1)
  raw = out["raw"].clone()  # render returns views of the scene's reusable buffers
  w = scene._work
  in_view = npy(w["g2d"][:, 5]) > 0
  assert 1000 < in_view.sum() < N
  weights = {k: npy(v) for k, v in scene.color.w.items()}
  d, s = go.Oracle("f32").color_mlp_fwd(weights, npy(scene.p["sh"]), npy(scene.p["position"]),
                                        np.array(camera_position(cam), np.float32), npy(scene.color.glo[1]),
                                        in_view=in_view.astype(np.float32), half=True)
  rgb = npy(w["rgb"])
  err = np.abs(rgb[in_view] - (d + s)[in_view])
  assert (err <= 4e-3 * np.maximum(np.abs((d + s)[in_view]), 0.25)).all(), err.max()
  # image == rasterising the oracle's colours with the same overlap list (colour enters the blend linearly)
  tb = w["tile"]
  ref_rgb = torch.from_numpy((d + s).astype(np.float32)).cuda()
  img, _, _, _, _ = ops.raster_fwd(scene.raster_cfg, w["g2d"], ref_rgb, w["depth"], tb.overlap_to_point, tb.tile_ranges, W, H)
  assert (raw - img).abs().max().item() < 6e-3
  # a view that is not a training image uses a zero GLO feature (reference scene/mlp_scene.py:356-360)
  out0 = scene.render(cam, image_idx=None)
  assert (out0["raw"] - raw).abs().max().item() > 1e-3


def test_training_fits_a_teacher_and_touches_only_blended_rows():
  teacher = _scene(5, )
  with torch.no_grad():
    for k, v in teacher.color.w.items():
      v.mul_(2.0)
  cams = _cams(3)
  targets = [teacher.render(c, image_idx=None)["image"].clone() for c in cams]
  student = _scene(9, reg_weights=(0.0, 0.0, 0.0))
  # same geometry, different features / weights: only the colour path has anything to learn
  for name in ("position", "log_scaling", "rotation", "alpha_logit"):
    student.p[name].copy_(teacher.p[name])
  tr = Trainer(student)
  before = {k: v.clone() for k, v in student.p.items()}
  w_before = student.color.flat.clone()
  losses = []
  for it in range(40):
    k = it % 3
    loss = tr.training_step([(cams[k], targets[k], k)], read_loss=True)
    losses.append(loss[1])
  assert np.isfinite(losses).all()
  first, last = np.mean(losses[:3]), np.mean(losses[-3:])
  assert last < 0.6 * first, (first, last)
  assert not torch.equal(student.color.flat, w_before), "MLP weights are trained"
  assert student.color.glo[:3].abs().max().item() > 0 and not student.color.glo[3].any(), "only the used GLO rows move"
  assert not student.grads.any() and not student.color.grad.any() and not student.color.glo_grad.any()
  changed = (student.p["sh"] != before["sh"]).any(dim=1)
  seen = student.stats["points_in_view"] > 0
  assert int(changed.sum().item()) > 100 and not (changed & ~seen).any()


def test_batch_accumulation_equals_sum_of_single_image_gradients():
  """two images in one step: feature gradients are ADDED (batch = sum, reference trainer/trainer.py:500-514)."""
  cams = _cams(2)
  gen = torch.Generator(device="cuda").manual_seed(4)
  targets = [torch.rand((H, W, 3), device="cuda", generator=gen) for _ in cams]
  singles = []
  for k in range(2):
    s = _scene(1, reg_weights=(0.0, 0.0, 0.0))
    s.forward_backward(cams[k], targets[k], first=True, image_idx=k)
    singles.append((s.g["sh"].clone(), s.color.grad.clone(), s.color.glo_grad.clone()))
  both = _scene(1, reg_weights=(0.0, 0.0, 0.0))
  both.forward_backward(cams[0], targets[0], first=True, image_idx=0)
  both.forward_backward(cams[1], targets[1], first=False, image_idx=1)
  # the raster backward's fp32 atomics reorder from run to run, and a last-bit change of d_rgb can flip an fp16 rounding
  # inside the colour backward: compare against the size of the gradients, not element by element
  def close(a, b, rel):
    assert (a - b).abs().max().item() <= rel * b.abs().max().item(), ((a - b).abs().max().item(), b.abs().max().item())
  close(both.g["sh"], singles[0][0] + singles[1][0], 2e-3)
  close(both.color.grad, singles[0][1] + singles[1][1], 2e-3)
  close(both.color.glo_grad, singles[0][2] + singles[1][2], 2e-3)


def test_densify_and_checkpoint_keep_the_feature_plane_and_colour_state():
  scene = _scene(2, reg_weights=(0.0, 0.0, 0.0))
  cams = _cams(2)
  gen = torch.Generator(device="cuda").manual_seed(8)
  target = torch.rand((H, W, 3), device="cuda", generator=gen)
  tr = Trainer(scene)
  for it in range(3):
    tr.training_step([(cams[it % 2], target, it % 2)])
  # split 50 rows, prune 100 others: children inherit the parent's feature row (reference gaussians/split.py:87-113)
  keep = torch.ones((N,), device="cuda", dtype=torch.bool)
  split_idx = torch.arange(200, 250, device="cuda")
  keep[split_idx] = False
  keep[1000:1100] = False
  parents = scene.p["sh"][split_idx].clone()
  survivor = scene.p["sh"][5].clone()
  old_n, new_n = scene.split_and_prune(keep, split_idx)
  assert new_n == N - 150 + 100 and scene.p["sh"].shape == (new_n, 16)
  assert torch.equal(scene.p["sh"][5], survivor)
  kids = scene.p["sh"][new_n - 100:].reshape(50, 2, 16)
  assert torch.equal(kids[:, 0], parents) and torch.equal(kids[:, 1], parents)
  tr.training_step([(cams[0], target, 0)])  # the resized scene trains on
  assert torch.isfinite(scene.params).all()
  # checkpoint round trip: same render, same colour-model state
  sd = scene.state_dict()
  clone = SplatScene.from_state_dict(sd)
  assert clone.color is not None and torch.equal(clone.color.flat, scene.color.flat) and torch.equal(clone.color.glo, scene.color.glo)
  a = scene.render(cams[1], image_idx=1)["raw"].clone()
  b = clone.render(cams[1], image_idx=1)["raw"]
  assert torch.equal(a, b)
