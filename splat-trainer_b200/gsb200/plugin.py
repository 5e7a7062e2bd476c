"""SHSceneConfig / SHScene — the scene plugin (L2) the reference's Trainer and controllers hold.

Implements the reference's `GaussianSceneConfig` / `GaussianScene` surface (scene/scene.py:17-94) member for member over
the operator shim (gsb200/compat/taichi_splatting -> libgsb200), with spherical-harmonic colour per splat in place of the
colour MLP: selected from hydra by
    scene:
      _target_: gsb200.plugin.SHSceneConfig
exactly where config/scene/mlp.yaml:6 names `splat_trainer.scene.MLPSceneConfig`.  The unchanged `Trainer`
(trainer/trainer.py:133,166,308,487,536,544,565,568), `TargetController` / `MCMCController`
(controller/target_controller.py:58,122; controller/point_state.py:107; controller/mcmc_controller.py:71,95) and the viewer
call only the members below.  Rendering goes through torch.autograd Functions whose backward runs the library's backward
kernels, so `loss.backward()` in the reference's loop (trainer/trainer.py:512) fills the per-point `.grad`s.

The fused, sync-free path of the same iteration is gsb200.scene.SplatScene + gsb200.trainer.Trainer (what bench.py
times); `SHScene.to_splat_scene()` / `from_splat_scene()` convert between the two, optimiser state included."""
import dataclasses
import math
import os
import sys
from dataclasses import dataclass, field
from typing import Optional

import torch
import torch.nn.functional as F

from . import compat as _compat

_compat.enable()

from taichi_splatting import Gaussians3D, RasterConfig, Rendering  # noqa: E402  (the operator shim, or the real package)
from taichi_splatting.optim import ParameterClass, VisibilityAwareLaProp, VisibilityOptimizer  # noqa: E402
from taichi_splatting.renderer import project_to_image, render_projected  # noqa: E402
from taichi_splatting.spherical_harmonics import evaluate_sh_at  # noqa: E402

from . import lib as L  # noqa: E402
from . import ops  # noqa: E402

SH0 = 0.282094791773878  # reference util/misc.py:41


def rgb_to_sh(rgb):
  return (rgb - 0.5) / SH0


def sh_to_rgb(sh):
  return sh * SH0 + 0.5


def _at(value, t):
  """a constant, or a schedule of training progress t in [0,1] (the reference's Varying objects are callables)."""
  return value(t) if callable(value) else value


def log_decay(start, factor):
  """the reference's `${log_decay:start, factor}` resolver (config/__init__.py:77-89): start · factor^t."""
  return lambda t: start * math.pow(factor, t)


def default_parameters():
  """config/scene/mlp.yaml:8-14; `feature` holds the SH coefficients here (rate: this repo's, see SceneConfig.lr_sh)."""
  return dict(position=dict(lr=0.3, type="local_vector"), log_scaling=dict(lr=0.08, type="scalar"),
              rotation=dict(lr=0.01, type="vector"), alpha_logit=dict(lr=0.1, type="scalar"),
              feature=dict(lr=0.01, type="vector"))


def default_reg_weight():
  """config/scene/mlp.yaml:16-19"""
  return dict(scale=log_decay(0.01, 0.01), opacity=log_decay(1.0, 100.0), aspect=1e-4)


_RASTER_KEYS = {f.name for f in dataclasses.fields(RasterConfig)}


def pop_raster_config(options):
  """split render(**options) into the RasterConfig fields and the rest (reference scene/util.py:13)."""
  cfg = {k: options.pop(k) for k in list(options) if k in _RASTER_KEYS}
  return RasterConfig(**cfg)


@dataclass
class SHSceneConfig:
  parameters: dict = field(default_factory=default_parameters)
  reg_weight: dict = field(default_factory=default_reg_weight)
  sh_degree: int = 3
  beta1: float = 0.8        # config/scene/mlp.yaml:25-31
  beta2: float = 0.95
  vis_beta: float = 0.999
  vis_smooth: float = 0.01
  grad_clip: Optional[float] = 2.0

  def optim_options(self):
    return dict(optimizer=VisibilityAwareLaProp, betas=(self.beta1, self.beta2), vis_beta=self.vis_beta,
                bias_correction=True, vis_smooth=self.vis_smooth, grad_clip=self.grad_clip)

  def groups(self, t):
    return {k: {a: _at(b, t) for a, b in g.items()} for k, g in self.parameters.items()}

  def from_color_gaussians(self, gaussians, camera_table, device, logger):
    """gaussians.feature: (N,3) RGB (the loaders' output, gaussians/loading.py:26-34) -> SH DC, higher bands zero; or
    (N,3,K') SH coefficients, cropped / zero-padded to this scene's degree."""
    K = (self.sh_degree + 1) ** 2
    n = gaussians.batch_size[0]
    f = gaussians.feature.to(torch.float32)
    sh = torch.zeros((n, 3, K))
    if f.dim() == 2:
      sh[:, :, 0] = rgb_to_sh(f[:, :3].cpu())
    else:
      k = min(K, f.shape[2])
      sh[:, :, :k] = f[:, :, :k].cpu()
    tensors = {k: getattr(gaussians, k).detach().to(device=device, dtype=torch.float32)
               for k in ("position", "log_scaling", "rotation", "alpha_logit")}
    tensors["alpha_logit"] = tensors["alpha_logit"].reshape(n, 1)
    tensors["feature"] = sh.to(device)
    tensors["visible"] = torch.zeros((n,), device=device)
    points = ParameterClass(tensors, parameter_groups=self.groups(0.0), **self.optim_options())
    return SHScene(points, self, camera_table, logger)

  def from_state_dict(self, state, camera_table, logger):
    points = ParameterClass.from_state_dict(state["points"], **self.optim_options())
    return SHScene(points, self, camera_table, logger)


class SHScene:
  def __init__(self, points, config, camera_table=None, logger=None):
    self.points, self.config, self.camera_table, self.logger = points, config, camera_table, logger
    self.update_learning_rate(0.0)

  # ---- properties -----------------------------------------------------------------------------------------
  @property
  def device(self):
    return self.points.position.device

  @property
  def num_points(self):
    return self.points.position.shape[0]

  def __repr__(self):
    return f"SHScene({self.num_points} points)"

  @property
  def gaussians(self):
    p = self.points
    return Gaussians3D(position=p.position, rotation=p.rotation, log_scaling=p.log_scaling, alpha_logit=p.alpha_logit,
                       feature=p.feature)

  def to_sh_gaussians(self):
    """already SH: the scene's own coefficients (N,3,K), detached (reference scene/mlp_scene.py:394-398 fits them)."""
    return self.gaussians.detach()

  @property
  def all_parameters(self):
    from tensordict import TensorDict
    return TensorDict(points=dict(self.points.tensors))

  def update_learning_rate(self, t):
    return self.points.update_groups(**self.config.groups(t))

  # ---- training-loop members -------------------------------------------------------------------------------
  @torch.no_grad()
  def zero_grad(self):
    self.points.visible.zero_()
    self.points.zero_grad()

  def render(self, camera_params, image_idx=None, point_mask=None, **options):
    """project -> SH colour -> tile map + rasterise -> clamp to [0,1] (scene/mlp_scene.py:410-427)."""
    config = pop_raster_config(options)
    g = self.gaussians
    g2d, depth, idx = project_to_image(g, camera_params, config)
    colors = evaluate_sh_at(g.feature, g.position, idx, camera_params.camera_position)
    rendering = render_projected(idx, g2d, colors, depth, camera_params, config, **options)
    return dataclasses.replace(rendering, points=rendering.points.replace(attributes=colors),
                               image=rendering.image.clamp(0, 1))

  @torch.no_grad()
  def query_visibility(self, camera_params):
    """(idx, visibility) of the splats a pixel blends (scene/mlp_scene.py:372-381): forward only, one channel."""
    config = RasterConfig(compute_visibility=True)
    g2d, depth, idx = project_to_image(self.gaussians, camera_params, config)
    feat = torch.zeros((idx.shape[0], 1), device=self.device)
    r = render_projected(idx, g2d, feat, depth, camera_params, config)
    vis = r.points.visible
    return vis.idx, vis.visibility

  @torch.no_grad()
  def add_rendering(self, image_idx, rendering):
    pts = rendering.points
    self.points.visible[pts.idx] += pts.visibility

  def reg_loss(self, rendering, progress):
    """Σ_k weight_k(t) · mean(w · term_k) over the rows the image blended (scene/mlp_scene.py:246-288, saturate
    util/misc.py:68-69); w = visibility for the visibility-aware optimisers."""
    pts = rendering.points.visible
    scale = torch.exp(self.points.log_scaling[pts.idx])
    w = pts.visibility if isinstance(self.points.optimizer, VisibilityOptimizer) else torch.ones((), device=self.device)
    norm_scale = scale.pow(2).sum(1) / pts.depths.pow(2).squeeze(-1)
    terms = dict(scale=norm_scale,
                 opacity=(1 - 1 / torch.exp(4.0 * pts.opacity)).pow(2.0) * norm_scale,
                 aspect=scale.max(1).values / scale.min(1).values)
    t = float(progress.t) if hasattr(progress, "t") else float(progress)
    weighted = {k: (v * w).mean() * _at(self.config.reg_weight[k], t) for k, v in terms.items() if k in self.config.reg_weight}
    if getattr(progress, "logging_step", False) and self.logger is not None:
      self.logger.log_values("train/reg", {k: float(v.item()) for k, v in weighted.items()})
    return sum(weighted.values())

  @torch.no_grad()
  def step(self, progress=None, log_details=False):
    """ParameterClass.step over the visible rows, then quaternion renormalise, log-scale clamp, zero_grad
    (scene/mlp_scene.py:214-239)."""
    vis_idx = self.points.visible.nonzero().squeeze(1)
    self.points.step(visibility=self.points.visible[vis_idx], indexes=vis_idx, basis=None)
    self.points.rotation.data = F.normalize(self.points.rotation.data, dim=1)
    self.points.log_scaling.data.clamp_(min=-8, max=8)
    self.zero_grad()

  @torch.no_grad()
  def split_and_prune(self, keep_mask, split_idx=None, uniform=None):
    """points = points[keep_mask] (parameters AND optimiser state), then two children per split row with zero state
    (scene/mlp_scene.py:301-313; children gaussians/split.py:87-113: random axis ~ scale, means at ±0.7 along it in the
    R·S basis, log-scale on that axis + log(1/√2), everything else duplicated)."""
    children = None
    if split_idx is not None and split_idx.numel() > 0:
      s = split_idx.shape[0]
      p = self.points
      if uniform is None:
        uniform = torch.rand((s,), device=self.device)
      pos = torch.empty((2 * s, 3), device=self.device)
      ls = torch.empty((2 * s, 3), device=self.device)
      ops.split_positions(split_idx, p.position.data, p.log_scaling.data, p.rotation.data, uniform, 0.7, pos, ls)
      children = dict(position=pos, log_scaling=ls)
      for k in self.points.keys():
        if k not in children:
          children[k] = getattr(p, k).data[split_idx].repeat_interleave(2, dim=0)
    self.points = self.points[keep_mask]
    if children is not None:
      self.points = self.points.append_tensors(children)

  def log_checkpoint(self, progress=None):
    if self.logger is None:
      return
    self.logger.log_histogram("params/opacity", torch.sigmoid(self.points.alpha_logit.detach()))
    self.logger.log_histogram("params/log_scale", self.points.log_scaling.detach())
    self.logger.log_histogram("params/feature", self.points.feature.detach())

  # ---- checkpointing (scene/mlp_scene.py:316-326) ------------------------------------------------------------------
  def state_dict(self):
    return dict(points=self.points.state_dict())

  def clone(self):
    return self.config.from_state_dict(self.state_dict(), self.camera_table, self.logger)

  # ---- the fused path ------------------------------------------------------------------------------------------------
  def to_splat_scene(self, **scene_config):
    """the same scene (parameters + optimiser state) as a gsb200.scene.SplatScene."""
    from .scene import SceneConfig, SplatScene
    p = self.points
    lrs = self.update_learning_rate(0.0)
    cfg = SceneConfig(sh_degree=self.config.sh_degree, lr_position=lrs["position"], lr_log_scaling=lrs["log_scaling"],
                      lr_rotation=lrs["rotation"], lr_alpha_logit=lrs["alpha_logit"], lr_sh=lrs["feature"],
                      optim=dict(beta1=self.config.beta1, beta2=self.config.beta2, vis_beta=self.config.vis_beta,
                                 vis_smooth=self.config.vis_smooth, grad_clip=self.config.grad_clip or 0.0), **scene_config)
    scene = SplatScene(dict(position=p.position.data, log_scaling=p.log_scaling.data, rotation=p.rotation.data,
                            alpha_logit=p.alpha_logit.data.reshape(-1), sh=p.feature.data), cfg, device=self.device)
    names = dict(position="position", log_scaling="log_scaling", rotation="rotation", alpha_logit="alpha_logit", sh="feature")
    for mine, theirs in names.items():
      st = p.tensor_state.get(theirs)
      if st:
        scene.m1[mine].copy_(st["exp_avg"].reshape(scene.m1[mine].shape))
        scene.m2[mine].copy_(st["exp_avg_sq"].reshape(scene.m2[mine].shape))
    shared = p.optimizer.shared
    if "step" in shared:
      scene.opt_step.copy_(shared["step"]); scene.running_vis.copy_(shared["running_vis"])
    scene.visible.copy_(p.visible.data)
    return scene


def register_with_reference():
  """make SHScene / SHSceneConfig (virtual) subclasses of the reference's ABCs so its beartype-checked signatures
  (e.g. TargetConfig.make_controller(scene: GaussianScene)) accept them.  No-op when splat_trainer is not importable."""
  try:
    from splat_trainer.scene.scene import GaussianScene, GaussianSceneConfig
  except Exception:
    return False
  GaussianScene.register(SHScene)
  GaussianSceneConfig.register(SHSceneConfig)
  return True


if "splat_trainer" in sys.modules or os.environ.get("GSB200_REGISTER_PLUGIN"):
  register_with_reference()
