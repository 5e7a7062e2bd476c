"""Types the reference reads at the `taichi_splatting` boundary (SURVEY.md §8b "operator interface").

Plain dataclasses over torch tensors (no tensordict dependency): `Gaussians3D`, `RasterConfig`, `RenderedPoints`,
`Rendering`.  Field and property names follow the reference's call sites, cited per member.
"""
import dataclasses
from dataclasses import dataclass
from typing import Optional

import torch


def _index(v, idx):
  return v[idx] if isinstance(v, torch.Tensor) else v


@dataclass
class Gaussians3D:
  """position (N,3), log_scaling (N,3), rotation (N,4) xyzw, alpha_logit (N,1), feature (N,F) | (N,3,K)
  (reference gaussians/loading.py:26-34, scene/io.py:13-65, scene/mlp_scene.py:403-407)."""
  position: torch.Tensor
  log_scaling: torch.Tensor
  rotation: torch.Tensor
  alpha_logit: torch.Tensor
  feature: torch.Tensor

  FIELDS = ("position", "log_scaling", "rotation", "alpha_logit", "feature")

  @property
  def batch_size(self):
    return (self.position.shape[0],)

  @property
  def device(self):
    return self.position.device

  def to_dict(self):
    return {k: getattr(self, k) for k in self.FIELDS}

  def to_tensordict(self):  # the reference only calls .replace(...).to(device) / .items() on the result
    return TensorBag(self.to_dict())

  @staticmethod
  def from_dict(d, batch_dims=1):
    return Gaussians3D(**{k: d[k] for k in Gaussians3D.FIELDS})

  from_tensordict = from_dict

  def apply(self, f):
    return Gaussians3D(**{k: f(v) for k, v in self.to_dict().items()})

  def to(self, *args, **kwargs):
    return self.apply(lambda t: t.to(*args, **kwargs))

  def detach(self):
    return self.apply(lambda t: t.detach())

  def __getitem__(self, idx):
    return self.apply(lambda t: t[idx])

  def alpha(self):
    return torch.sigmoid(self.alpha_logit)

  def scale(self):
    return torch.exp(self.log_scaling)

  def translated(self, t):
    return dataclasses.replace(self, position=self.position + t)

  def scaled(self, s):
    import math
    return dataclasses.replace(self, position=self.position * s, log_scaling=self.log_scaling + math.log(s))


class TensorBag(dict):
  """The sliver of TensorDict the reference uses on point tensors: replace / to / select / item access."""

  def replace(self, **kw):
    out = TensorBag(self)
    out.update(kw)
    return out

  def to(self, *args, **kwargs):
    return TensorBag({k: v.to(*args, **kwargs) for k, v in self.items()})

  def select(self, *names):
    return TensorBag({k: self[k] for k in names})

  @property
  def batch_size(self):
    return (next(iter(self.values())).shape[0],)


@dataclass(frozen=True)
class RasterConfig:
  """Names passed by the reference (trainer/trainer.py:308-310, scene/mlp_scene.py:373) + SURVEY Appendix A defaults."""
  tile_size: int = 16
  margin_tiles: int = 3
  gaussian_scale: float = 3.0
  antialias: bool = False
  blur_cov: float = 0.3
  clamp_max_alpha: float = 0.99
  alpha_threshold: float = 1.0 / 255.0
  saturate_threshold: float = 0.9999
  beta: float = 1.0
  depth16: bool = False
  use_alpha_blending: bool = True
  compute_point_heuristic: bool = False
  compute_visibility: bool = False
  pixel_stride: tuple = (2, 2)


@dataclass
class RenderedPoints:
  """Per-point outputs of one rendering (reference consumers: controller/point_state.py:36-50,
  scene/mlp_scene.py:243-244,269-279,380-381; trainer/trainer.py:507).  All tensors have V rows."""
  idx: torch.Tensor                 # (V,) int64 rows of the scene
  gaussians2d: torch.Tensor         # (V,6) mean.xy, conic a b c, opacity
  depths: torch.Tensor              # (V,1)
  prune_cost: Optional[torch.Tensor] = None   # (V,) filled by the raster BACKWARD (trainer/trainer.py:511-514)
  split_score: Optional[torch.Tensor] = None  # (V,)
  visibility: Optional[torch.Tensor] = None   # (V,) Σ α·T
  attributes: object = None

  @property
  def batch_size(self):
    return (self.idx.shape[0],)

  @property
  def opacity(self):
    return self.gaussians2d[:, 5]

  @property
  def screen_scale(self):
    """(V,2) std-devs in px of the 2-D Gaussian along its major / minor axis (controller/point_state.py:37)."""
    a, b, c = self.gaussians2d[:, 2], self.gaussians2d[:, 3], self.gaussians2d[:, 4]
    det = (a * c - b * b).clamp_min(1e-20)
    cxx, cyy, cxy = c / det, a / det, -b / det
    mid = 0.5 * (cxx + cyy)
    disc = (mid * mid - (cxx * cyy - cxy * cxy)).clamp_min(0).sqrt()
    return torch.stack([(mid + disc).clamp_min(0).sqrt(), (mid - disc).clamp_min(0).sqrt()], dim=1)

  @property
  def visible_mask(self):
    return self.visibility > 0

  @property
  def visible(self):
    return self[self.visible_mask]

  @property
  def num_visible(self):
    return int(self.visible_mask.sum().item())

  def __getitem__(self, sel):
    f = lambda v: _index(v, sel)
    attrs = self.attributes
    if attrs is not None and hasattr(attrs, "__getitem__") and not isinstance(attrs, (dict, str)):
      try:
        attrs = attrs[sel]
      except Exception:
        pass
    return RenderedPoints(idx=f(self.idx), gaussians2d=f(self.gaussians2d), depths=f(self.depths),
                          prune_cost=f(self.prune_cost), split_score=f(self.split_score),
                          visibility=f(self.visibility), attributes=attrs)

  def replace(self, **kw):
    return dataclasses.replace(self, **kw)

  def detach(self):
    d = lambda v: v.detach() if isinstance(v, torch.Tensor) else v
    return RenderedPoints(**{f.name: d(getattr(self, f.name)) for f in dataclasses.fields(self)})


@dataclass
class Rendering:
  """Result bundle of render_projected (reference consumers: trainer/trainer.py:320,356-359,468,507,519-528;
  scene/mlp_scene.py:422-424; trainer/evaluation.py:27,49).  A dataclass because the reference calls
  dataclasses.replace on it."""
  image: torch.Tensor               # (H,W,C)
  image_weight: torch.Tensor        # (H,W) = 1 − final transmittance
  points: RenderedPoints
  camera: object
  config: RasterConfig
  median_depth_image: Optional[torch.Tensor] = None

  @property
  def image_size(self):
    return self.camera.image_size

  @property
  def median_ndc_image(self):
    """median depth mapped to [0,1] between the near and far planes in inverse depth"""
    d = self.median_depth_image
    near, far = self.camera.near_plane, self.camera.far_plane
    ndc = (1.0 / d.clamp_min(1e-12) - 1.0 / near) / (1.0 / far - 1.0 / near)
    return torch.where(d > 0, ndc, torch.ones_like(d))

  @property
  def point_depth(self):
    return self.points.depths

  @property
  def visible_mask(self):
    return self.points.visible_mask

  @property
  def visible_indices(self):
    return self.points.idx[self.points.visible_mask]

  def detach(self):
    d = lambda v: v.detach() if isinstance(v, torch.Tensor) else v
    return Rendering(image=d(self.image), image_weight=d(self.image_weight), points=self.points.detach(),
                     camera=self.camera, config=self.config, median_depth_image=d(self.median_depth_image))
