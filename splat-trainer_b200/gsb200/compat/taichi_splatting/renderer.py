"""project_to_image / render_projected / render_gaussians over libgsb200 (reference call sites
scene/mlp_scene.py:375-377,415-418; scripts/test_split.py:30).  Each differentiable op is a torch.autograd.Function
whose backward calls the matching `_bwd` entry point of the C-ABI; nothing is computed with torch ops except the
row gather / scatter between the N-row scene and the V in-view rows (the reference's own `indexes` contract)."""
import torch

from gsb200 import lib as L
from gsb200 import ops

from .data_types import Gaussians3D, RasterConfig, RenderedPoints, Rendering


def _lib_camera(cam):
  T = cam.T_camera_world.detach().to("cpu", torch.float64).numpy()
  fx, fy, cx, cy = [float(x) for x in cam.projection.detach().cpu().tolist()]
  W, H = int(cam.image_size[0]), int(cam.image_size[1])
  return L.make_camera(T, fx, fy, cx, cy, W, H, float(cam.near_plane), float(cam.far_plane))


def _lib_config(config: RasterConfig, median=False):
  if config.tile_size != 16:
    raise L.GsError("libgsb200 supports tile_size 16 only")
  return L.default_raster_config(antialias=int(config.antialias), blur_cov=float(config.blur_cov),
                                 alpha_threshold=float(config.alpha_threshold),
                                 clamp_max_alpha=float(config.clamp_max_alpha),
                                 saturate_threshold=float(config.saturate_threshold),
                                 compute_visibility=int(config.compute_visibility),
                                 compute_point_heuristic=int(config.compute_point_heuristic),
                                 render_median_depth=int(median))


class _Project(torch.autograd.Function):
  @staticmethod
  def forward(ctx, position, log_scaling, rotation, alpha_logit, cam, cfg):
    pos, ls, rot = position.contiguous(), log_scaling.contiguous(), rotation.contiguous()
    al = alpha_logit.reshape(-1).contiguous()
    g2d, depth, _ = ops.project_fwd(cam, cfg, pos, ls, rot, al)
    idx = (g2d[:, 5] > 0).nonzero().squeeze(1)  # host sync, like the reference's own compaction
    ctx.cam, ctx.cfg, ctx.al_shape = cam, cfg, alpha_logit.shape
    ctx.save_for_backward(pos, ls, rot, al, g2d, idx)
    ctx.mark_non_differentiable(idx)
    return g2d[idx], depth[idx].unsqueeze(1), idx

  @staticmethod
  def backward(ctx, d_g2d, d_depth, _):
    pos, ls, rot, al, g2d, idx = ctx.saved_tensors
    n = pos.shape[0]
    dg = torch.zeros((n, 6), device=pos.device)
    dd = torch.zeros((n,), device=pos.device)
    if d_g2d is not None:
      dg[idx] = d_g2d.to(torch.float32)
    if d_depth is not None:
      dd[idx] = d_depth.reshape(-1).to(torch.float32)
    g = ops.project_bwd(ctx.cam, ctx.cfg, pos, ls, rot, al, None, 0, g2d, dg, None, d_depth=dd)
    return g["position"], g["log_scaling"], g["rotation"], g["alpha_logit"].reshape(ctx.al_shape), None, None


def project_to_image(gaussians: Gaussians3D, camera_params, config: RasterConfig = RasterConfig()):
  """-> (gaussians2d (V,6), depth (V,1), indexes (V,) int64): the in-view rows of the scene, compacted."""
  cam, cfg = _lib_camera(camera_params), _lib_config(config)
  return _Project.apply(gaussians.position, gaussians.log_scaling, gaussians.rotation, gaussians.alpha_logit, cam, cfg)


class _Raster(torch.autograd.Function):
  @staticmethod
  def forward(ctx, g2d, features, depth, cfg, size, flags, heur):
    W, H = size
    g2d = g2d.contiguous().to(torch.float32)
    feat = features.contiguous().to(torch.float32)
    d1 = depth.reshape(-1).contiguous().to(torch.float32)
    if feat.shape[1] not in (1, 3, 4):
      raise L.GsError(f"render_projected: {feat.shape[1]} feature channels (libgsb200 instantiates 1, 3, 4)")
    o2p, ranges, _, _ = ops.tile_map(cfg, g2d, d1, W, H)  # exact sizing: one count read-back, as the reference does
    image, final_T, n_contrib, vis, med = ops.raster_fwd(cfg, g2d, feat, d1, o2p, ranges, W, H,
                                                         visibility=flags["visibility"], median=flags["median"])
    ctx.cfg, ctx.size, ctx.heur, ctx.ch = cfg, size, heur, feat.shape[1]
    ctx.save_for_backward(g2d, feat, o2p, ranges, final_T, n_contrib)
    weight = 1.0 - final_T
    vis = vis if vis is not None else torch.zeros((g2d.shape[0],), device=g2d.device)
    med = med if med is not None else torch.zeros((0,), device=g2d.device)
    ctx.mark_non_differentiable(weight, vis, med)
    return image, weight, vis, med

  @staticmethod
  def backward(ctx, d_image, *_):
    g2d, feat, o2p, ranges, final_T, n_contrib = ctx.saved_tensors
    W, H = ctx.size
    acc = ops.raster_bwd(ctx.cfg, g2d, feat, o2p, ranges, W, H, final_T, n_contrib, d_image.contiguous().to(torch.float32))
    c = ctx.ch
    if ctx.heur is not None:  # the tensors handed out in forward are filled in place (read after backward)
      ctx.heur["prune_cost"].copy_(acc[:, 6 + c])
      ctx.heur["split_score"].copy_(acc[:, 7 + c])
    return acc[:, 0:6].contiguous(), acc[:, 6:6 + c].contiguous(), None, None, None, None, None


def render_projected(indexes, gaussians2d, features, depth, camera_params, config: RasterConfig = RasterConfig(),
                     render_median_depth: bool = False) -> Rendering:
  W, H = int(camera_params.image_size[0]), int(camera_params.image_size[1])
  cfg = _lib_config(config, render_median_depth)
  v = gaussians2d.shape[0]
  heur = None
  if config.compute_point_heuristic:
    heur = dict(prune_cost=torch.zeros((v,), device=gaussians2d.device), split_score=torch.zeros((v,), device=gaussians2d.device))
  flags = dict(visibility=bool(config.compute_visibility), median=bool(render_median_depth))
  image, weight, vis, med = _Raster.apply(gaussians2d, features, depth, cfg, (W, H), flags, heur)
  points = RenderedPoints(idx=indexes, gaussians2d=gaussians2d, depths=depth,
                          prune_cost=None if heur is None else heur["prune_cost"],
                          split_score=None if heur is None else heur["split_score"],
                          visibility=vis if config.compute_visibility else None)
  return Rendering(image=image, image_weight=weight, points=points, camera=camera_params, config=config,
                   median_depth_image=med if render_median_depth else None)


def render_gaussians(gaussians: Gaussians3D, camera_params, config: RasterConfig = RasterConfig(), use_sh: bool = False,
                     render_median_depth: bool = False) -> Rendering:
  """project + colour + rasterise.  feature (N,C) is used as is; (N,3,K) with use_sh is evaluated with evaluate_sh_at."""
  g2d, depth, idx = project_to_image(gaussians, camera_params, config)
  if use_sh or gaussians.feature.dim() == 3:
    from .spherical_harmonics import evaluate_sh_at
    feat = evaluate_sh_at(gaussians.feature, gaussians.position, idx, camera_params.camera_position)
  else:
    feat = gaussians.feature[idx]
  return render_projected(idx, g2d, feat, depth, camera_params, config, render_median_depth=render_median_depth)
