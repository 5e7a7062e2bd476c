"""The hot path as ONE training iteration of the CPU oracle (oracle/gs_oracle.c through gs_oracle.py).

TEST INFRASTRUCTURE ONLY (see gs_oracle.py): the chain the -m gpu parity tests, tools/psnr_parity.py and bench.py's
cpu_baseline / --impl reference legs run next to the CUDA path.  It restates, stage for stage, what
gsb200.scene.SplatScene.forward_backward + step do for an SH scene — i.e. the reference's Trainer.training_step
(trainer/trainer.py:500-545): render -> loss (MSE + multi-scale SSIM, trainer/trainer.py:450-476) -> backward ->
reg_loss (scene/mlp_scene.py:246-288) -> add_rendering (controller/point_state.py:34-50) -> ParameterClass.step
(scene/mlp_scene.py:214-239).
"""
import numpy as np

from . import gs_oracle as go

ATTRS = ("position", "log_scaling", "rotation", "alpha_logit", "sh")
# reference config/scene/mlp.yaml:8-14 (group kinds and learning rates); `sh` is this repo's SH colour group
KIND = dict(position=2, log_scaling=0, rotation=1, alpha_logit=0, sh=1)
POST = dict(position=(0, 0.0, 0.0), log_scaling=(2, -8.0, 8.0), rotation=(1, 0.0, 0.0), alpha_logit=(0, 0.0, 0.0), sh=(0, 0.0, 0.0))


def psnr(image, target):
  """reference trainer/evaluation.py:29-31 compute_psnr: 10 log10(1 / mse) on [0,1] images."""
  mse = float(np.mean((np.asarray(image, np.float64) - np.asarray(target, np.float64)) ** 2))
  return 10.0 * np.log10(1.0 / max(mse, 1e-20))


class OracleScene:
  def __init__(self, gaussians, sh_degree, lrs, w_l1=0.0, w_mse=10.0, w_ssim=1.0, ssim_levels=3,
               reg_weights=(0.01, 1.0, 1e-4), split_alpha=0.01, prune_alpha=0.1, blur_cov=0.3, optim=None):
    """lrs: dict attribute -> learning rate (the same numbers SceneConfig holds)."""
    self.orc = go.Oracle("f32")
    self.deg = sh_degree
    K = (sh_degree + 1) ** 2
    self.g = {k: np.ascontiguousarray(v, np.float32).copy() for k, v in gaussians.items() if k in ATTRS}
    self.g["sh"] = np.ascontiguousarray(self.g["sh"][:, :, :K])
    self.lrs = dict(lrs)
    self.w_l1, self.w_mse, self.w_ssim, self.ssim_levels = w_l1, w_mse, w_ssim, ssim_levels
    self.reg_weights = tuple(reg_weights)
    self.split_alpha, self.prune_alpha = split_alpha, prune_alpha
    self.cfg = go.make_config(blur_cov=blur_cov)
    self.ocfg = go.make_optim_config(**(optim or {}))
    n = self.n = self.g["position"].shape[0]
    z = lambda *s: np.zeros(s, np.float32)
    self.grads = {k: np.zeros_like(v) for k, v in self.g.items()}
    self.m1 = {k: np.zeros_like(v) for k, v in self.g.items()}
    self.m2 = dict(position=z(n), log_scaling=z(n, 3), rotation=z(n), alpha_logit=z(n), sh=z(n))
    self.running, self.step_count = z(n), z(n)
    self.stats = dict(prune_cost=z(n), split_score=z(n), max_scale_px=z(n), points_in_view=np.zeros(n, np.int16),
                      visibility=z(n), scene_visible=z(n))
    self.last = {}

  def render(self, cam):
    o, g, cfg = self.orc, self.g, self.cfg
    W, H = cam.width, cam.height
    g2d, depth, rgb = o.project_fwd(cam, cfg, g["position"], g["log_scaling"], g["rotation"], g["alpha_logit"], g["sh"], self.deg)
    o2p, ranges, total, _ = o.tile_map(cfg, g2d, depth, W, H)
    image, fT, nc, _, _ = o.raster_fwd(cfg, g2d, rgb, depth, o2p, ranges, W, H, visibility=False)
    self.last = dict(g2d=g2d, depth=depth, rgb=rgb, o2p=o2p, ranges=ranges, total=total, image=image, final_T=fT, n_contrib=nc)
    return image

  def forward_backward(self, cam, target):
    """target: (H,W,3) uint8 or float32 in [0,1].  Gradients are ADDED into self.grads (batch = sum).
    Returns (Σ|d|, Σd²) like SplatScene.forward_backward."""
    o, g, cfg = self.orc, self.g, self.cfg
    W, H = cam.width, cam.height
    image = self.render(cam)
    L = self.last
    u8 = target.dtype == np.uint8
    tf = target.astype(np.float32) / np.float32(255.0) if u8 else np.asarray(target, np.float32)
    dimg, sums = o.loss_l1_mse(image, None if u8 else tf, self.w_l1, self.w_mse, target_u8=target if u8 else None)
    if self.w_ssim != 0.0:
      loss_ssim, dssim, ssim0 = o.ssim_loss(np.clip(image, 0, 1), tf, levels=self.ssim_levels, weight=self.w_ssim)
      dimg = dimg + np.where((image >= 0) & (image <= 1), dssim, 0).astype(np.float32)
      L["ssim_loss"], L["ssim0"] = loss_ssim, ssim0
    acc = o.raster_bwd(cfg, L["g2d"], L["rgb"], L["o2p"], L["ranges"], W, H, L["final_T"], L["n_contrib"], dimg)
    vis = np.ascontiguousarray(acc[:, 11])
    d_depth = None
    d_ls = None
    if any(w != 0.0 for w in self.reg_weights):
      d_op = np.ascontiguousarray(acc[:, 5]); d_depth = np.zeros_like(vis); d_ls = np.zeros((self.n, 3), np.float32)
      L["reg_sums"] = o.reg_loss(self.reg_weights, L["g2d"], L["depth"], g["log_scaling"], vis, d_op, d_depth, d_ls)
      acc[:, 5] = d_op
    grads = o.project_bwd(cam, cfg, g["position"], g["log_scaling"], g["rotation"], g["alpha_logit"], g["sh"], self.deg,
                          L["g2d"], acc[:, 0:6], acc[:, 6:9], d_depth, row_vis=vis)
    if d_ls is not None:
      grads["log_scaling"] = grads["log_scaling"] + d_ls
    for k in ATTRS:
      self.grads[k] += grads[k]
    o.accumulate_stats(None, L["g2d"], vis, np.ascontiguousarray(acc[:, 9:11]), self.split_alpha, self.prune_alpha, self.stats)
    L["acc"], L["d_image"] = acc, dimg
    return sums

  def step(self):
    o = self.orc
    vis = self.stats["scene_visible"]
    o.optim_vis_update(self.ocfg, vis, self.running, self.step_count)
    bls, bq = self.g["log_scaling"].copy(), self.g["rotation"].copy()
    for name in ATTRS:
      post = POST[name]
      o.optim_step(self.ocfg, KIND[name], self.lrs[name], self.g[name], self.grads[name], self.m1[name], self.m2[name], vis,
                   self.running, self.step_count, bls, bq, post[0], post[1], post[2], 1)
    vis[:] = 0

  def training_step(self, batch):
    """batch: list of (oracle camera, target).  One optimiser step over the batch (trainer/trainer.py:531-545)."""
    total = np.zeros(2, np.float64)
    for cam, target in batch:
      total += np.asarray(self.forward_backward(cam, target), np.float64)
    self.step()
    return total
