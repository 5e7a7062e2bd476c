"""SplatScene — the fused training hot path over packed splat parameters (host side).

Mirrors the reference scene interface for this path (reference scene/scene.py:17-83; the MLPScene
implementation scene/mlp_scene.py:214-313,411-427): render(), add_rendering (statistics), step(),
zero_grad(), split_and_prune(), num_points — with an SH colour model instead of the colour MLP.

HBM layout (one allocation each, attribute planes 256-B aligned):
  params / grads / exp_avg : [position 3N | log_scaling 3N | rotation 4N | alpha_logit N | sh 3K·N] (+ `visible` N at
  the tail of grads, so ONE all-reduce moves gradients and visibility under camera-batch data parallelism)
  exp_avg_sq               : [position N | log_scaling 3N | rotation N | alpha_logit N | sh N]
Per-image work buffers (g2d, depth, rgb, acc, image, final_T, n_contrib, d_image, tile-map workspace) are
allocated once and reused; no stage reads a count back to the host.
"""
import dataclasses
import math

import torch

from . import lib as L
from . import ops

ATTRS = ("position", "log_scaling", "rotation", "alpha_logit", "sh")


@dataclasses.dataclass
class SceneConfig:
  sh_degree: int = 3
  # per-group learning rates = the reference's config/scene/mlp.yaml:8-14 (position .3 local_vector, log_scaling .08,
  # rotation .01 vector, alpha_logit .1, feature 5.0 vector).  `sh` is this repo's SH colour group (the reference's
  # scene has MLP point features instead): its rate is ours.
  lr_position: float = 0.3       # local_vector: in units of the splat's own R·S basis
  lr_log_scaling: float = 0.08   # scalar
  lr_rotation: float = 0.01      # vector
  lr_alpha_logit: float = 0.1    # scalar
  lr_sh: float = 0.01            # vector
  optim: dict = dataclasses.field(default_factory=dict)  # overrides of make_optim_config (LaProp, .8/.95, ...)
  w_l1: float = 0.0              # reference trainer/config.py:67-70, config/trainer/default.yaml:52-55
  w_mse: float = 10.0
  w_ssim: float = 1.0            # ssim_weight / ssim_levels: reference config/trainer/default.yaml:52,55
  ssim_levels: int = 3
  split_alpha: float = 0.01      # reference controller/point_state.py:49-50
  prune_alpha: float = 0.1
  antialias: bool = False        # reference trainer/trainer.py:308-310: blur_cov = 0.0 if antialias else 0.3
  blur_cov: float = 0.3
  max_overlaps: int = 0          # capacity of the tile-overlap list; 0 = calibrate on first use
  capacity: int = 0              # rows the buffers are laid out for (0 = the initial count; grown ×1.5 when exceeded)
  # reg_weight (scale, opacity, aspect): reference config/scene/mlp.yaml:16-19 at t = 0 (the log_decay schedules are the
  # caller's: assign scene.config.reg_weights per step).  (0, 0, 0) disables the two regularisation sweeps.
  reg_weights: tuple = (0.01, 1.0, 1e-4)
  # colour model: False = SH coefficients per splat (SHScene); True = the reference's default MLPScene colour path
  # (per-splat feature 16 + per-image GLO embedding 32 -> ColorModel, scene/mlp_scene.py:352-368): the last attribute
  # plane then holds the point features (n, 16) and `sh_degree` is unused
  # rasteriser of the training path: "async" = staged records gathered by cp.async with the heaviest-first persistent
  # tile queue (csrc/raster_rec.cu); "gather" = the plain gather-staged kernels (csrc/raster.cu)
  raster: str = "async"
  color_mlp: bool = False
  num_images: int = 0            # rows of the GLO table (training images)
  lr_feature: float = 5.0        # vector group, config/scene/mlp.yaml:14
  lr_color: float = 1e-2         # ColorModel Adam (lr_diffuse / lr_specular at t = 0, config/scene/mlp.yaml:40-41)
  lr_glo: float = 2.0            # lr_glo_feature, config/scene/mlp.yaml:21
  reg_specular: float = 1e-5     # reg_weight.specular, config/scene/mlp.yaml:19


def camera_position(cam):
  """camera centre in world coordinates from the world->camera transform: -R^T t."""
  T = [cam.T[i] for i in range(16)]
  R = [[T[4 * r + c] for c in range(3)] for r in range(3)]
  t = [T[3], T[7], T[11]]
  return [-(R[0][k] * t[0] + R[1][k] * t[1] + R[2][k] * t[2]) for k in range(3)]


def _plane_offsets(n, dims):
  offs, o = [], 0
  for d in dims:
    offs.append(o)
    o += (n * d + 63) // 64 * 64  # 256-B aligned planes
  return offs, o


class SplatScene:
  def __init__(self, gaussians, config=None, device="cuda"):
    """gaussians: dict position (N,3), log_scaling (N,3), rotation (N,4) xyzw, alpha_logit (N,), sh (N,3,K) — or,
    with config.color_mlp, feature (N,16) instead of sh (kept in the plane named "sh")."""
    self.config = config or SceneConfig()
    self.device = torch.device(device)
    if self.device.type != "cuda":
      raise L.GsError("SplatScene needs a CUDA device (no CPU fallback)")
    L.load()
    self.K = (self.config.sh_degree + 1) ** 2
    self.ocfg = L.make_optim_config(**self.config.optim)
    self.color = None
    if self.config.color_mlp:
      from .color import ColorModule
      c = self.config
      self.color = ColorModule(c.num_images, self.device, lr=c.lr_color, lr_glo=c.lr_glo, specular_reg=c.reg_specular,
                               weights=gaussians.get("color_weights"))
    self._work = None
    self._alloc(int(gaussians["position"].shape[0]))
    for name in ATTRS:
      src = gaussians["feature" if (name == "sh" and self.color is not None) else name]
      if name == "sh" and self.color is None:
        src = src[:, :, :self.K]
      self.p[name].copy_(torch.as_tensor(src, dtype=torch.float32).reshape(self.p[name].shape))
    self._work = None
    self.timer = None  # optional StageTimer (bench.py): records a CUDA event after every stage
    self.raster_cfg = L.default_raster_config(antialias=int(self.config.antialias),
                                              blur_cov=0.0 if self.config.antialias else self.config.blur_cov,
                                              compute_visibility=1, compute_point_heuristic=1)

  # ---- storage ---------------------------------------------------------------------------------------
  def _dims(self):
    return (3, 3, 4, 1, 16 if self.color is not None else 3 * self.K)

  def _alloc(self, n, capacity=0):
    """One fp32 allocation per buffer kind, planes laid out for `capacity` rows (>= n) so the scene can grow to it
    without re-allocating: densification compacts into the second copy of each buffer (ping-pong) and swaps.
    capacity 0 = SceneConfig.capacity, else n."""
    dev = self.device
    cap = max(int(capacity or self.config.capacity or 0), n)
    self.n, self.capacity = n, cap
    dims = self._dims()
    offs, total = _plane_offsets(cap, dims)
    self._offs, self._total = offs, total
    sq_offs, sq_total = _plane_offsets(cap, (1, 3, 1, 1, 1))
    self._sq_offs = sq_offs
    self.params = torch.zeros((total,), device=dev)
    self.grads = torch.zeros((total + (cap + 63) // 64 * 64,), device=dev)  # + visible at the tail
    self.exp_avg = torch.zeros((total,), device=dev)
    self.exp_avg_sq_flat = torch.zeros((sq_total,), device=dev)
    self._spare = None  # (params, exp_avg, exp_avg_sq) compaction targets, allocated by the first densification
    self.running_vis = torch.zeros((cap,), device=dev)[:n]
    self.opt_step = torch.zeros((cap,), device=dev)[:n]
    self.stats = self.new_point_state(n)
    self._view(n)

  def _view(self, n):
    """row views [0, n) of the capacity-sized planes"""
    self.n = n
    dims = self._dims()
    shapes = dict(position=(n, 3), log_scaling=(n, 3), rotation=(n, 4), alpha_logit=(n,),
                  sh=(n, 16) if self.color is not None else (n, 3, self.K))

    def views(flat, offsets, dims_, shp):
      return {name: flat[o:o + n * d].view(shp[name]) for name, o, d in zip(ATTRS, offsets, dims_)}

    self.p = views(self.params, self._offs, dims, shapes)
    self.g = views(self.grads, self._offs, dims, shapes)
    self.m1 = views(self.exp_avg, self._offs, dims, shapes)
    sq_shapes = dict(position=(n,), log_scaling=(n, 3), rotation=(n,), alpha_logit=(n,), sh=(n,))
    self.m2 = views(self.exp_avg_sq_flat, self._sq_offs, (1, 3, 1, 1, 1), sq_shapes)
    self.visible = self.grads[self._total:self._total + n]
    self._vis_idx = None

  def new_point_state(self, n):
    """PointState.new_zeros (reference controller/point_state.py:19-32)."""
    dev = self.device
    return dict(prune_cost=torch.zeros((n,), device=dev), split_score=torch.zeros((n,), device=dev),
                max_scale_px=torch.zeros((n,), device=dev), points_in_view=torch.zeros((n,), device=dev, dtype=torch.int16),
                visibility=torch.zeros((n,), device=dev))

  @property
  def num_points(self):
    return self.n

  def gaussians(self):
    """copies of the parameter planes; the last one is named `feature` for an MLP colour scene (gsb200.io with_sh=False)."""
    out = {k: v.clone() for k, v in self.p.items()}
    if self.color is not None:
      out["feature"] = out.pop("sh")
    return out

  # ---- per-image work buffers --------------------------------------------------------------------------
  def _buffers(self, cam):
    """per-image work buffers: allocated once for (image size, row capacity); a densification only re-slices them"""
    W, H = cam.width, cam.height
    w = self._work
    if w is not None and w["W"] == W and w["H"] == H and w["n"] == self.n:
      return w
    dev, n, cap = self.device, self.n, self.capacity
    stride = ops.acc_stride(3)
    base = w["base"] if (w is not None and w["W"] == W and w["H"] == H and w["base"]["cap"] == cap) else None
    if base is None:
      levels = max(self.config.ssim_levels, 1)
      base = dict(cap=cap, g2d=torch.empty((cap, 6), device=dev), depth=torch.empty((cap,), device=dev),
                  rgb=torch.empty((cap, 3), device=dev), acc=torch.zeros((cap, stride), device=dev),
                  vis=torch.zeros((cap,), device=dev),
                  tinfo=torch.empty((L.load().gs_tile_info_bytes(cap, W, H),), device=dev, dtype=torch.uint8),
                  image=torch.empty((H, W, 3), device=dev), final_T=torch.empty((H, W), device=dev),
                  n_contrib=torch.empty((H, W), device=dev, dtype=torch.int32), d_image=torch.empty((H, W, 3), device=dev),
                  loss_sums=torch.zeros((2,), device=dev), reg_sums=torch.zeros((4,), device=dev),
                  ssim_sums=torch.zeros((levels,), device=dev),
                  ssim_ws=ops.ssim_workspace(H, W, 3, levels, dev) if self.config.w_ssim != 0.0 else None,
                  aux=ops.RasterAux(cap, W, H, dev) if self.config.raster == "async" else None,
                  tile=w["tile"] if (w is not None and w["W"] == W and w["H"] == H and w["tile"] is not None
                                     and w["tile"].m >= n) else None)
    w = dict(W=W, H=H, n=n, base=base, g2d=base["g2d"][:n], depth=base["depth"][:n], rgb=base["rgb"][:n], acc=base["acc"][:n],
             vis=base["vis"][:n], tinfo=ops.TileInfo(n, W, H, dev, buffer=base["tinfo"]), tile=base["tile"], aux=base["aux"])
    for k in ("image", "final_T", "n_contrib", "d_image", "loss_sums", "reg_sums", "ssim_sums", "ssim_ws"):
      w[k] = base[k]
    self._work = w
    return w

  def _tile_buffers(self, w, cam):
    """Overlap-list capacity: config.max_overlaps, else calibrated once from this camera's exact count × 1.5
    (one host read at start-up; the steady-state path never reads a count back)."""
    if w["tile"] is None:
      cap = self.config.max_overlaps
      if cap <= 0:
        cap = int(int(w["tinfo"].counts.sum(dtype=torch.int64).item()) * 1.5) + 65536
      w["tile"] = w["base"]["tile"] = ops.TileMapBuffers(self.capacity, cap, cam.width, cam.height, self.device)
    return w["tile"]

  def overflowed(self):
    """True if the last tile map exceeded the overlap capacity (host sync; call off the hot path)."""
    w = self._work
    return bool(w and w["tile"] is not None and int(w["tile"].counters[1].item()) != 0)

  def grow_overlaps(self, factor=2.0):
    w = self._work
    if w and w["tile"] is not None:
      cap = int(w["tile"].max_overlaps * factor)
      w["tile"] = w["base"]["tile"] = ops.TileMapBuffers(self.capacity, cap, w["W"], w["H"], self.device)

  def _mark(self, name):
    if self.timer is not None:
      self.timer.mark(name)

  # ---- forward ---------------------------------------------------------------------------------------
  def _forward(self, cam, median=False, visibility=None, image_idx=None):
    w = self._buffers(cam)
    p = self.p
    self._mark("begin")
    ops.project_fwd(cam, self.raster_cfg, p["position"], p["log_scaling"], p["rotation"], p["alpha_logit"],
                    None if self.color is not None else p["sh"], self.config.sh_degree,
                    out=(w["g2d"], w["depth"], w["rgb"]), tile_info=w["tinfo"],
                    records=w["aux"].records if (w["aux"] is not None and self.color is None) else None)
    self._mark("project_fwd")
    if self.color is not None:  # MLPScene.eval_colors for the in-view rows (reference scene/mlp_scene.py:415-417)
      self.color.forward(p["sh"], p["position"], camera_position(cam), image_idx, w["g2d"][:, 5], w["rgb"])
      if w["aux"] is not None:
        ops.raster_records(self.raster_cfg, w["g2d"], w["rgb"], w["aux"].records)
      self._mark("color_fwd")
    tb = self._tile_buffers(w, cam)
    ops.tile_map(self.raster_cfg, w["g2d"], w["depth"], cam.width, cam.height, buffers=tb, tile_info=w["tinfo"])
    self._mark("tile_map")
    _, _, _, vis, med = ops.raster_fwd(self.raster_cfg, w["g2d"], w["rgb"], w["depth"], tb.overlap_to_point,
                                       tb.tile_ranges, cam.width, cam.height, visibility=visibility, median=median,
                                       out=(w["image"], w["final_T"], w["n_contrib"]), aux=w["aux"])
    self._mark("raster_fwd")
    return w, tb, vis, med

  def render(self, cam, median_depth=False, compute_visibility=False, image_idx=None):
    """Forward only (reference scene/mlp_scene.py:411-427): returns dict(image (H,W,3) clamped to [0,1] like the
    reference's post_activation scene/color_model.py:154-160, raw, weight, visibility, median_depth)."""
    w, tb, vis, med = self._forward(cam, median=median_depth, visibility=True if compute_visibility else None,
                                    image_idx=image_idx)
    return dict(image=w["image"].clamp(0, 1), raw=w["image"], weight=1.0 - w["final_T"], visibility=vis,
                median_depth=med, g2d=w["g2d"], depth=w["depth"])

  def query_visibility(self, cam):
    """MLPScene.query_visibility (reference scene/mlp_scene.py:372-381): forward-only render for (idx, visibility) of
    the splats a pixel blends.  Returns (idx int64 ascending, visibility) — one host sync for the count, like the
    reference's nonzero()."""
    _, _, vis, _ = self._forward(cam, visibility=True)
    idx = (vis > 0).nonzero().squeeze(1)
    return idx, vis[idx]

  def evaluate(self, cam, target, image_idx=None):
    """Evaluation.metrics (reference trainer/evaluation.py:29-46, util/colors.py compute_psnr): psnr / l1 / ssim of the
    clamped render against `target` ((H,W,3) uint8 or float32 on the device), as one (3,) device tensor — forward only,
    the loss kernels' sums, no gradient."""
    w, _, _, _ = self._forward(cam, image_idx=image_idx)
    H, W = w["H"], w["W"]
    _, sums = ops.loss_l1_mse(w["image"], target, 0.0, 0.0, d_pred=w["d_image"])
    ss = ops.ssim_loss(w["image"], target, levels=1, weight=0.0, workspace=w["ssim_ws"])
    npx = float(H * W * 3)
    mse = sums[1] / npx
    return torch.stack([10.0 * torch.log10(1.0 / mse.clamp_min(1e-20)), sums[0] / npx,
                        ss[0] / float(max(ops.ssim_counts(H, W, 3, 1)[0], 1))])

  def ssim_value(self):
    """(weighted multi-scale SSIM loss, SSIM of level 0) of the last forward_backward image — device scalars."""
    w, c = self._work, self.config
    counts = torch.tensor(ops.ssim_counts(w["H"], w["W"], 3, c.ssim_levels), device=self.device, dtype=torch.float32)
    ssim = w["ssim_sums"] / counts.clamp_min(1.0)
    return c.w_ssim * (1.0 - ssim).mean(), ssim[0]

  def reg_value(self):
    """Σ weight_k · reg_k of the last forward_backward image (device scalar; reference reg_loss's return value)."""
    s, wts = self._work["reg_sums"], self.config.reg_weights
    return (wts[0] * s[0] + wts[1] * s[1] + wts[2] * s[2]) / s[3].clamp_min(1.0)

  # ---- MCMC controller (reference controller/mcmc_controller.py:69-100) ---------------------------------------------
  stats_partial = None  # set by the Trainer under data parallelism while each rank holds only a slice of the statistics

  state_partial = None  # likewise for the optimiser state under the row-sharded optimiser

  def _require_full_stats(self):
    if self.stats_partial:
      raise L.GsError(f"scene statistics are incomplete on this rank: {self.stats_partial}")
    if self.state_partial:
      raise L.GsError(f"optimiser state is incomplete on this rank: {self.state_partial}")

  def mcmc_noise(self, noise_level, min_views=5, opacity_threshold=0.1, randn=None):
    """noise branch (:91-100).  randn: (N,3) N(0,1) draws (None: torch's CUDA generator; seed it identically on every
    rank).  The reference writes through a boolean-masked TensorDict copy; the intended effect — the noise is added to
    the positions — is what is implemented."""
    self._require_full_stats()
    if randn is None:
      randn = torch.randn((self.n, 3), device=self.device)
    ops.mcmc_noise(self.p["alpha_logit"], self.p["log_scaling"], self.p["rotation"], self.stats["points_in_view"], min_views,
                   opacity_threshold, noise_level, randn, self.p["position"])

  def mcmc_find_prune_split(self, opacity_threshold=0.1, max_scale_px=200.0, min_split_px=0.0):
    """prune / respawn branch (:76-89): prune large or transparent splats and split as many by top split_score."""
    self._require_full_stats()
    st = self.stats
    opacity = torch.sigmoid(self.p["alpha_logit"])
    prune_mask = (st["max_scale_px"] > max_scale_px) | (opacity < opacity_threshold)
    n = int(prune_mask.sum().item())
    score = torch.where(prune_mask | (st["max_scale_px"] < min_split_px), torch.zeros_like(st["split_score"]), st["split_score"])
    split_mask = ops.select_topk_mask(score, n, descending=1)
    return split_mask, prune_mask

  # ---- one training image: forward + loss + backward + statistics ------------------------------------------
  def forward_backward(self, cam, target, first=False, exchange=None, slot=0, target_ready=None, image_idx=None):
    """Reference trainer/trainer.py:500-514 for one image (render → loss → backward → add_rendering), fused:
    gradients are ADDED into self.g (batch = sum), statistics into self.stats / self.visible.
    target: (H,W,3) float32 in [0,1] or uint8 on the device.  Returns loss_sums (device, 2): Σ|d|, Σd².
    first=True (the step's first image, gradient buffer all-zero: the optimiser clears the rows it consumes): rows are
    WRITTEN instead of read-modified-written.
    exchange (an ImageExchange, camera-batch DP): the image's gradient rows and statistics are packed into record
    `slot` instead of being applied here; ImageExchange.fold() applies every rank's records in image order.
    target_ready: CUDA event of a side-stream upload of `target`; the stream waits for it right before the loss."""
    w, tb, _, _ = self._forward(cam, image_idx=image_idx)
    c = self.config
    if target_ready is not None:
      torch.cuda.current_stream().wait_event(target_ready)
    if c.w_ssim != 0.0:  # L1/MSE + multi-scale SSIM (reference trainer/trainer.py:450-498), one fused call
      ops.photometric_loss(w["image"], target, c.w_l1, c.w_mse, c.w_ssim, c.ssim_levels, w["d_image"], w["loss_sums"],
                           w["ssim_sums"], w["ssim_ws"])
    else:
      w["loss_sums"].zero_()
      ops.loss_l1_mse(w["image"], target, c.w_l1, c.w_mse, d_pred=w["d_image"], sums=w["loss_sums"])
    self._mark("loss")
    acc = w["acc"]
    ops.raster_bwd(self.raster_cfg, w["g2d"], w["rgb"], tb.overlap_to_point, tb.tile_ranges, cam.width, cam.height,
                   w["final_T"], w["n_contrib"], w["d_image"], acc=acc, vis_plane=w["vis"], aux=w["aux"])
    self._mark("raster_bwd")
    p = self.p
    reg = any(x != 0.0 for x in c.reg_weights)
    if reg:
      # reg_loss (reference scene/mlp_scene.py:246-288) over the rows this image blended.  One sweep for the sums (the
      # gradients need the blended-row count); the gradients w.r.t. the projected opacity / depth / log_scaling are
      # formed per row inside the projection backward, which already holds everything they need
      ops.reg_loss(c.reg_weights, w["g2d"], w["depth"], p["log_scaling"], w["vis"], sums=w["reg_sums"], compute_sums=True)
      self._mark("reg")
    accumulate = not (first or exchange is not None)
    ops.project_bwd(cam, self.raster_cfg, p["position"], p["log_scaling"], p["rotation"], p["alpha_logit"],
                    None if self.color is not None else p["sh"], c.sh_degree, w["g2d"], acc[:, 0:6],
                    None if self.color is not None else acc[:, 6:9], None, row_vis=w["vis"], grads=self.g,
                    accumulate=accumulate, reg=(c.reg_weights, w["reg_sums"], w["depth"]) if reg else None)
    if self.color is not None:  # colour gradient of the blended rows -> point features, GLO row, MLP weights
      self._mark("project_bwd")
      self.color.backward(p["sh"], p["position"], camera_position(cam), image_idx, w["vis"], acc[:, 6:9], self.g["sh"],
                          accumulate, reg_sums=w["reg_sums"] if reg else None)
      self._mark("color_bwd")
    else:
      self._mark("project_bwd")
    if exchange is not None:
      # the dense gradient buffer holds this image's rows only (pack moves them out and zeroes them again)
      exchange.pack(slot, w, acc)
      self._mark("pack")
      return w["loss_sums"]
    # the statistics pass is the accumulator's last reader: it clears the rows this image touched, so `acc` / `vis`
    # are all-zero again for the next image without a 48·N-byte memset
    ops.accumulate_stats(None, w["g2d"], w["vis"], acc[:, 9:11], c.split_alpha, c.prune_alpha, self.stats,
                         scene_visible=self.visible, clear_acc=acc, clear_vis=w["vis"])
    self._mark("stats")
    return w["loss_sums"]

  # ---- optimiser step (reference scene/mlp_scene.py:214-239) -----------------------------------------------------
  def step(self, rows=None):
    """rows = (begin, end): step only that row slice (row-sharded optimiser under DP, begin a multiple of 128; the
    other rows' state on this rank is then stale until Trainer.sync_statistics gathers it)."""
    c, o = self.config, self.ocfg
    vis, rv, st = self.visible, self.running_vis, self.opt_step
    p, g, m1, m2 = self.p, self.g, self.m1, self.m2
    n = self.n
    if rows is not None and rows != (0, self.n):
      b, e = rows
      n = e - b
      vis, rv, st = vis[b:e], rv[b:e], st[b:e]
      p, g, m1, m2 = ({k: v[b:e] for k, v in d.items()} for d in (p, g, m1, m2))
    self._mark("optim_begin")
    if self._vis_idx is None or self._vis_idx.n != n:
      self._vis_idx = ops.VisibleIndexes(n, self.device)
    # visibility state + device-side visible.nonzero() (reference mlp_scene.py:217) in one kernel, no host sync
    ix = self._vis_idx.begin(o, vis, rv, st)
    # two launches instead of five: position's local basis is built from log_scaling / rotation BEFORE their update
    # (mlp_scene.py:219), so those two groups go in the second launch; groups within a launch are independent
    G = ops.optim_group
    ops.optim_step_groups(o, n, [
      G(L.GS_GROUP_LOCAL_VECTOR, c.lr_position, p["position"], g["position"], m1["position"], m2["position"],
        basis_ls=p["log_scaling"], basis_q=p["rotation"]),
      G(L.GS_GROUP_SCALAR, c.lr_alpha_logit, p["alpha_logit"], g["alpha_logit"], m1["alpha_logit"], m2["alpha_logit"]),
      G(L.GS_GROUP_VECTOR, c.lr_feature if self.color is not None else c.lr_sh, p["sh"], g["sh"], m1["sh"], m2["sh"])],
      vis, rv, st, indexes=ix)
    ops.optim_step_groups(o, n, [
      G(L.GS_GROUP_SCALAR, c.lr_log_scaling, p["log_scaling"], g["log_scaling"], m1["log_scaling"], m2["log_scaling"],
        post=L.GS_POST_CLAMP, post_lo=-8.0, post_hi=8.0),
      G(L.GS_GROUP_VECTOR, c.lr_rotation, p["rotation"], g["rotation"], m1["rotation"], m2["rotation"],
        post=L.GS_POST_NORMALIZE)], vis, rv, st, indexes=ix)
    if self.color is not None:
      self.color.step()
    self._mark("optim")
    vis.zero_()  # zero_grad(): gradient rows were cleared by the optimiser kernels (mlp_scene.py:155-161)

  def zero_grad(self):
    self.grads.zero_()

  # ---- densification (reference scene/mlp_scene.py:301-313, gaussians/split.py:87-113) ---------------------------
  def split_and_prune(self, keep_mask, split_idx=None, uniform=None, sep=0.7):
    """points = points[keep_mask] (parameters AND optimiser state) then append 2 children per split_idx row with
    zero optimiser state.  keep_mask: (N,) bool; split_idx: (S,) int64 rows of the OLD numbering.
    uniform: (S,) U[0,1) draws for the split axis (random_axis=True); None draws them from torch's CUDA generator.
    The kept rows are compacted into the spare copy of each buffer (allocated once, at capacity) and the copies are
    swapped: no per-densification allocation of the 59·4·N-byte buffers (the allocator churn of the reference,
    controller/target_controller.py:125-126).  The number of kept rows is read back once (densification is off the
    per-iteration path; the reference reads four scalars here, controller/point_state.py:83-91)."""
    old_n = self.n
    s = 0 if split_idx is None else int(split_idx.shape[0])
    slots, count = ops.compact_slots(keep_mask)
    kept = int(count.item())
    new_n = kept + 2 * s
    old = dict(p=self.p, m1=self.m1, m2=self.m2, running_vis=self.running_vis, opt_step=self.opt_step, stats=self.stats)
    if new_n > self.capacity:  # grow: fresh buffers laid out for 1.5x the rows (the old ones stay alive through `old`)
      self._alloc(new_n, capacity=max(int(1.5 * new_n), self.capacity))
    else:
      if self._spare is None:
        self._spare = (torch.zeros_like(self.params), torch.zeros_like(self.exp_avg), torch.zeros_like(self.exp_avg_sq_flat))
      spare = self._spare
      self._spare = (self.params, self.exp_avg, self.exp_avg_sq_flat)
      self.params, self.exp_avg, self.exp_avg_sq_flat = spare
      cap = self.capacity
      self.running_vis = torch.zeros((cap,), device=self.device)[:new_n]
      self.opt_step = torch.zeros((cap,), device=self.device)[:new_n]
      self.stats = self.new_point_state(new_n)
      self._view(new_n)
      if s > 0:  # children start with zero optimiser state: clear their rows of the recycled moment buffers
        for name in ATTRS:
          self.m1[name][kept:].zero_()
          self.m2[name][kept:].zero_()
    for name in ATTRS:
      ops.compact_rows(old["p"][name], slots, self.p[name])
      ops.compact_rows(old["m1"][name], slots, self.m1[name])
      ops.compact_rows(old["m2"][name], slots, self.m2[name])
    ops.compact_rows(old["running_vis"], slots, self.running_vis)
    ops.compact_rows(old["opt_step"], slots, self.opt_step)
    for k, v in old["stats"].items():
      ops.compact_rows(v, slots, self.stats[k])
    if s > 0:
      if uniform is None:
        uniform = torch.rand((s,), device=self.device)
      op = old["p"]
      ops.split_positions(split_idx, op["position"], op["log_scaling"], op["rotation"], uniform, sep,
                          self.p["position"], self.p["log_scaling"], out_offset=kept)
      for name in ("rotation", "alpha_logit", "sh"):
        ops.gather_rows(op[name], self.p[name], split_idx, repeat=2, out_offset=kept)
    return old_n, new_n

  def find_split_prune(self, t, target_points, prune_rate=0.025, min_views=5, max_scale_px=200.0, min_split_px=0.0):
    """TargetController.find_split_prune_indexes (reference controller/target_controller.py:73-96) with the two
    full argsorts replaced by radix selects.  Returns (split_mask, prune_mask)."""
    self._require_full_stats()
    st = self.stats
    n = self.n
    exceeds = st["max_scale_px"] > max_scale_px
    n_prune = int(math.ceil(prune_rate * n * (1 - t)))
    seen = st["points_in_view"] >= min_views
    prune_cost = torch.where(seen, st["prune_cost"], torch.full_like(st["prune_cost"], float("inf")))
    prune_mask = ops.select_topk_mask(prune_cost, n_prune, descending=0) | exceeds
    target_split = (target_points - n) + int(prune_mask.sum().item())
    split_score = st["split_score"]
    split_score[prune_mask] = 0.0
    if min_split_px > 0:
      split_score[st["max_scale_px"] < min_split_px] = 0.0
    split_mask = ops.select_topk_mask(split_score, max(target_split, 0), descending=1)
    both = split_mask & prune_mask
    return split_mask ^ both, prune_mask ^ both

  def densify(self, t, target_points, uniform=None, uniform_rows=None, **kw):
    """densify_and_prune (reference controller/point_state.py:76-110) + new PointState (target_controller.py:122).
    uniform: (S,) split-axis draws in split order; uniform_rows: (N,) one draw per OLD row (the split rows take theirs —
    what replicated ranks use, so identical statistics give identical children)."""
    split_mask, prune_mask = self.find_split_prune(t, target_points, **kw)
    split_idx = split_mask.nonzero().squeeze(1)
    if uniform is None and uniform_rows is not None:
      uniform = uniform_rows[split_idx].contiguous()
    keep = ~(split_mask | prune_mask)
    out = self.split_and_prune(keep, split_idx, uniform=uniform)
    self.stats = self.new_point_state(self.n)
    return out

  # ---- checkpoint: the reference's nested layout (trainer/trainer.py:188-194, scene/mlp_scene.py:316-322) -----------
  _CKPT = dict(position="position", log_scaling="log_scaling", rotation="rotation", alpha_logit="alpha_logit", sh="feature")

  def learning_rates(self):
    c = self.config
    return dict(position=c.lr_position, log_scaling=c.lr_log_scaling, rotation=c.lr_rotation, alpha_logit=c.lr_alpha_logit,
                feature=c.lr_feature if self.color is not None else c.lr_sh)

  def state_dict(self):
    """dict(points=dict(tensors, parameter_groups, state, shared)[, color=...]) — the layout of the reference's
    MLPScene.state_dict (`points` = ParameterClass.state_dict()), plain tensors / scalars only (reference checkpoints
    load with weights_only=True, trainer/checkpoint.py:38).  gsb200.plugin.SHScene reads and writes the same layout."""
    n = self.n
    kinds = dict(position="local_vector", log_scaling="scalar", rotation="vector", alpha_logit="scalar", feature="vector")
    lrs = self.learning_rates()
    tensors = {self._CKPT[k]: v.clone() for k, v in self.p.items()}
    tensors["alpha_logit"] = tensors["alpha_logit"].reshape(n, 1)
    tensors["visible"] = self.visible.clone()
    state = {self._CKPT[k]: dict(exp_avg=self.m1[k].clone().reshape(tensors[self._CKPT[k]].shape), exp_avg_sq=self.m2[k].clone())
             for k in ATTRS}
    o = self.ocfg
    points = dict(tensors=tensors, parameter_groups={k: dict(lr=float(lrs[k]), type=kinds[k]) for k in kinds},
                  state=state, shared=dict(step=self.opt_step.clone(), running_vis=self.running_vis.clone()),
                  optim=dict(mode=int(o.mode), beta1=float(o.beta1), beta2=float(o.beta2), eps=float(o.eps),
                             vis_beta=float(o.vis_beta), vis_smooth=float(o.vis_smooth), grad_clip=float(o.grad_clip),
                             bias_correction=int(o.bias_correction)))
    c = self.config
    sd = dict(points=points, sh_degree=int(c.sh_degree), reg_weights=[float(x) for x in c.reg_weights],
              loss=dict(w_l1=float(c.w_l1), w_mse=float(c.w_mse), w_ssim=float(c.w_ssim), ssim_levels=int(c.ssim_levels)))
    if self.color is not None:  # colour model + GLO table + their optimiser state (reference all_parameters, mlp_scene.py:291-298)
      sd["color"] = self.color.state_dict()
    return sd

  @staticmethod
  def from_state_dict(sd, config=None, device="cuda"):
    pts = sd["points"]
    t = pts["tensors"]
    mlp = "color" in sd
    if config is None:
      lr = {k: float(g["lr"]) for k, g in pts["parameter_groups"].items()}
      loss = sd.get("loss", {})
      config = SceneConfig(sh_degree=int(sd.get("sh_degree", round(t["feature"].shape[-1] ** 0.5) - 1 if not mlp else 3)),
                           color_mlp=mlp, num_images=int(sd["color"]["glo"].shape[0]) if mlp else 0,
                           lr_position=lr["position"], lr_log_scaling=lr["log_scaling"], lr_rotation=lr["rotation"],
                           lr_alpha_logit=lr["alpha_logit"], optim=dict(pts.get("optim", {})),
                           reg_weights=tuple(sd.get("reg_weights", (0.01, 1.0, 1e-4))), **loss)
      if mlp:
        config.lr_feature = lr["feature"]
      else:
        config.lr_sh = lr["feature"]
    g = dict(position=t["position"], log_scaling=t["log_scaling"], rotation=t["rotation"],
             alpha_logit=t["alpha_logit"].reshape(-1))
    g["feature" if mlp else "sh"] = t["feature"]
    scene = SplatScene(g, config, device)
    if mlp:
      scene.color.load_state_dict(sd["color"])
    for k in ATTRS:
      st = pts.get("state", {}).get(SplatScene._CKPT[k])
      if st:
        scene.m1[k].copy_(st["exp_avg"].reshape(scene.m1[k].shape)); scene.m2[k].copy_(st["exp_avg_sq"].reshape(scene.m2[k].shape))
    shared = pts.get("shared", {})
    if "step" in shared:
      scene.opt_step.copy_(shared["step"]); scene.running_vis.copy_(shared["running_vis"])
    if "visible" in t:
      scene.visible.copy_(t["visible"])
    return scene
