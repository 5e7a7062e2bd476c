"""ctypes binding of the CPU oracle (oracle/gs_oracle.c).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs.  The product package never imports this module.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))


class Camera(C.Structure):  # layout == gs_camera (include/gsb200.h)
  _fields_ = [("T", C.c_float * 16), ("fx", C.c_float), ("fy", C.c_float), ("cx", C.c_float), ("cy", C.c_float),
              ("width", C.c_int32), ("height", C.c_int32), ("near_plane", C.c_float), ("far_plane", C.c_float)]


class RasterConfig(C.Structure):  # layout == gs_raster_config
  _fields_ = [("tile_size", C.c_int32), ("alpha_threshold", C.c_float), ("clamp_max_alpha", C.c_float),
              ("saturate_threshold", C.c_float), ("blur_cov", C.c_float), ("antialias", C.c_int32),
              ("compute_visibility", C.c_int32), ("compute_point_heuristic", C.c_int32),
              ("render_median_depth", C.c_int32), ("frustum_clamp", C.c_float)]


class OptimConfig(C.Structure):  # layout == gs_optim_config
  _fields_ = [("mode", C.c_int32), ("beta1", C.c_float), ("beta2", C.c_float), ("eps", C.c_float),
              ("vis_beta", C.c_float), ("vis_smooth", C.c_float), ("grad_clip", C.c_float),
              ("bias_correction", C.c_int32)]


def make_config(**kw):
  cfg = RasterConfig(tile_size=16, alpha_threshold=1.0 / 255.0, clamp_max_alpha=0.99, saturate_threshold=0.9999,
                     blur_cov=0.3, antialias=0, compute_visibility=1, compute_point_heuristic=1,
                     render_median_depth=0, frustum_clamp=1.3)
  for k, v in kw.items():
    setattr(cfg, k, v)
  return cfg


def make_camera(T, fx, fy, cx, cy, width, height, near=0.1, far=100.0):
  cam = Camera()
  T = np.asarray(T, dtype=np.float32).reshape(16)
  for i in range(16):
    cam.T[i] = float(T[i])
  cam.fx, cam.fy, cam.cx, cam.cy = float(fx), float(fy), float(cx), float(cy)
  cam.width, cam.height = int(width), int(height)
  cam.near_plane, cam.far_plane = float(near), float(far)
  return cam


def make_optim_config(mode=1, beta1=0.8, beta2=0.95, eps=1e-8, vis_beta=0.999, vis_smooth=0.01, grad_clip=2.0,
                      bias_correction=1):
  return OptimConfig(mode=mode, beta1=beta1, beta2=beta2, eps=eps, vis_beta=vis_beta, vis_smooth=vis_smooth,
                     grad_clip=grad_clip, bias_correction=bias_correction)


def build(force=False):
  names = ["libgs_oracle_f32.so", "libgs_oracle_f64.so"]
  src = os.path.join(_HERE, "gs_oracle.c")
  stale = force or any(not os.path.exists(os.path.join(_HERE, n)) or
                       os.path.getmtime(os.path.join(_HERE, n)) < os.path.getmtime(src) for n in names)
  if stale:
    subprocess.run(["make", "-C", _HERE, "-B"], check=True, capture_output=True)


def acc_stride(channels):
  return (9 + channels + 3) // 4 * 4


class Oracle:
  """precision: 'f32' (the oracle) or 'f64' (validation of the oracle itself)."""

  def __init__(self, precision="f32"):
    build()
    self.lib = C.CDLL(os.path.join(_HERE, f"libgs_oracle_{precision}.so"))
    self.real = np.float32 if precision == "f32" else np.float64
    assert self.lib.gso_real_bytes() == np.dtype(self.real).itemsize
    self.lib.gso_tile_map.restype = C.c_int64
    self.lib.gso_compact_slots.restype = C.c_int64
    self.lib.gso_exp_lerp.restype = C.c_float if precision == "f32" else C.c_double
    self.lib.gso_exp_lerp.argtypes = [self.lib.gso_exp_lerp.restype] * 3

  def set_threads(self, n):
    self.lib.gso_set_threads(C.c_int(int(n)))

  def get_threads(self):
    return int(self.lib.gso_get_threads())

  # -- helpers -------------------------------------------------------------------------------
  def r(self, a):
    return np.ascontiguousarray(a, dtype=self.real)

  @staticmethod
  def p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)

  def creal(self, x):
    return (C.c_float if self.real == np.float32 else C.c_double)(x)

  # -- (1) projection --------------------------------------------------------------------------
  def project_fwd(self, cam, cfg, position, log_scaling, rotation, alpha_logit, sh=None, sh_degree=0):
    n = position.shape[0]
    position, log_scaling, rotation, alpha_logit = map(self.r, (position, log_scaling, rotation, alpha_logit))
    sh = None if sh is None else self.r(sh)
    g2d = np.zeros((n, 6), self.real); depth = np.zeros(n, self.real); rgb = np.zeros((n, 3), self.real)
    self.lib.gso_project_fwd(C.byref(cam), C.byref(cfg), C.c_int64(n), self.p(position), self.p(log_scaling),
                             self.p(rotation), self.p(alpha_logit), self.p(sh), C.c_int(sh_degree),
                             self.p(g2d), self.p(depth), self.p(rgb))
    return g2d, depth, rgb

  def project_bwd(self, cam, cfg, position, log_scaling, rotation, alpha_logit, sh, sh_degree, g2d,
                  d_g2d, d_rgb, d_depth=None, grads=None, row_vis=None):
    n = position.shape[0]
    position, log_scaling, rotation, alpha_logit, g2d = map(self.r, (position, log_scaling, rotation, alpha_logit, g2d))
    sh = None if sh is None else self.r(sh)
    d_g2d = self.r(d_g2d); d_rgb = None if d_rgb is None else self.r(d_rgb)
    d_depth = None if d_depth is None else self.r(d_depth)
    row_vis = None if row_vis is None else self.r(row_vis)
    accumulate = grads is not None
    if grads is None:
      grads = dict(position=np.zeros((n, 3), self.real), log_scaling=np.zeros((n, 3), self.real),
                   rotation=np.zeros((n, 4), self.real), alpha_logit=np.zeros(n, self.real),
                   sh=None if sh is None else np.zeros_like(sh))
    self.lib.gso_project_bwd(C.byref(cam), C.byref(cfg), C.c_int64(n), self.p(position), self.p(log_scaling),
                             self.p(rotation), self.p(alpha_logit), self.p(sh), C.c_int(sh_degree), self.p(g2d),
                             self.p(d_g2d), C.c_int(d_g2d.shape[1]), self.p(d_rgb),
                             C.c_int(0 if d_rgb is None else d_rgb.shape[1]), self.p(d_depth),
                             self.p(row_vis), C.c_int(1), self.p(grads["position"]), self.p(grads["log_scaling"]), self.p(grads["rotation"]),
                             self.p(grads["alpha_logit"]), self.p(grads["sh"]), C.c_int(int(accumulate)))
    return grads

  def sh_eval_fwd(self, indexes, sh, sh_degree, position, cam_pos):
    sh, position, cam_pos = map(self.r, (sh, position, cam_pos))
    idx = None if indexes is None else np.ascontiguousarray(indexes, np.int64)
    v = position.shape[0] if idx is None else idx.shape[0]
    rgb = np.zeros((v, 3), self.real)
    self.lib.gso_sh_eval_fwd(C.c_int64(v), self.p(idx), self.p(sh), C.c_int(sh_degree), self.p(position),
                             self.p(cam_pos), self.p(rgb))
    return rgb

  def sh_eval_bwd(self, indexes, sh, sh_degree, position, cam_pos, d_rgb):
    sh, position, cam_pos, d_rgb = map(self.r, (sh, position, cam_pos, d_rgb))
    idx = None if indexes is None else np.ascontiguousarray(indexes, np.int64)
    v = position.shape[0] if idx is None else idx.shape[0]
    g_sh = np.zeros_like(sh); g_pos = np.zeros_like(position)
    self.lib.gso_sh_eval_bwd(C.c_int64(v), self.p(idx), self.p(sh), C.c_int(sh_degree), self.p(position),
                             self.p(cam_pos), self.p(d_rgb), self.p(g_sh), self.p(g_pos))
    return g_sh, g_pos

  def sh_basis(self, deg, dirs):
    dirs = self.r(dirs)
    out = np.zeros((dirs.shape[0], 16), self.real)
    self.lib.gso_sh_basis(C.c_int(deg), C.c_int64(dirs.shape[0]), self.p(dirs), self.p(out))
    return out

  # -- (2) tile mapping (always float32) -------------------------------------------------------
  def tile_count(self, cfg, g2d, depth, W, H):
    g2d = np.ascontiguousarray(g2d, np.float32); depth = np.ascontiguousarray(depth, np.float32)
    m = g2d.shape[0]
    counts = np.zeros(m, np.int32); total = C.c_int64(0)
    self.lib.gso_tile_count(C.byref(cfg), C.c_int64(m), self.p(g2d), self.p(depth), C.c_int(W), C.c_int(H),
                            self.p(counts), C.byref(total))
    return counts, total.value

  def tile_map(self, cfg, g2d, depth, W, H, max_overlaps=None):
    g2d = np.ascontiguousarray(g2d, np.float32); depth = np.ascontiguousarray(depth, np.float32)
    m = g2d.shape[0]
    if max_overlaps is None:
      max_overlaps = max(1, self.tile_count(cfg, g2d, depth, W, H)[1])
    T = ((W + 15) // 16) * ((H + 15) // 16)
    o2p = np.zeros(max_overlaps, np.int32); keys = np.zeros(max_overlaps, np.uint64)
    ranges = np.zeros((T, 2), np.int32)
    total = self.lib.gso_tile_map(C.byref(cfg), C.c_int64(m), self.p(g2d), self.p(depth), C.c_int(W), C.c_int(H),
                                  C.c_int64(max_overlaps), self.p(o2p), self.p(keys), self.p(ranges))
    k = min(total, max_overlaps)
    return o2p[:k], ranges, total, keys[:k]

  def radix_sort_pairs(self, keys, vals, begin_bit, end_bit):
    keys = np.ascontiguousarray(keys, np.uint32); vals = np.ascontiguousarray(vals, np.uint32)
    ko = np.zeros_like(keys); vo = np.zeros_like(vals)
    self.lib.gso_radix_sort_pairs(self.p(keys), self.p(vals), C.c_int64(keys.shape[0]), C.c_int(begin_bit),
                                  C.c_int(end_bit), self.p(ko), self.p(vo))
    return ko, vo

  # -- (3)/(4) rasterisation -------------------------------------------------------------------
  def raster_fwd(self, cfg, g2d, features, depth, o2p, ranges, W, H, visibility=True, median=False):
    g2d, features = self.r(g2d), self.r(features)
    depth = None if depth is None else self.r(depth)
    Cn = features.shape[1]
    image = np.zeros((H, W, Cn), self.real); fT = np.zeros((H, W), self.real); nc = np.zeros((H, W), np.int32)
    vis = np.zeros(g2d.shape[0], self.real) if visibility else None
    med = np.zeros((H, W), self.real) if median else None
    o2p = np.ascontiguousarray(o2p, np.int32); ranges = np.ascontiguousarray(ranges, np.int32)
    self.lib.gso_raster_fwd(C.byref(cfg), C.c_int(Cn), C.c_int64(g2d.shape[0]), self.p(g2d), self.p(features),
                            self.p(depth), self.p(o2p), self.p(ranges), C.c_int(W), C.c_int(H), self.p(image),
                            self.p(fT), self.p(nc), self.p(vis), self.p(med))
    return image, fT, nc, vis, med

  def raster_bwd(self, cfg, g2d, features, o2p, ranges, W, H, final_T, n_contrib, d_image):
    g2d, features, final_T, d_image = map(self.r, (g2d, features, final_T, d_image))
    Cn = features.shape[1]
    stride = acc_stride(Cn)
    acc = np.zeros((g2d.shape[0], stride), self.real)
    o2p = np.ascontiguousarray(o2p, np.int32); ranges = np.ascontiguousarray(ranges, np.int32)
    n_contrib = np.ascontiguousarray(n_contrib, np.int32)
    self.lib.gso_raster_bwd(C.byref(cfg), C.c_int(Cn), C.c_int64(g2d.shape[0]), self.p(g2d), self.p(features),
                            self.p(o2p), self.p(ranges), C.c_int(W), C.c_int(H), self.p(final_T), self.p(n_contrib),
                            self.p(d_image), self.p(acc), C.c_int(stride))
    return acc

  def loss_l1_mse(self, pred, target, w_l1, w_mse, target_u8=None):
    pred = self.r(pred)
    target = None if target is None else self.r(target)
    target_u8 = None if target_u8 is None else np.ascontiguousarray(target_u8, np.uint8)
    d_pred = np.zeros_like(pred); sums = np.zeros(2, self.real)
    self.lib.gso_loss_l1_mse(C.c_int64(pred.size), self.p(pred), self.p(target), self.p(target_u8),
                             self.creal(w_l1), self.creal(w_mse), self.p(d_pred), self.p(sums))
    return d_pred, sums

  # -- (5) optimiser / stats / densify -----------------------------------------------------------
  def optim_vis_update(self, ocfg, visible, running_vis, step):
    self.lib.gso_optim_vis_update(C.byref(ocfg), C.c_int64(visible.shape[0]), self.p(visible), self.p(running_vis),
                                  self.p(step))

  def optim_step(self, ocfg, kind, lr, param, grad, exp_avg, exp_avg_sq, visible, running_vis, step,
                 basis_ls=None, basis_q=None, post=0, post_lo=0.0, post_hi=0.0, zero_grad=1):
    n = param.shape[0]
    dim = 1 if param.ndim == 1 else int(np.prod(param.shape[1:]))
    self.lib.gso_optim_step(C.byref(ocfg), C.c_int(kind), C.c_int64(n), C.c_int(dim), self.creal(lr),
                            self.p(param), self.p(grad), self.p(exp_avg), self.p(exp_avg_sq), self.p(visible),
                            self.p(running_vis), self.p(step), self.p(basis_ls), self.p(basis_q), C.c_int(post),
                            self.creal(post_lo), self.creal(post_hi), C.c_int(zero_grad))

  def point_basis(self, log_scaling, rotation, eps=1e-4):
    ls, q = self.r(log_scaling), self.r(rotation)
    out = np.zeros((ls.shape[0], 3, 3), self.real)
    for i in range(ls.shape[0]):
      self.lib.gso_point_basis(self.p(ls[i]), self.p(q[i]), self.creal(eps), self.p(out[i]))
    return out

  def exp_lerp(self, t, a, b):
    return self.lib.gso_exp_lerp(t, a, b)

  def accumulate_stats(self, idx, g2d, vis, heur, split_alpha, prune_alpha, state):
    g2d, vis, heur = self.r(g2d), self.r(vis), self.r(heur)
    idx = None if idx is None else np.ascontiguousarray(idx, np.int64)
    self.lib.gso_accumulate_stats(C.c_int64(g2d.shape[0]), self.p(idx), self.p(g2d), self.p(vis), C.c_int(1),
                                  self.p(heur), C.c_int(2), self.creal(split_alpha), self.creal(prune_alpha),
                                  self.p(state["prune_cost"]), self.p(state["split_score"]),
                                  self.p(state["max_scale_px"]), self.p(state["points_in_view"]),
                                  self.p(state["visibility"]), self.p(state.get("scene_visible")))

  def select_topk_mask(self, values, k, descending):
    values = np.ascontiguousarray(values, np.float32)
    mask = np.zeros(values.shape[0], np.uint8)
    self.lib.gso_select_topk_mask(self.p(values), C.c_int64(values.shape[0]), C.c_int64(k), C.c_int(int(descending)),
                                  self.p(mask))
    return mask.astype(bool)

  def compact_slots(self, keep):
    keep = np.ascontiguousarray(keep, np.uint8)
    slots = np.zeros(keep.shape[0], np.int32)
    cnt = self.lib.gso_compact_slots(self.p(keep), C.c_int64(keep.shape[0]), self.p(slots))
    return slots, cnt

  def compact_rows(self, src, slots, count):
    src = np.ascontiguousarray(src)
    row_bytes = src.dtype.itemsize * int(np.prod(src.shape[1:])) if src.ndim > 1 else src.dtype.itemsize
    dst = np.zeros((count,) + src.shape[1:], src.dtype)
    self.lib.gso_compact_rows(self.p(src), self.p(dst), C.c_int64(src.shape[0]), C.c_int(row_bytes), self.p(slots))
    return dst

  def split_positions(self, split_idx, position, log_scaling, rotation, uniform, sep=0.7):
    position, log_scaling, rotation = map(self.r, (position, log_scaling, rotation))
    split_idx = np.ascontiguousarray(split_idx, np.int64)
    uniform = None if uniform is None else self.r(uniform)
    s = split_idx.shape[0]
    op = np.zeros((2 * s, 3), self.real); ols = np.zeros((2 * s, 3), self.real)
    self.lib.gso_split_positions(C.c_int64(s), self.p(split_idx), self.p(position), self.p(log_scaling),
                                 self.p(rotation), self.p(uniform), self.creal(sep), self.p(op), self.p(ols),
                                 C.c_int64(0))
    return op, ols

  def mcmc_noise(self, alpha_logit, log_scaling, rotation, points_in_view, min_views, opacity_threshold, noise_level,
                 randn, position):
    """in place on `position` (must be this oracle's real dtype and contiguous)."""
    al, ls, q, rn = map(self.r, (alpha_logit, log_scaling, rotation, randn))
    piv = np.ascontiguousarray(points_in_view, np.int16)
    assert position.dtype == self.real and position.flags["C_CONTIGUOUS"]
    self.lib.gso_mcmc_noise(C.c_int64(al.shape[0]), self.p(al), self.p(ls), self.p(q), self.p(piv), C.c_int(min_views),
                            self.creal(opacity_threshold), self.creal(noise_level), self.p(rn), self.p(position))

  def reg_loss(self, weights, g2d, depth, log_scaling, vis, d_opacity=None, d_depth=None, d_log_scaling=None):
    """weights = (scale, opacity, aspect).  Returns sums (4,); gradient arrays (own dtype, contiguous (n,)/(n,3))
    are updated in place when given."""
    g2d, depth, ls, vis = map(self.r, (g2d, depth, log_scaling, vis))
    sums = np.zeros(4, self.real)
    self.lib.gso_reg_loss(C.c_int64(g2d.shape[0]), self.creal(weights[0]), self.creal(weights[1]), self.creal(weights[2]),
                          self.p(g2d), self.p(depth), self.p(ls), self.p(vis), self.p(sums), self.p(d_opacity), C.c_int(1),
                          self.p(d_depth), self.p(d_log_scaling))
    return sums

  def ssim_loss(self, pred, target, levels=3, weight=1.0, want_grad=True):
    """pred/target (H,W,C).  Returns (weight * loss, weight * d loss/d pred or None, ssim of level 0)."""
    pred, target = self.r(pred), self.r(target)
    H, W, Cc = pred.shape
    d = np.zeros_like(pred) if want_grad else None
    s0 = self.creal(0)
    self.lib.gso_ssim_loss.restype = C.c_float if self.real == np.float32 else C.c_double
    loss = self.lib.gso_ssim_loss(C.c_int(H), C.c_int(W), C.c_int(Cc), C.c_int(levels), self.p(pred), self.p(target),
                                  self.creal(weight), self.p(d), C.byref(s0))
    return float(loss), d, float(s0.value)

  # -- (9) colour MLP --------------------------------------------------------------
  COLOR_WEIGHTS = ("base_w1", "base_b1", "base_w2", "base_b2", "dir_wsh", "dir_bsh", "dir_w1", "dir_b1", "dir_w2", "dir_b2")

  def rsh(self, degree, dirs):
    """real spherical harmonics of unit vectors (m,3) -> (m, (degree+1)^2), reference scene/mlp/rsh.py convention."""
    dirs = self.r(dirs)
    out = np.zeros((dirs.shape[0], (degree + 1) ** 2), self.real)
    for i in range(dirs.shape[0]):
      self.lib.gso_rsh(C.c_int(degree), self.p(dirs[i]), self.p(out[i]))
    return out

  def _color_args(self, weights, feature, position, cam_pos, glo):
    W = [self.r(weights[k]) for k in self.COLOR_WEIGHTS]
    feature, position, cam_pos, glo = map(self.r, (feature, position, cam_pos, glo))
    pf, gf = feature.shape[1], glo.reshape(-1).shape[0]
    H = W[2].shape[1]
    S = W[4].shape[1]
    deg = int(round(S ** 0.5)) - 1
    assert W[0].shape == (2 * H, pf + gf) and W[4].shape == (2 * (pf + gf), (deg + 1) ** 2)
    ptrs = (C.c_void_p * 10)(*[w.ctypes.data for w in W])
    return W, ptrs, feature, position, cam_pos, glo, pf, gf, H, deg

  def color_mlp_fwd(self, weights, feature, position, cam_pos, glo, in_view=None, half=False):
    """weights: dict COLOR_WEIGHTS -> arrays shaped like the reference's nn.Linear parameters.
    Returns (diffuse (n,3), specular (n,3)); rows with in_view <= 0 stay zero."""
    W, ptrs, feature, position, cam_pos, glo, pf, gf, H, deg = self._color_args(weights, feature, position, cam_pos, glo)
    n = feature.shape[0]
    iv = None if in_view is None else self.r(in_view)
    diffuse, specular = np.zeros((n, 3), self.real), np.zeros((n, 3), self.real)
    self.lib.gso_color_mlp_fwd(C.c_int64(n), C.c_int(pf), C.c_int(gf), C.c_int(H), C.c_int(deg), ptrs, self.p(feature),
                               self.p(position), self.p(cam_pos), self.p(glo), self.p(iv), C.c_int(1), C.c_int(int(half)),
                               self.p(diffuse), self.p(specular))
    return diffuse, specular

  def color_mlp_bwd(self, weights, feature, position, cam_pos, glo, d_diffuse, d_specular, row_w=None, half=False):
    """Returns (d_feature (n,pf), d_glo (gf,), dW dict shaped like weights)."""
    W, ptrs, feature, position, cam_pos, glo, pf, gf, H, deg = self._color_args(weights, feature, position, cam_pos, glo)
    n = feature.shape[0]
    rw = None if row_w is None else self.r(row_w)
    dd, ds = self.r(d_diffuse), self.r(d_specular)
    d_feature, d_glo = np.zeros((n, pf), self.real), np.zeros((gf,), self.real)
    dW = [np.zeros_like(w) for w in W]
    dptrs = (C.c_void_p * 10)(*[w.ctypes.data for w in dW])
    self.lib.gso_color_mlp_bwd(C.c_int64(n), C.c_int(pf), C.c_int(gf), C.c_int(H), C.c_int(deg), ptrs, self.p(feature),
                               self.p(position), self.p(cam_pos), self.p(glo), self.p(rw), C.c_int(1), C.c_int(int(half)),
                               self.p(dd), self.p(ds), self.p(d_feature), self.p(d_glo), dptrs)
    return d_feature, d_glo, dict(zip(self.COLOR_WEIGHTS, dW))

  # -- (10) point clustering -------------------------------------------------------
  def assign_clusters(self, x, centroids):
    x, c = self.r(x), self.r(centroids)
    labels = np.zeros(x.shape[0], np.int32)
    self.lib.gso_assign_clusters(C.c_int64(x.shape[0]), C.c_int(x.shape[1]), C.c_int(c.shape[0]), self.p(x), self.p(c),
                                 self.p(labels))
    return labels

  def kmeans(self, x, centroids, iters):
    """-> (labels of the last assignment (n,) int32, centroids after the last update (k,d))."""
    x, c = self.r(x), self.r(centroids).copy()
    labels = np.zeros(x.shape[0], np.int32)
    self.lib.gso_kmeans(C.c_int64(x.shape[0]), C.c_int(x.shape[1]), C.c_int(c.shape[0]), C.c_int(iters), self.p(x), self.p(c),
                        self.p(labels))
    return labels, c

  # -- (11) estimate_scale ----------------------------------------------------------
  def knn_scale(self, points, num_neighbors=3):
    pts = self.r(points)
    out = np.zeros(pts.shape[0], self.real)
    self.lib.gso_knn_scale(C.c_int64(pts.shape[0]), self.p(pts), C.c_int(num_neighbors), self.p(out))
    return out

  # -- (12) per-view cluster features / frustum counts ----------------------------------
  def view_features(self, labels, num_clusters, vis, indexes=None, threshold=0.01):
    vis = self.r(vis)
    labels = np.ascontiguousarray(labels, np.int32)
    idx = None if indexes is None else np.ascontiguousarray(indexes, np.int64)
    out = np.zeros(num_clusters, self.real)
    self.lib.gso_view_features(C.c_int64(vis.shape[0]), self.p(vis), self.p(idx), self.p(labels), self.creal(threshold),
                               C.c_int(num_clusters), self.p(out))
    return out

  def point_visibility(self, points, cameras):
    pts, cams = self.r(points), self.r(cameras)
    counts = np.zeros(pts.shape[0], np.int32)
    self.lib.gso_point_visibility(C.c_int64(pts.shape[0]), self.p(pts), C.c_int(cams.shape[0]), self.p(cams), self.p(counts))
    return counts
