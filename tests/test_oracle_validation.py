"""Validation of the CPU oracle itself (CPU only).

The oracle's kernel stages are 'parity unpinned' (no reference source/tests at the taichi-splatting
boundary), so before it is trusted as the checker for the CUDA path it is checked against
  (i)   an independent float64 numpy projection (direct J W Σ Wᵀ Jᵀ matrix form),
  (ii)  an independent float64 brute-force renderer with no tiles and no tile sort,
  (iii) central finite differences of the whole chain params -> image (f64 build),
  (iv)  the f32 build against the f64 build.
"""
import numpy as np
import pytest

from oracle import gs_oracle as go
from util import random_camera, random_gaussians


def cam_arrays(cam):
  T = np.array(list(cam.T), np.float64).reshape(4, 4)
  return T, cam.fx, cam.fy, cam.cx, cam.cy, cam.width, cam.height


def quat_R(q):
  x, y, z, w = q / np.linalg.norm(q)
  return np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y)],
                   [2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x)],
                   [2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)]])


def numpy_project(cam, cfg, g):
  """independent restatement: full 3x3 covariance, then J W Σ Wᵀ Jᵀ."""
  T, fx, fy, cx, cy, W, H = cam_arrays(cam)
  n = g["position"].shape[0]
  out = np.zeros((n, 6)); depth = np.full(n, np.inf)
  for i in range(n):
    p = T[:3, :3] @ g["position"][i].astype(np.float64) + T[:3, 3]
    if not (cam.near_plane < p[2] < cam.far_plane):
      continue
    R = quat_R(g["rotation"][i].astype(np.float64))
    S = np.diag(np.exp(g["log_scaling"][i].astype(np.float64)))
    Sigma = R @ S @ S @ R.T
    limx, limy = cfg.frustum_clamp * 0.5 * W / fx, cfg.frustum_clamp * 0.5 * H / fy
    tx, ty = np.clip(p[0] / p[2], -limx, limx), np.clip(p[1] / p[2], -limy, limy)
    J = np.array([[fx / p[2], 0, -fx * tx / p[2]], [0, fy / p[2], -fy * ty / p[2]]])
    cov = J @ T[:3, :3] @ Sigma @ T[:3, :3].T @ J.T
    det0 = np.linalg.det(cov)
    cov = cov + cfg.blur_cov * np.eye(2)
    det1 = np.linalg.det(cov)
    op = 1 / (1 + np.exp(-float(g["alpha_logit"][i])))
    if cfg.antialias:
      op *= np.sqrt(max(det0 / det1, 0))
    if op <= cfg.alpha_threshold:
      continue
    con = np.linalg.inv(cov)
    out[i] = [fx * p[0] / p[2] + cx, fy * p[1] / p[2] + cy, con[0, 0], con[0, 1], con[1, 1], op]
    depth[i] = p[2]
  return out, depth


def brute_render(g2d, depth, feats, W, H, cfg):
  """no tiles: every pixel visits every splat in (float32 depth bits, index) order."""
  n = g2d.shape[0]
  order = np.lexsort((np.arange(n), depth.astype(np.float32).view(np.uint32)))
  X, Y = np.meshgrid(np.arange(W) + 0.5, np.arange(H) + 0.5)
  T = np.ones((H, W)); img = np.zeros((H, W, feats.shape[1])); vis = np.zeros(n)
  alive = np.ones((H, W), bool)
  tmin = 1.0 - float(np.float32(cfg.saturate_threshold))
  for i in order:
    if not g2d[i, 5] > 0:
      continue
    dx, dy = X - g2d[i, 0], Y - g2d[i, 1]
    power = -0.5 * (g2d[i, 2] * dx * dx + g2d[i, 4] * dy * dy) - g2d[i, 3] * dx * dy
    a = np.minimum(cfg.clamp_max_alpha, g2d[i, 5] * np.exp(np.minimum(power, 0)))
    m = alive & (power <= 0) & (a >= cfg.alpha_threshold)
    w = np.where(m, a * T, 0.0)
    img += w[..., None] * feats[i]
    vis[i] = w.sum()
    T = np.where(m, T * (1 - a), T)
    alive &= ~(m & (T < tmin))
  return img, T, vis


@pytest.mark.parametrize("antialias", [0, 1])
def test_projection_matches_independent_numpy(antialias):
  orc = go.Oracle("f64")
  cam = random_camera()
  cfg = go.make_config(antialias=antialias)
  g = random_gaussians(400, seed=1)
  g2d, depth, _ = orc.project_fwd(cam, cfg, g["position"], g["log_scaling"], g["rotation"], g["alpha_logit"])
  ref, ref_depth = numpy_project(cam, cfg, g)
  inview = g2d[:, 5] > 0
  assert inview.sum() > 100
  # oracle culls more (rows touching no tile); wherever it kept a row the numbers must agree
  np.testing.assert_allclose(g2d[inview], ref[inview], rtol=1e-9, atol=1e-9)
  np.testing.assert_allclose(depth[inview], ref_depth[inview], rtol=1e-12)
  # and whatever it dropped beyond numpy's cull really touches no pixel centre of the image
  dropped = (~inview) & (ref[:, 5] > 0)
  img, _, vis = brute_render(ref, ref_depth, np.ones((400, 1)), cam.width, cam.height, cfg)
  assert np.all(vis[dropped] == 0)


@pytest.mark.parametrize("size", [(96, 64), (101, 75)])
def test_raster_matches_bruteforce(size):
  W, H = size
  orc = go.Oracle("f64")
  cam = random_camera(W, H)
  cfg = go.make_config()
  g = random_gaussians(500, seed=2, scale=0.12)
  g2d, depth, rgb = orc.project_fwd(cam, cfg, g["position"], g["log_scaling"], g["rotation"], g["alpha_logit"],
                                    g["sh"], 3)
  o2p, ranges, total, _ = orc.tile_map(cfg, g2d, depth, W, H)
  assert total > 500
  image, fT, nc, vis, _ = orc.raster_fwd(cfg, g2d, rgb, depth, o2p, ranges, W, H)
  ref_img, ref_T, ref_vis = brute_render(g2d, depth, rgb, W, H, cfg)
  np.testing.assert_allclose(image, ref_img, atol=1e-10)
  np.testing.assert_allclose(fT, ref_T, atol=1e-12)
  np.testing.assert_allclose(vis, ref_vis, atol=1e-9)


def _chain(orc, cam, cfg, g, Wimg, wdepth, sh_degree=3):
  W, H = cam.width, cam.height
  g2d, depth, rgb = orc.project_fwd(cam, cfg, g["position"], g["log_scaling"], g["rotation"], g["alpha_logit"],
                                    g["sh"], sh_degree)
  o2p, ranges, _, _ = orc.tile_map(cfg, g2d, depth, W, H)
  image, fT, nc, vis, _ = orc.raster_fwd(cfg, g2d, rgb, depth, o2p, ranges, W, H)
  inview = g2d[:, 5] > 0
  loss = float((image * Wimg).sum() + (np.where(inview, depth, 0.0) * wdepth).sum())
  return loss, (g2d, depth, rgb, o2p, ranges, fT, nc)


@pytest.mark.parametrize("antialias", [0, 1])
def test_gradients_match_finite_differences(antialias):
  orc = go.Oracle("f64")
  W, H = 64, 48
  cam = random_camera(W, H)
  cfg = go.make_config(antialias=antialias)
  n = 120
  g = {k: v.astype(np.float64) for k, v in random_gaussians(n, seed=3, scale=0.15).items()}
  g["alpha_logit"] = np.clip(g["alpha_logit"], -2, 1.5)  # keep away from the 0.99 clamp
  rng = np.random.default_rng(0)
  Wimg = rng.normal(0, 1, (H, W, 3)); wdepth = rng.normal(0, 0.1, n)
  loss, (g2d, depth, rgb, o2p, ranges, fT, nc) = _chain(orc, cam, cfg, g, Wimg, wdepth)
  acc = orc.raster_bwd(cfg, g2d, rgb, o2p, ranges, W, H, fT, nc, Wimg)
  grads = orc.project_bwd(cam, cfg, g["position"], g["log_scaling"], g["rotation"], g["alpha_logit"], g["sh"], 3,
                          g2d, acc[:, 0:6], acc[:, 6:9], wdepth)
  inview = np.nonzero(g2d[:, 5] > 0)[0]
  assert len(inview) > 30
  h = 1e-6
  checked = 0
  bad = []
  for name in ["position", "log_scaling", "rotation", "alpha_logit", "sh"]:
    key = name
    flat = g[key].reshape(n, -1)
    gflat = grads[name].reshape(n, -1)
    for i in inview[:25]:
      for j in range(min(flat.shape[1], 6)):
        old = flat[i, j]
        flat[i, j] = old + h; lp, _ = _chain(orc, cam, cfg, g, Wimg, wdepth)
        flat[i, j] = old - h; lm, _ = _chain(orc, cam, cfg, g, Wimg, wdepth)
        flat[i, j] = old
        fd = (lp - lm) / (2 * h)
        an = gflat[i, j]
        checked += 1
        if abs(fd - an) > 1e-5 + 1e-4 * max(abs(fd), abs(an)):
          bad.append((name, int(i), j, fd, an))
  # thresholds (alpha cut, saturation) make the image piecewise smooth; allow a couple of FD hits on a kink
  assert len(bad) <= max(2, checked // 200), bad[:10]


def test_f32_oracle_matches_f64():
  o32, o64 = go.Oracle("f32"), go.Oracle("f64")
  W, H = 96, 64
  cam = random_camera(W, H); cfg = go.make_config()
  g = random_gaussians(600, seed=4, scale=0.1)
  out = {}
  for name, orc in (("f32", o32), ("f64", o64)):
    g2d, depth, rgb = orc.project_fwd(cam, cfg, g["position"], g["log_scaling"], g["rotation"], g["alpha_logit"],
                                      g["sh"], 3)
    out[name] = (g2d, depth, rgb)
  a, b = out["f32"], out["f64"]
  same = (a[0][:, 5] > 0) == (b[0][:, 5] > 0)
  assert same.mean() > 0.995
  m = (a[0][:, 5] > 0) & (b[0][:, 5] > 0)
  np.testing.assert_allclose(a[0][m], b[0][m], rtol=2e-4, atol=2e-4)
  np.testing.assert_allclose(a[2][m], b[2][m], rtol=1e-4, atol=1e-5)
  # raster on IDENTICAL float32 projected rows: only the per-pixel arithmetic differs
  g2d32, depth32, rgb32 = a
  o2p, ranges, _, _ = o32.tile_map(cfg, g2d32, depth32, W, H)
  rng = np.random.default_rng(1)
  dimg = rng.normal(0, 1, (H, W, 3))
  res = {}
  for name, orc in (("f32", o32), ("f64", o64)):
    image, fT, nc, vis, _ = orc.raster_fwd(cfg, g2d32, rgb32, depth32, o2p, ranges, W, H)
    acc = orc.raster_bwd(cfg, g2d32, rgb32, o2p, ranges, W, H, fT, nc, dimg)
    res[name] = (image, fT, vis, acc)
  bad = np.abs(res["f32"][0] - res["f64"][0]).max(axis=-1) > 1e-4
  assert bad.mean() < 1e-3  # pixels sitting exactly on the 1/255 cut may flip
  ok = ~bad
  np.testing.assert_allclose(res["f32"][0][ok], res["f64"][0][ok], atol=1e-4)
  scale = np.abs(res["f64"][3]).max(axis=0) + 1e-6
  err = np.abs(res["f32"][3] - res["f64"][3]) / scale
  assert np.quantile(err, 0.999) < 1e-3
