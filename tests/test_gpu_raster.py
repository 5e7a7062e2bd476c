"""GPU parity: rasteriser forward / backward, projection backward and the whole chain vs the CPU oracle."""
import numpy as np
import pytest
import torch

from gsb200 import ops
from gpu_util import cu, npy, oracle, to_lib_camera, to_lib_config
from oracle import gs_oracle as go
from util import random_camera, random_gaussians

pytestmark = pytest.mark.gpu

IMG_TOL = 1e-4  # north_star: max abs 1e-4 relative to the image range [0,1]


def _setup(W, H, n, seed, channels=3, scale=0.1, sh_degree=3, **cfgkw):
  orc = oracle()
  cam = random_camera(W, H)
  cfg = go.make_config(**cfgkw)
  g = random_gaussians(n, seed=seed, scale=scale)
  gc = {k: cu(v) for k, v in g.items()}
  lcam, lcfg = to_lib_camera(cam), to_lib_config(cfg)
  g2d, depth, rgb = ops.project_fwd(lcam, lcfg, gc["position"], gc["log_scaling"], gc["rotation"], gc["alpha_logit"],
                                    gc["sh"], sh_degree)
  if channels == 3:
    feat = rgb
  else:
    rng = np.random.default_rng(seed)
    feat = cu(rng.uniform(0, 1, (n, channels)).astype(np.float32))
  o2p, ranges, counters, _ = ops.tile_map(lcfg, g2d, depth, W, H)
  total = int(counters[0].item())
  return dict(orc=orc, cam=cam, cfg=cfg, lcam=lcam, lcfg=lcfg, g=g, gc=gc, g2d=g2d, depth=depth, feat=feat, o2p=o2p,
              ranges=ranges, total=total, W=W, H=H)


@pytest.mark.parametrize("channels", [1, 3, 4])
@pytest.mark.parametrize("size,n", [((96, 64), 600), ((101, 75), 3000), ((320, 240), 20000), ((640, 480), 100000)])
def test_raster_fwd_matches_oracle(channels, size, n):
  W, H = size
  s = _setup(W, H, n, seed=30, channels=channels)
  image, fT, nc, vis, med = ops.raster_fwd(s["lcfg"], s["g2d"], s["feat"], s["depth"], s["o2p"], s["ranges"], W, H,
                                           visibility=True, median=True)
  rimage, rfT, rnc, rvis, rmed = s["orc"].raster_fwd(s["cfg"], npy(s["g2d"]), npy(s["feat"]), npy(s["depth"]),
                                                     npy(s["o2p"])[:s["total"]], npy(s["ranges"]), W, H,
                                                     visibility=True, median=True)
  image, fT, nc, vis, med = map(npy, (image, fT, nc, vis, med))
  # the blend decision is taken in the exponent domain with IEEE ops shared by both sides (DESIGN.md "Spec R"), so no
  # pixel may flip: strict max-abs over the whole image, identical contributor counts and median depths
  assert np.abs(image - rimage).max() < IMG_TOL
  assert np.abs(fT - rfT).max() < IMG_TOL
  assert (nc != rnc).mean() < 1e-5
  scale = max(rvis.max(), 1.0)
  assert np.abs(vis - rvis).max() / scale < 1e-4
  assert (med != rmed).mean() < 1e-5


def test_raster_fwd_no_visibility_early_exit_same_image():
  W, H = 160, 120
  s = _setup(W, H, 8000, seed=31, scale=0.2)
  a = ops.raster_fwd(s["lcfg"], s["g2d"], s["feat"], s["depth"], s["o2p"], s["ranges"], W, H, visibility=True)
  b = ops.raster_fwd(s["lcfg"], s["g2d"], s["feat"], s["depth"], s["o2p"], s["ranges"], W, H, visibility=False)
  assert torch.equal(a[0], b[0]) and torch.equal(a[1], b[1]) and torch.equal(a[2], b[2])


@pytest.mark.parametrize("channels", [1, 3, 4])
def test_raster_bwd_matches_oracle(channels):
  W, H = 128, 96
  s = _setup(W, H, 4000, seed=32, channels=channels, scale=0.12)
  image, fT, nc, _, _ = ops.raster_fwd(s["lcfg"], s["g2d"], s["feat"], s["depth"], s["o2p"], s["ranges"], W, H)
  rng = np.random.default_rng(1)
  d_image = rng.normal(0, 1, (H, W, channels)).astype(np.float32)
  acc = ops.raster_bwd(s["lcfg"], s["g2d"], s["feat"], s["o2p"], s["ranges"], W, H, fT, nc, cu(d_image))
  # the oracle backward consumes the SAME forward state, so only the backward arithmetic is compared
  racc = s["orc"].raster_bwd(s["cfg"], npy(s["g2d"]), npy(s["feat"]), npy(s["o2p"])[:s["total"]], npy(s["ranges"]), W, H,
                             npy(fT), npy(nc), d_image)
  acc = npy(acc)
  assert acc.shape == racc.shape
  ncol = 9 + channels
  scale = np.abs(racc[:, :ncol]).max(axis=0) + 1e-12
  err = np.abs(acc[:, :ncol] - racc[:, :ncol]) / scale
  # sums of up to thousands of fp32 terms in a different order + ex2.approx; cut-boundary pixels may flip
  assert np.quantile(err, 0.999) < 1e-4, np.quantile(err, [0.5, 0.99, 0.999, 1.0])
  assert err.max() < 5e-3, err.max(axis=0)
  # visibility column equals the forward flag's visibility
  vis_f = npy(ops.raster_fwd(s["lcfg"], s["g2d"], s["feat"], s["depth"], s["o2p"], s["ranges"], W, H, visibility=True)[3])
  assert np.abs(acc[:, 8 + channels] - vis_f).max() / max(vis_f.max(), 1.0) < 1e-4


def test_raster_bwd_linear_in_d_image():
  """size-independent property: the backward is linear in dL/dimage (heuristic columns excepted)."""
  W, H = 200, 150
  s = _setup(W, H, 15000, seed=33)
  _, fT, nc, _, _ = ops.raster_fwd(s["lcfg"], s["g2d"], s["feat"], s["depth"], s["o2p"], s["ranges"], W, H)
  g = torch.Generator(device="cuda").manual_seed(0)
  d1 = torch.randn((H, W, 3), device="cuda", generator=g)
  d2 = torch.randn((H, W, 3), device="cuda", generator=g)
  run = lambda d: ops.raster_bwd(s["lcfg"], s["g2d"], s["feat"], s["o2p"], s["ranges"], W, H, fT, nc, d)
  a1, a2, a12 = run(d1), run(d2), run(2.0 * d1 - 3.0 * d2)
  lin = 2.0 * a1[:, :9] - 3.0 * a2[:, :9]
  scale = a12[:, :9].abs().amax(dim=0) + 1e-12
  assert ((a12[:, :9] - lin).abs() / scale).max().item() < 1e-4


@pytest.mark.parametrize("antialias", [0, 1])
@pytest.mark.parametrize("sh_degree", [0, 3])
def test_project_bwd_matches_oracle(antialias, sh_degree):
  W, H = 128, 96
  orc = oracle()
  cam = random_camera(W, H); cfg = go.make_config(antialias=antialias)
  n = 3000
  g = random_gaussians(n, seed=34, scale=0.12)
  K = (sh_degree + 1) ** 2
  g["sh"] = np.ascontiguousarray(g["sh"][:, :, :K])
  gc = {k: cu(v) for k, v in g.items()}
  lcam, lcfg = to_lib_camera(cam), to_lib_config(cfg)
  g2d, depth, rgb = ops.project_fwd(lcam, lcfg, gc["position"], gc["log_scaling"], gc["rotation"], gc["alpha_logit"],
                                    gc["sh"], sh_degree)
  rng = np.random.default_rng(2)
  acc = rng.normal(0, 1, (n, 12)).astype(np.float32)
  acc[:, 11] = rng.uniform(-1, 1, n)  # row_vis: about half the rows are skipped
  d_depth = rng.normal(0, 1, n).astype(np.float32)
  acc_c = cu(acc)
  grads = ops.project_bwd(lcam, lcfg, gc["position"], gc["log_scaling"], gc["rotation"], gc["alpha_logit"], gc["sh"],
                          sh_degree, g2d, acc_c[:, 0:6], acc_c[:, 6:9], cu(d_depth), row_vis=acc_c[:, 11])
  rgrads = orc.project_bwd(cam, cfg, g["position"], g["log_scaling"], g["rotation"], g["alpha_logit"], g["sh"],
                           sh_degree, npy(g2d), acc[:, 0:6], acc[:, 6:9], d_depth, row_vis=acc[:, 11])
  inview = npy(g2d)[:, 5] > 0
  assert (inview & (acc[:, 11] > 0)).sum() > 300
  for name in ["position", "log_scaling", "rotation", "alpha_logit", "sh"]:
    a, r = npy(grads[name]).reshape(n, -1), rgrads[name].reshape(n, -1)
    scale = np.abs(r).max(axis=0) + 1e-12
    err = np.abs(a - r) / scale
    assert np.quantile(err, 0.999) < 1e-4, (name, np.quantile(err, [0.5, 0.99, 0.999, 1.0]))
    assert err.max() < 1e-2, (name, err.max())
    skipped = ~(inview & (acc[:, 11] > 0))
    assert np.all(a[skipped] == 0), name
  # accumulate=True adds on top
  grads2 = ops.project_bwd(lcam, lcfg, gc["position"], gc["log_scaling"], gc["rotation"], gc["alpha_logit"], gc["sh"],
                           sh_degree, g2d, acc_c[:, 0:6], acc_c[:, 6:9], cu(d_depth), row_vis=acc_c[:, 11],
                           grads={k: v.clone() for k, v in grads.items()}, accumulate=True)
  for name in grads:
    assert torch.allclose(grads2[name], 2 * grads[name], rtol=1e-6, atol=0)


def test_full_chain_matches_oracle_chain():
  """whole path params -> image -> grads on the GPU vs the whole path in the oracle (each with its own
  intermediates).  Image max-abs within 1e-4; parameter gradients within 1e-4 of each column's range for 99.9 %
  of the entries (projection rounding can move a row across a cull boundary)."""
  W, H = 160, 120
  orc = oracle()
  cam = random_camera(W, H); cfg = go.make_config()
  n = 6000
  g = random_gaussians(n, seed=35, scale=0.1)
  rng = np.random.default_rng(3)
  target = rng.uniform(0, 1, (H, W, 3)).astype(np.float32)
  # --- oracle
  rg2d, rdepth, rrgb = orc.project_fwd(cam, cfg, g["position"], g["log_scaling"], g["rotation"], g["alpha_logit"],
                                       g["sh"], 3)
  ro2p, rranges, rtotal, _ = orc.tile_map(cfg, rg2d, rdepth, W, H)
  rimage, rfT, rnc, _, _ = orc.raster_fwd(cfg, rg2d, rrgb, rdepth, ro2p, rranges, W, H)
  rdimg, rsums = orc.loss_l1_mse(rimage, target, 0.2, 10.0)
  racc = orc.raster_bwd(cfg, rg2d, rrgb, ro2p, rranges, W, H, rfT, rnc, rdimg)
  rgrads = orc.project_bwd(cam, cfg, g["position"], g["log_scaling"], g["rotation"], g["alpha_logit"], g["sh"], 3, rg2d,
                           racc[:, 0:6], racc[:, 6:9], None, row_vis=racc[:, 11])
  # --- GPU
  gc = {k: cu(v) for k, v in g.items()}
  lcam, lcfg = to_lib_camera(cam), to_lib_config(cfg)
  acc = torch.full((n, 12), 7.0, device="cuda")  # project_fwd must clear in-view rows itself
  g2d, depth, rgb = ops.project_fwd(lcam, lcfg, gc["position"], gc["log_scaling"], gc["rotation"], gc["alpha_logit"],
                                    gc["sh"], 3, acc=acc)
  o2p, ranges, counters, _ = ops.tile_map(lcfg, g2d, depth, W, H)
  image, fT, nc, _, _ = ops.raster_fwd(lcfg, g2d, rgb, depth, o2p, ranges, W, H)
  dimg, sums = ops.loss_l1_mse(image, cu(target), 0.2, 10.0)
  ops.raster_bwd(lcfg, g2d, rgb, o2p, ranges, W, H, fT, nc, dimg, acc=acc)
  grads = ops.project_bwd(lcam, lcfg, gc["position"], gc["log_scaling"], gc["rotation"], gc["alpha_logit"], gc["sh"], 3,
                          g2d, acc[:, 0:6], acc[:, 6:9], None, row_vis=acc[:, 11])
  assert np.abs(npy(image) - rimage).max() < IMG_TOL
  np.testing.assert_allclose(npy(sums), rsums, rtol=1e-4)
  for name in ["position", "log_scaling", "rotation", "alpha_logit", "sh"]:
    a, r = npy(grads[name]).reshape(n, -1), rgrads[name].reshape(n, -1)
    scale = np.abs(r).max(axis=0) + 1e-12
    e = np.abs(a - r) / scale
    assert np.quantile(e, 0.999) < 1e-4, (name, np.quantile(e, [0.5, 0.99, 0.999, 1.0]))


def test_loss_matches_oracle_float_and_u8():
  orc = oracle()
  rng = np.random.default_rng(4)
  pred = rng.uniform(-0.2, 1.2, (37, 53, 3)).astype(np.float32)
  tf = rng.uniform(0, 1, pred.shape).astype(np.float32)
  tu = rng.integers(0, 256, pred.shape, dtype=np.uint8)
  for target in (tf, tu):
    d, s = ops.loss_l1_mse(cu(pred), cu(target), 0.3, 10.0)
    rd, rs = orc.loss_l1_mse(pred, tf if target is tf else None, 0.3, 10.0, target_u8=None if target is tf else tu)
    np.testing.assert_allclose(npy(d), rd, rtol=1e-5, atol=1e-9)
    np.testing.assert_allclose(npy(s), rs, rtol=1e-5)
