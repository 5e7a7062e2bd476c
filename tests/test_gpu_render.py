"""GPU parity tests (run on the B200 with -m gpu): libgsb200 through its C-ABI vs the CPU oracle.

Tolerances (north_star): tile-overlap lists and sort order BIT-EXACT; images within 1e-4 max-abs of the
image range [0,1]; gradients within 1e-4 of each gradient column's range (thresholded pixels excepted,
see the individual tests).
"""
import numpy as np
import pytest
import torch

from gsb200 import lib as L
from gsb200 import ops
from gpu_util import cu, npy, oracle, to_lib_camera, to_lib_config
from oracle import gs_oracle as go
from util import random_camera, random_gaussians

pytestmark = pytest.mark.gpu


def _project_gpu(cam, cfg, g, sh_degree=3, acc=None):
  gc = {k: cu(v) for k, v in g.items()}
  sh = gc["sh"][:, :, :(sh_degree + 1) ** 2].contiguous()
  out = ops.project_fwd(to_lib_camera(cam), to_lib_config(cfg), gc["position"], gc["log_scaling"], gc["rotation"],
                        gc["alpha_logit"], sh, sh_degree, acc=acc)
  return gc, sh, out


@pytest.mark.parametrize("antialias", [0, 1])
@pytest.mark.parametrize("sh_degree", [0, 1, 2, 3])
def test_project_fwd_matches_oracle(antialias, sh_degree):
  orc = oracle()
  cam = random_camera(160, 120)
  cfg = go.make_config(antialias=antialias)
  g = random_gaussians(5000, seed=10 + sh_degree)
  K = (sh_degree + 1) ** 2
  sh = np.ascontiguousarray(g["sh"][:, :, :K])
  rg2d, rdepth, rrgb = orc.project_fwd(cam, cfg, g["position"], g["log_scaling"], g["rotation"], g["alpha_logit"], sh,
                                       sh_degree)
  _, _, (g2d, depth, rgb) = _project_gpu(cam, cfg, g, sh_degree)
  g2d, depth, rgb = npy(g2d), npy(depth), npy(rgb)
  in_ref, in_gpu = rg2d[:, 5] > 0, g2d[:, 5] > 0
  assert in_ref.sum() > 1000
  # FMA contraction on the GPU may flip a row that sits exactly on a cull boundary
  assert (in_ref != in_gpu).mean() < 2e-3
  m = in_ref & in_gpu
  np.testing.assert_allclose(g2d[m, :2], rg2d[m, :2], rtol=1e-5, atol=2e-3)  # pixels
  np.testing.assert_allclose(g2d[m, 2:5], rg2d[m, 2:5], rtol=2e-3, atol=1e-6)  # conic (1/det amplifies rounding)
  np.testing.assert_allclose(g2d[m, 5], rg2d[m, 5], rtol=1e-4, atol=1e-6)
  np.testing.assert_allclose(depth[m], rdepth[m], rtol=4e-6)  # FMA contraction
  np.testing.assert_allclose(rgb[m], rrgb[m], rtol=1e-5, atol=1e-5)
  out = ~in_gpu
  assert np.all(np.isinf(depth[out])) and np.all(g2d[out, 5] == 0) and np.all(rgb[out] == 0)


@pytest.mark.parametrize("size,n", [((96, 64), 700), ((101, 75), 1500), ((640, 480), 20000), ((33, 17), 50)])
def test_tile_map_bit_exact(size, n):
  """tile-overlap list, sort order and ranges equal the oracle's literal (tile<<32|depth, row) sort."""
  W, H = size
  orc = oracle()
  cam = random_camera(W, H)
  cfg = go.make_config()
  g = random_gaussians(n, seed=20, scale=0.1)
  _, _, (g2d, depth, _) = _project_gpu(cam, cfg, g)
  lcfg = to_lib_config(cfg)
  counts, counters = ops.tile_count(lcfg, g2d, depth, W, H)
  rcounts, rtotal = orc.tile_count(cfg, npy(g2d), npy(depth), W, H)
  assert np.array_equal(npy(counts), rcounts)
  assert int(counters[0].item()) == rtotal
  o2p, ranges, counters, _ = ops.tile_map(lcfg, g2d, depth, W, H)
  ro2p, rranges, rtotal, _ = orc.tile_map(cfg, npy(g2d), npy(depth), W, H)
  c = npy(counters)
  assert c[0] == rtotal and c[1] == 0 and c[2] == rtotal
  assert rtotal > n // 4
  assert np.array_equal(npy(o2p)[:rtotal], ro2p)
  assert np.array_equal(npy(ranges), rranges)


def test_tile_map_depth_ties_and_duplicates():
  """identical depths: ties must resolve by row index (stable sort)."""
  W, H = 64, 48
  orc = oracle()
  cfg = go.make_config()
  rng = np.random.default_rng(0)
  m = 600
  g2d = np.zeros((m, 6), np.float32)
  g2d[:, 0] = rng.uniform(-5, W + 5, m); g2d[:, 1] = rng.uniform(-5, H + 5, m)
  g2d[:, 2] = g2d[:, 4] = rng.uniform(0.005, 0.5, m); g2d[:, 3] = 0
  g2d[:, 5] = rng.uniform(0.0, 1.0, m)
  depth = rng.choice(np.array([1.0, 1.5, 2.0, 7.25], np.float32), m)
  g2d[::7, 5] = 0  # culled rows
  depth[::7] = np.inf
  o2p, ranges, counters, _ = ops.tile_map(to_lib_config(cfg), cu(g2d), cu(depth), W, H)
  ro2p, rranges, rtotal, _ = orc.tile_map(cfg, g2d, depth, W, H)
  assert int(counters[0].item()) == rtotal
  assert np.array_equal(npy(o2p)[:rtotal], ro2p)
  assert np.array_equal(npy(ranges), rranges)


def test_tile_map_empty_and_all_culled():
  cfg = to_lib_config(go.make_config())
  W, H = 48, 32
  g2d = torch.zeros((0, 6), device="cuda"); depth = torch.zeros((0,), device="cuda")
  o2p, ranges, counters, _ = ops.tile_map(cfg, g2d, depth, W, H, max_overlaps=16)
  torch.cuda.synchronize()
  assert int(counters[0].item()) == 0 and int(ranges.abs().sum().item()) == 0
  g2d = torch.tensor([[0, 0, 1, 0, 1, 0]] * 10, device="cuda", dtype=torch.float32)
  depth = torch.full((10,), float("inf"), device="cuda")
  o2p, ranges, counters, _ = ops.tile_map(cfg, g2d, depth, W, H, max_overlaps=16)
  assert int(counters[0].item()) == 0 and int(ranges.abs().sum().item()) == 0


def test_tile_map_overflow_flag():
  W, H = 96, 64
  cam = random_camera(W, H); cfg = go.make_config()
  g = random_gaussians(800, seed=21, scale=0.1)
  _, _, (g2d, depth, _) = _project_gpu(cam, cfg, g)
  _, counters = ops.tile_count(to_lib_config(cfg), g2d, depth, W, H)
  total = int(counters[0].item())
  cap = total // 2
  o2p, ranges, counters, _ = ops.tile_map(to_lib_config(cfg), g2d, depth, W, H, max_overlaps=cap)
  c = npy(counters)
  assert c[0] == total and c[1] == 1 and c[2] == cap
  r = npy(ranges)
  assert r.max() <= cap and np.all(r[:, 0] <= r[:, 1])


@pytest.mark.parametrize("n", [0, 1, 31, 4096, 4097, 100003])
@pytest.mark.parametrize("bits", [(0, 32), (0, 13), (8, 24), (0, 8)])
def test_radix_sort_pairs_bit_exact(n, bits):
  orc = oracle()
  rng = np.random.default_rng(n + bits[1])
  keys = rng.integers(0, 2 ** 32, n, dtype=np.uint64).astype(np.uint32)
  if n > 10:
    keys[: n // 3] = keys[0]  # many duplicates -> stability matters
  vals = np.arange(n, dtype=np.uint32)
  rk, rv = orc.radix_sort_pairs(keys, vals, bits[0], bits[1])
  k, v = ops.radix_sort_pairs(cu(keys.view(np.int32)), cu(vals.view(np.int32)), bits[0], bits[1])
  assert np.array_equal(npy(k).view(np.uint32), rk)
  assert np.array_equal(npy(v).view(np.uint32), rv)


def test_radix_sort_device_count():
  orc = oracle()
  rng = np.random.default_rng(5)
  n, live = 9000, 5555
  keys = rng.integers(0, 2 ** 20, n, dtype=np.uint64).astype(np.uint32)
  vals = np.arange(n, dtype=np.uint32)
  n_dev = torch.tensor([live], device="cuda", dtype=torch.int32)
  k, v = ops.radix_sort_pairs(cu(keys.view(np.int32)), cu(vals.view(np.int32)), 0, 20, n_dev=n_dev)
  rk, rv = orc.radix_sort_pairs(keys[:live], vals[:live], 0, 20)
  assert np.array_equal(npy(k).view(np.uint32)[:live], rk)
  assert np.array_equal(npy(v).view(np.uint32)[:live], rv)


def test_project_counts_feed_tile_map():
  """gs_project_fwd's tile info carries gs_tile_count's counts, and gs_tile_map(tile_info_in) gives the identical map."""
  W, H = 200, 150
  cam = random_camera(W, H); cfg = go.make_config()
  g = random_gaussians(9000, seed=22, scale=0.1)
  gc = {k: cu(v) for k, v in g.items()}
  lcfg = to_lib_config(cfg)
  info = ops.TileInfo(9000, W, H, "cuda")
  info.buffer.fill_(0xff)
  g2d, depth, _ = ops.project_fwd(to_lib_camera(cam), lcfg, gc["position"], gc["log_scaling"], gc["rotation"],
                                  gc["alpha_logit"], gc["sh"], 3, tile_info=info)
  counts = info.counts
  ref_counts, counters = ops.tile_count(lcfg, g2d, depth, W, H)
  assert torch.equal(counts, ref_counts)
  a = ops.tile_map(lcfg, g2d, depth, W, H)
  b = ops.tile_map(lcfg, g2d, depth, W, H, tile_info=info)
  total = int(a[2][0].item())
  assert torch.equal(a[0][:total], b[0][:total]) and torch.equal(a[1], b[1]) and torch.equal(a[2], b[2])
  assert int(a[2][3].item()) == int((counts > 0).sum().item())
