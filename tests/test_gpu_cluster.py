"""GPU parity of the point clustering (include/gsb200.h section 10) against the CPU oracle and the golden vectors made
by the reference's own kmeans_iter / assign_clusters (tests/golden/make_golden_kmeans.py)."""
import os

import numpy as np
import pytest
import torch

from gsb200 import lib as L
from gsb200 import ops
from gpu_util import cu, npy
from oracle import gs_oracle as go

pytestmark = pytest.mark.gpu
G = np.load(os.path.join(os.path.dirname(__file__), "golden", "kmeans.npz"))


def test_assign_clusters_bit_exact_vs_reference_golden():
  labels = ops.assign_clusters(cu(G["probe"]), cu(G["probe_centroids"]))
  assert np.array_equal(npy(labels), G["probe_labels"])


def test_kmeans_matches_reference_golden():
  x, init = cu(G["x"]), cu(G["init"])
  labels, cent = ops.kmeans(x, iters=1, centroids=init)
  assert np.array_equal(npy(labels), G["labels_1"]), "same centroids in: identical assignment"
  np.testing.assert_allclose(npy(cent), G["centroids_1"], rtol=1e-5, atol=1e-5)
  # later iterations: the centroid sums are fp32 reductions in arbitrary order (as the reference's scatter_add_ on CUDA),
  # so a point on a cell boundary may flip: bound the flips, compare centroids loosely
  labels, cent = ops.kmeans(x, iters=10, centroids=init)
  assert (npy(labels) != G["labels_10"]).mean() < 2e-3
  np.testing.assert_allclose(npy(cent), G["centroids_10"], rtol=1e-3, atol=1e-3)


@pytest.mark.parametrize("n,d,k", [(1, 3, 1), (1000, 3, 1), (5003, 3, 64), (70001, 3, 1024), (4099, 2, 17), (3000, 8, 33)])
def test_assign_bit_exact_vs_oracle(n, d, k):
  rng = np.random.default_rng(n + d + k)
  x = rng.normal(0, 3, (n, d)).astype(np.float32)
  c = rng.normal(0, 3, (k, d)).astype(np.float32)
  if k > 4:
    c[3] = c[1]          # duplicate centroid: the lower index wins
    c[2] = np.nan        # an emptied cluster is never nearest
  if n > 10:
    x[7] = c[0]          # zero distance
  labels = npy(ops.assign_clusters(cu(x), cu(c)))
  ref = go.Oracle("f32").assign_clusters(x, c)
  assert np.array_equal(labels, ref)
  if k > 4:
    assert not np.isin(labels, (2, 3)).any()


def test_kmeans_properties_at_scale_and_empty_clusters():
  gen = torch.Generator(device="cuda").manual_seed(0)
  n, k = 300_000, 256
  x = torch.randn((n, 3), device="cuda", generator=gen) * 5
  init = x[:k].clone()
  init[9] = init[4]  # cluster 9 empties in the first round
  labels, cent = ops.kmeans(x, iters=5, centroids=init)
  assert torch.isnan(cent[9]).all() and not (labels == 9).any()
  live = [j for j in range(k) if j != 9]
  assert torch.isfinite(cent[live]).all()
  # fixed-point property of one more round: centroids == mean of the points assigned to them
  l2 = ops.assign_clusters(x, cent)
  counts = torch.bincount(l2.long(), minlength=k).float()
  sums = torch.zeros((k, 3), device="cuda").index_add_(0, l2.long(), x)
  l3, c3 = ops.kmeans(x, iters=1, centroids=cent)
  assert torch.equal(l3, l2)
  torch.testing.assert_close(c3[live], (sums / counts[:, None])[live], rtol=1e-4, atol=1e-4)
  # inertia does not increase from round to round
  def inertia(c, l):
    return ((x - c[l.long()]) ** 2).sum().item()
  la, ca = ops.kmeans(x, iters=1, centroids=init)
  lb, cb = ops.kmeans(x, iters=4, centroids=init)
  assert inertia(cb, ops.assign_clusters(x, cb)) <= inertia(ca, ops.assign_clusters(x, ca)) * (1 + 1e-6)


def test_bad_arguments_are_refused():
  x = torch.zeros((10, 9), device="cuda")
  with pytest.raises(L.GsError):
    ops.assign_clusters(x, torch.zeros((4, 9), device="cuda"))


def test_estimate_scale_matches_reference_golden():
  for nn in (3, 5):
    ours = npy(ops.estimate_scale(cu(G["knn_points"]), nn))
    np.testing.assert_allclose(ours, G[f"knn_scale_{nn}"], rtol=2e-6, atol=1e-7)


@pytest.mark.parametrize("n,kind", [(2, "uniform"), (5, "uniform"), (4000, "uniform"), (6000, "clumps"), (3000, "plane"),
                                    (2500, "outliers"), (3000, "duplicates")])
def test_estimate_scale_bit_exact_vs_brute_force_oracle(n, kind):
  rng = np.random.default_rng(n)
  if kind == "uniform":
    p = rng.uniform(-3, 3, (n, 3))
  elif kind == "clumps":
    p = rng.normal(0, 4, (10, 3))[rng.integers(0, 10, n)] + rng.normal(0, 0.05, (n, 3))
  elif kind == "plane":
    p = np.concatenate([rng.uniform(-5, 5, (n, 2)), np.zeros((n, 1))], axis=1)
  elif kind == "outliers":
    p = rng.normal(0, 1, (n, 3)); p[:5] = rng.uniform(200, 400, (5, 3)) * np.array([1, -1, 1])
  else:
    p = rng.uniform(-2, 2, (n, 3)); p[n // 2:] = p[:n - n // 2]
  p = p.astype(np.float32)
  ours = npy(ops.estimate_scale(cu(p), 3))
  ref = go.Oracle("f32").knn_scale(p, 3)
  # the k+1 smallest squared distances are the same multiset; sqrt/mean on the GPU are correctly rounded like libm's
  np.testing.assert_allclose(ours, ref, rtol=3e-7, atol=0)
  if n > 10:
    assert (ours == ref).mean() > 0.99


def test_estimate_scale_at_scale_sanity():
  gen = torch.Generator(device="cuda").manual_seed(1)
  n = 1_000_000
  p = torch.rand((n, 3), device="cuda", generator=gen)
  s = ops.estimate_scale(p, 3)
  # uniform density n per unit volume: mean distance to the 3 nearest neighbours ~ 0.74 * n^(-1/3)
  assert torch.isfinite(s).all()
  assert 0.6 < s.mean().item() * n ** (1 / 3) < 0.9
  sub = torch.arange(0, n, 5000, device="cuda")
  d = torch.cdist(p[sub], p, compute_mode="donot_use_mm_for_euclid_dist").topk(4, dim=1, largest=False).values[:, 1:].mean(dim=1)
  torch.testing.assert_close(s[sub], d, rtol=1e-4, atol=1e-6)
