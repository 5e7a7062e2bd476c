"""Pins the clustering oracle (oracle/gs_oracle.c section 10) to golden vectors produced by the reference's own
kmeans_iter / assign_clusters (tests/golden/make_golden_kmeans.py)."""
import os

import numpy as np

from oracle import gs_oracle as go

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "kmeans.npz"))


def test_kmeans_iterations_match_reference():
  orc = go.Oracle("f32")
  for iters in (1, 2, 10):
    labels, cent = orc.kmeans(G["x"], G["init"], iters)
    assert np.array_equal(labels, G[f"labels_{iters}"]), iters
    np.testing.assert_allclose(cent, G[f"centroids_{iters}"], rtol=1e-5, atol=1e-6)


def test_assign_clusters_matches_reference():
  orc = go.Oracle("f32")
  assert np.array_equal(orc.assign_clusters(G["probe"], G["probe_centroids"]), G["probe_labels"])


def test_emptied_cluster_is_nan_and_never_nearest():
  orc = go.Oracle("f32")
  init = G["init"].copy()
  init[5] = init[2]
  labels, cent = orc.kmeans(G["x"], init, 3)
  assert np.isnan(cent[5]).all() and not (labels == 5).any()
  assert np.isfinite(np.delete(cent, 5, axis=0)).all()


def test_knn_scale_matches_reference_estimate_scale():
  orc = go.Oracle("f32")
  for nn in (3, 5):
    ours = orc.knn_scale(G["knn_points"], nn)
    # the reference's dense expression sums (x-y)^2 over the last axis with torch's reduction; same values up to 1 ulp
    np.testing.assert_allclose(ours, G[f"knn_scale_{nn}"], rtol=2e-6, atol=1e-7)
  assert ours[7] < 1e-3 or ours[100] < ours.mean(), "the duplicated point has a zero-distance neighbour"
