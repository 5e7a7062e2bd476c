"""Oracle of the per-view cluster features and the frustum visibility counts (SURVEY 8f-3 remainder) pinned to vectors
produced by the reference's own PointClusters.view_features / point_visibility (tests/golden/make_golden_visibility.py)."""
import os

import numpy as np

from oracle import gs_oracle as go

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "visibility.npz")


def camera_rows(g):
  rows = []
  for T, (fx, fy, cx, cy) in zip(g["cam_T"], g["cam_intrinsics"]):
    K = np.zeros((4, 4), np.float32)
    K[0, 0], K[1, 1], K[0, 2], K[1, 2], K[2, 2], K[3, 3] = fx, fy, cx, cy, 1, 1
    rows.append(np.concatenate([(K @ T.astype(np.float32)).reshape(-1), [g["W"], g["H"], g["near"], g["far"]]]).astype(np.float32))
  return np.stack(rows)


def test_view_features_match_reference():
  g = np.load(GOLD)
  orc = go.Oracle("f32")
  f = orc.view_features(g["labels"], int(g["k"]), g["vis"], g["idx"])
  np.testing.assert_allclose(f, g["features"], rtol=1e-5, atol=1e-6)  # the reference's scatter_add order is unspecified
  f5 = orc.view_features(g["labels"], int(g["k"]), g["vis"], g["idx"], threshold=0.5)
  np.testing.assert_allclose(f5, g["features_thr05"], rtol=1e-5, atol=1e-6)
  # the same entries given as a dense visibility plane
  plane = np.zeros(g["labels"].shape[0], np.float32)
  plane[g["idx"]] = g["vis"]
  np.testing.assert_allclose(orc.view_features(g["labels"], int(g["k"]), plane), g["features"], rtol=1e-5, atol=1e-6)


def test_point_visibility_matches_reference():
  g = np.load(GOLD)
  counts = go.Oracle("f32").point_visibility(g["points"], camera_rows(g))
  assert (counts != g["counts"]).mean() < 1e-3  # a point on a frustum face may flip with the matmul's summation order
  assert np.abs(counts - g["counts"]).max() <= 1
