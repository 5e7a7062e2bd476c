"""GPU: libgsb200 against the golden vectors produced by the reference's own code (tests/golden/)."""
import numpy as np
import pytest
import torch

from gsb200 import ops
from gsb200.scene import SceneConfig, SplatScene
from gpu_util import cu, npy
from test_golden_oracle import gold, stats_inputs

pytestmark = pytest.mark.gpu


def test_sh_eval_matches_rsh_cart_3():
  """colour - 0.5 with one-hot coefficients isolates each basis function of the CUDA SH evaluation."""
  g = gold("sh_basis.npz")
  dirs = g["dirs"]
  n = dirs.shape[0]
  pos = cu(dirs * 2.0)  # camera at the origin -> direction = dirs
  for k in range(16):
    sh = np.zeros((n, 3, 16), np.float32)
    sh[:, 0, k] = 1.0
    rgb = npy(ops.sh_eval_fwd(None, cu(sh), 3, pos, np.zeros(3, np.float32)))
    np.testing.assert_allclose(rgb[:, 0] - 0.5, g["Y"][:, k], rtol=1e-5, atol=2e-6, err_msg=f"Y{k}")
    assert np.all(rgb[:, 1:] == 0.5)


def test_take_n_matches_reference():
  g = gold("take_n.npz")
  v = cu(g["values"])
  for k in (0, 1, 37, 2000, 4000):
    for desc in (0, 1):
      assert np.array_equal(npy(ops.select_topk_mask(v, k, desc)), g[f"k{k}_d{desc}"])


def test_split_matches_reference():
  g = gold("split.npz")
  n = g["in_position"].shape[0]
  idx = torch.arange(n, device="cuda", dtype=torch.int64)
  for mode, uniform in (("argmax", None), ("random", g["uniform"])):
    op = torch.zeros((2 * n, 3), device="cuda"); ols = torch.zeros((2 * n, 3), device="cuda")
    ops.split_positions(idx, cu(g["in_position"]), cu(g["in_log_scaling"]), cu(g["in_rotation"]),
                        None if uniform is None else cu(uniform), 0.7, op, ols)
    np.testing.assert_allclose(npy(op), g[f"{mode}_position"], rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(npy(ols), g[f"{mode}_log_scaling"], rtol=1e-6, atol=1e-6)
    dst = torch.zeros((2 * n, 48), device="cuda")
    ops.gather_rows(cu(g["in_feature"]), dst, idx, repeat=2)
    assert np.array_equal(npy(dst), g[f"{mode}_feature"])


def test_point_state_add_rendering_matches_reference():
  g = gold("point_state.npz")
  g2d, heur, state = stats_inputs(g)
  dev = {k: cu(v) for k, v in state.items()}
  ops.accumulate_stats(cu(g["idx"]), cu(g2d), cu(g["vis"]), cu(heur), 0.01, 0.1, dev)
  assert np.array_equal(npy(dev["points_in_view"]), g["after_points_in_view"])
  for k in ("prune_cost", "split_score", "visibility"):
    np.testing.assert_allclose(npy(dev[k]), g[f"after_{k}"], rtol=5e-6, atol=5e-6, err_msg=k)
  np.testing.assert_allclose(npy(dev["max_scale_px"]), g["after_max_scale_px"], rtol=1e-4)


def test_find_split_prune_and_densify_match_reference():
  """SplatScene.find_split_prune == TargetController.find_split_prune_indexes on the reference's own run, and
  densify() keeps the count invariant N' = target_points."""
  g = gold("find_split_prune.npz")
  n = g["in_prune_cost"].shape[0]
  rng = np.random.default_rng(0)
  gs = dict(position=rng.normal(0, 1, (n, 3)), log_scaling=rng.normal(-2, 0.5, (n, 3)), rotation=rng.normal(0, 1, (n, 4)),
            alpha_logit=rng.normal(0, 1, n), sh=rng.normal(0, 1, (n, 3, 16)))
  scene = SplatScene({k: v.astype(np.float32) for k, v in gs.items()}, SceneConfig(sh_degree=3))
  for k in scene.stats:
    scene.stats[k].copy_(cu(g[f"in_{k}"]))
  split, prune = scene.find_split_prune(float(g["t"]), int(g["target_points"]))
  assert np.array_equal(npy(prune), g["prune_mask"])
  assert np.array_equal(npy(split), g["split_mask"])
  for k in scene.stats:
    scene.stats[k].copy_(cu(g[f"in_{k}"]))
  before = {k: npy(v).copy() for k, v in scene.p.items()}
  scene.running_vis.copy_(torch.arange(n, device="cuda", dtype=torch.float32))
  old_n, new_n = scene.densify(float(g["t"]), int(g["target_points"]))
  assert old_n == n and new_n == int(g["target_points"]) == scene.num_points
  keep = ~(g["split_mask"] | g["prune_mask"])
  kept = int(keep.sum())
  for k in before:
    assert np.array_equal(npy(scene.p[k])[:kept], before[k][keep]), k
  assert np.array_equal(npy(scene.running_vis)[:kept], np.arange(n, dtype=np.float32)[keep])
  assert np.all(npy(scene.running_vis)[kept:] == 0)
  # children duplicate the parent's rotation / opacity / SH (repeat_interleave, gaussians/split.py:42-44)
  sidx = np.nonzero(g["split_mask"])[0]
  assert np.array_equal(npy(scene.p["rotation"])[kept:], np.repeat(before["rotation"][sidx], 2, axis=0))
  assert np.array_equal(npy(scene.p["sh"])[kept:], np.repeat(before["sh"][sidx], 2, axis=0))
  assert all(int(v.abs().sum().item()) == 0 for v in scene.stats.values())
