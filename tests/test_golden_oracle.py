"""CPU: the oracle's PINNED parts against golden vectors produced by running the reference's own code
(tests/golden/make_golden.py)."""
import os

import numpy as np
import pytest

from oracle import gs_oracle as go

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def gold(name):
  return np.load(os.path.join(GOLD, name))


@pytest.fixture(scope="module")
def orc():
  return go.Oracle("f32")


def test_sh_basis_matches_rsh_cart_3(orc):
  g = gold("sh_basis.npz")
  Y = orc.sh_basis(3, g["dirs"])
  np.testing.assert_allclose(Y, g["Y"], rtol=1e-5, atol=1e-6)
  # known answers on the axes: Y1 = -c1*y, Y2 = c1*z, Y3 = -c1*x
  c1 = 0.48860251190292
  np.testing.assert_allclose(Y[0, 1:4], [0, 0, -c1], atol=1e-7)
  np.testing.assert_allclose(Y[1, 1:4], [-c1, 0, 0], atol=1e-7)
  np.testing.assert_allclose(Y[2, 1:4], [0, c1, 0], atol=1e-7)


def test_sh_colour_dc_is_rgb_to_sh_inverse(orc):
  """colour = 0.5 + sh0*c (util/misc.py:41-49): evaluating the oracle with DC = rgb_to_sh(rgb) returns rgb."""
  g = gold("sh_basis.npz")
  rgb = g["rgb"]
  sh = np.zeros((rgb.shape[0], 3, 16), np.float32)
  sh[:, :, 0] = g["rgb_to_sh"]
  pos = np.random.default_rng(0).normal(0, 1, (rgb.shape[0], 3)).astype(np.float32)
  out = orc.sh_eval_fwd(None, sh, 3, pos, np.zeros(3, np.float32))
  np.testing.assert_allclose(out, rgb, atol=1e-6)
  assert abs(float(g["sh0"]) - 0.282094791773878) < 1e-7


def test_exp_lerp_matches_reference(orc):
  g = gold("exp_lerp.npz")
  for t in (0.0, 0.01, 0.1, 0.5, 1.0):
    ours = np.array([orc.exp_lerp(t, float(a), float(b)) for a, b in zip(g["a"], g["b"])], np.float32)
    np.testing.assert_allclose(ours, g[f"t{t}"], rtol=2e-6, atol=2e-6)
  np.testing.assert_allclose(g["t0.5"][:4], g["a"][:4], rtol=1e-6)  # exp_lerp(t, a, a) = a


def test_take_n_matches_reference(orc):
  g = gold("take_n.npz")
  for k in (0, 1, 37, 2000, 4000):
    for desc in (0, 1):
      ref = g[f"k{k}_d{desc}"]
      assert ref.sum() == k
      assert np.array_equal(orc.select_topk_mask(g["values"], k, desc), ref)


def test_split_matches_reference(orc):
  g = gold("split.npz")
  n = g["in_position"].shape[0]
  B = orc.point_basis(g["in_log_scaling"], g["in_rotation"])
  np.testing.assert_allclose(B, g["basis"], rtol=1e-5, atol=1e-7)
  idx = np.arange(n, dtype=np.int64)
  for mode, uniform in (("argmax", None), ("random", g["uniform"])):
    pos, ls = orc.split_positions(idx, g["in_position"], g["in_log_scaling"], g["in_rotation"], uniform, 0.7)
    np.testing.assert_allclose(pos, g[f"{mode}_position"], rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(ls, g[f"{mode}_log_scaling"], rtol=1e-6, atol=1e-6)
    # the other attributes are repeat_interleave'd unchanged
    for k in ("rotation", "alpha_logit", "feature"):
      assert np.array_equal(g[f"{mode}_{k}"], np.repeat(g[f"in_{k}"], 2, axis=0))


def stats_inputs(g):
  """projected rows whose larger std-dev equals the reference's screen_scale.max(1) (axis-aligned conic)."""
  sc = g["screen_scale"].astype(np.float64)
  m = sc.shape[0]
  g2d = np.zeros((m, 6), np.float32)
  g2d[:, 2] = 1.0 / sc[:, 0] ** 2
  g2d[:, 4] = 1.0 / sc[:, 1] ** 2
  g2d[:, 5] = 0.5
  heur = np.stack([g["prune_cost"], g["split_score"]], axis=1).astype(np.float32)
  state = {k: g[f"before_{k}"].copy() for k in ("prune_cost", "split_score", "max_scale_px", "points_in_view", "visibility")}
  return g2d, heur, state


def test_point_state_add_rendering_matches_reference(orc):
  g = gold("point_state.npz")
  g2d, heur, state = stats_inputs(g)
  orc.accumulate_stats(g["idx"], g2d, g["vis"], heur, 0.01, 0.1, state)
  assert np.array_equal(state["points_in_view"], g["after_points_in_view"])
  for k in ("prune_cost", "split_score", "visibility"):
    np.testing.assert_allclose(state[k], g[f"after_{k}"], rtol=3e-6, atol=3e-6, err_msg=k)
  # the scale is re-derived from the conic (eigenvalue of its inverse): a few fp32 ulps of cancellation
  np.testing.assert_allclose(state["max_scale_px"], g["after_max_scale_px"], rtol=1e-4)


def find_split_prune_oracle(orc, stats, t, target_points, prune_rate=0.025, min_views=5, max_scale_px=200.0):
  """TargetController.find_split_prune_indexes (reference controller/target_controller.py:73-96) on top of the
  oracle's take_n — the host logic SplatScene.find_split_prune mirrors."""
  n = stats["prune_cost"].shape[0]
  exceeds = stats["max_scale_px"] > max_scale_px
  n_prune = int(np.ceil(prune_rate * n * (1 - t)))
  seen = stats["points_in_view"] >= min_views
  prune_cost = np.where(seen, stats["prune_cost"], np.inf).astype(np.float32)
  prune_mask = orc.select_topk_mask(prune_cost, n_prune, 0) | exceeds
  target_split = (target_points - n) + int(prune_mask.sum())
  split_score = stats["split_score"].copy()
  split_score[prune_mask] = 0.0
  split_mask = orc.select_topk_mask(split_score, max(target_split, 0), 1)
  both = split_mask & prune_mask
  return split_mask ^ both, prune_mask ^ both


def test_find_split_prune_matches_reference(orc):
  g = gold("find_split_prune.npz")
  stats = {k: g[f"in_{k}"] for k in ("prune_cost", "split_score", "max_scale_px", "points_in_view", "visibility")}
  split, prune = find_split_prune_oracle(orc, stats, float(g["t"]), int(g["target_points"]))
  assert np.array_equal(prune, g["prune_mask"])
  assert np.array_equal(split, g["split_mask"])
  n = stats["prune_cost"].shape[0]
  assert n - prune.sum() + split.sum() == int(g["target_points"])  # densify count invariant


def test_mcmc_noise_matches_reference(orc):
  """controller/mcmc_controller.py:91-100 run with the reference's own soft_lt / sample_gaussians on recorded normals."""
  g = gold("mcmc.npz")
  pos = g["position"].copy()
  orc.mcmc_noise(g["alpha_logit"][:, 0], g["log_scaling"], g["rotation"], g["points_in_view"], int(g["min_views"]),
                 float(g["opacity_threshold"]), float(g["noise_level"]), g["randn"], pos)
  moved = g["points_in_view"] > int(g["min_views"])
  assert moved.sum() > 100 and (~moved).sum() > 100
  assert np.array_equal(pos[~moved], g["position"][~moved])
  np.testing.assert_allclose(pos, g["out_position"], rtol=2e-5, atol=2e-6)
  assert np.abs(pos[moved] - g["position"][moved]).max() > 1e-3


def test_mcmc_prune_respawn_masks_match_reference(orc):
  """controller/mcmc_controller.py:76-89: prune mask and the take_n respawn mask (same count)."""
  g = gold("mcmc.npz")
  op = 1.0 / (1.0 + np.exp(-g["alpha_logit"][:, 0].astype(np.float64)))
  prune = (g["max_scale_px"] > 200.0) | (op.astype(np.float32) < float(g["opacity_threshold"]))
  assert np.array_equal(prune, g["prune_mask"])
  score = np.where(prune | (g["max_scale_px"] < float(g["min_split_px"])), 0.0, g["split_score"]).astype(np.float32)
  split = orc.select_topk_mask(score, int(prune.sum()), descending=1)
  assert np.array_equal(split, g["split_mask"])


def test_reg_loss_matches_reference(orc):
  """scene/mlp_scene.py:246-265 (compute_reg) and its autograd gradients with the yaml weights."""
  g = gold("reg_loss.npz")
  m = g["opacity"].shape[0]
  g2d = np.zeros((m, 6), np.float32); g2d[:, 5] = g["opacity"]
  w = (float(g["w_scale"]), float(g["w_opacity"]), float(g["w_aspect"]))
  d_op, d_dep, d_ls = np.zeros(m, np.float32), np.zeros(m, np.float32), np.zeros((m, 3), np.float32)
  sums = orc.reg_loss(w, g2d, g["depths"][:, 0], g["log_scaling"], g["weight"], d_op, d_dep, d_ls)
  assert sums[3] == m
  np.testing.assert_allclose(sums[0] / m, g["reg_scale"], rtol=1e-5)
  np.testing.assert_allclose(sums[1] / m, g["reg_opacity"], rtol=1e-5)
  np.testing.assert_allclose(sums[2] / m, g["reg_aspect"], rtol=1e-5)
  total = (w[0] * sums[0] + w[1] * sums[1] + w[2] * sums[2]) / m
  np.testing.assert_allclose(total, g["total"], rtol=1e-5)
  np.testing.assert_allclose(d_op, g["d_opacity"], rtol=2e-4, atol=1e-9)
  np.testing.assert_allclose(d_dep, g["d_depths"][:, 0], rtol=2e-4, atol=1e-9)
  np.testing.assert_allclose(d_ls, g["d_log_scaling"], rtol=2e-4, atol=1e-9)
  # rows no pixel blended (vis == 0) are outside the mean and get no gradient
  vis = g["weight"].copy(); vis[::3] = 0
  d2 = np.zeros(m, np.float32)
  s2 = orc.reg_loss(w, g2d, g["depths"][:, 0], g["log_scaling"], vis, d2, None, None)
  assert s2[3] == (vis > 0).sum() and np.all(d2[::3] == 0)
