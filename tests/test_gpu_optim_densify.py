"""GPU parity: optimiser, statistics, top-k selection, compaction and split vs the CPU oracle."""
import numpy as np
import pytest
import torch

from gsb200 import lib as L
from gsb200 import ops
from gpu_util import cu, npy, oracle, to_lib_optim
from oracle import gs_oracle as go

pytestmark = pytest.mark.gpu


def _quat(rng, n):
  q = rng.normal(0, 1, (n, 4)).astype(np.float32)
  return q


@pytest.mark.parametrize("indexed", [False, True])
@pytest.mark.parametrize("mode", [0, 1])
@pytest.mark.parametrize("kind,dim,post", [(0, 1, 0), (0, 3, 2), (1, 4, 1), (1, 48, 0), (1, 16, 0), (2, 3, 0), (0, 48, 0)])
def test_optim_step_matches_oracle(mode, kind, dim, post, indexed):
  orc = oracle()
  rng = np.random.default_rng(100 + 10 * kind + dim)
  n = 3001
  ocfg = go.make_optim_config(mode=mode)
  param = rng.normal(0, 1, (n, dim)).astype(np.float32)
  visible = np.where(rng.uniform(0, 1, n) < 0.6, rng.uniform(0.01, 30, n), 0).astype(np.float32)
  running = rng.uniform(0.0, 5, n).astype(np.float32)
  step = rng.integers(0, 50, n).astype(np.float32)
  exp_avg = rng.normal(0, 0.1, (n, dim)).astype(np.float32)
  exp_avg_sq = rng.uniform(0, 0.1, (n, dim) if kind == 0 else (n,)).astype(np.float32)
  bls = rng.normal(-2, 1, (n, 3)).astype(np.float32); bq = _quat(rng, n)
  state = dict(param=param, exp_avg=exp_avg, exp_avg_sq=exp_avg_sq, running=running, step=step)
  ref = {k: v.copy() for k, v in state.items()}
  dev = {k: cu(v) for k, v in state.items()}
  vis_c = cu(visible)
  for it in range(3):
    grad = (rng.normal(0, 1, (n, dim)) * rng.choice([1e-3, 1.0, 50.0], (n, 1))).astype(np.float32)
    rgrad = grad.copy(); dgrad = cu(grad)
    orc.optim_vis_update(ocfg, visible, ref["running"], ref["step"])
    orc.optim_step(ocfg, kind, 0.05, ref["param"], rgrad, ref["exp_avg"], ref["exp_avg_sq"], visible, ref["running"],
                   ref["step"], bls, bq, post, -0.5, 0.5, 1)
    locfg = to_lib_optim(ocfg)
    ops.optim_vis_update(locfg, vis_c, dev["running"], dev["step"])
    ix = ops.VisibleIndexes(n, "cuda").update(vis_c) if indexed else None
    if indexed:
      assert int(ix.count.item()) == (visible > 0).sum()
      assert np.array_equal(npy(ix.indexes)[:int(ix.count.item())], np.nonzero(visible > 0)[0])
    ops.optim_step(locfg, kind, 0.05, dev["param"], dgrad, dev["exp_avg"], dev["exp_avg_sq"], vis_c, dev["running"],
                   dev["step"], cu(bls), cu(bq), post, -0.5, 0.5, True, indexes=ix)
    for k in state:
      np.testing.assert_allclose(npy(dev[k]), ref[k], rtol=2e-4, atol=2e-6, err_msg=f"{k} it{it}")
    # consumed gradient rows are cleared, untouched rows keep theirs
    g = npy(dgrad)
    assert np.all(g[visible > 0] == 0) and np.array_equal(g[visible <= 0], grad[visible <= 0])
    assert np.array_equal(g, rgrad)
  untouched = visible <= 0
  assert np.array_equal(npy(dev["param"])[untouched], param[untouched])


@pytest.mark.parametrize("indexed", [False, True])
def test_grouped_launch_equals_one_launch_per_group(indexed):
  """gs_optim_step_groups (the training step's two launches) == gs_optim_step per group, bit for bit; a call whose
  groups would race on a basis plane is refused."""
  rng = np.random.default_rng(7)
  n = 20011
  locfg = to_lib_optim(go.make_optim_config(mode=1))
  dims = dict(position=3, log_scaling=3, rotation=4, alpha_logit=1, sh=48)
  kinds = dict(position=2, log_scaling=0, rotation=1, alpha_logit=0, sh=1)
  posts = dict(log_scaling=(2, -8.0, 8.0), rotation=(1, 0.0, 0.0))
  host = {}
  for name, d in dims.items():
    host[name] = dict(param=rng.normal(0, 1, (n, d)).astype(np.float32), grad=rng.normal(0, 1, (n, d)).astype(np.float32),
                      m1=rng.normal(0, 0.1, (n, d)).astype(np.float32),
                      m2=rng.uniform(0, 0.1, (n, d) if kinds[name] == 0 else (n,)).astype(np.float32))
  vis = cu(np.where(rng.uniform(0, 1, n) < 0.3, rng.uniform(0.01, 30, n), 0).astype(np.float32))
  rv, st = cu(rng.uniform(0.1, 5, n).astype(np.float32)), cu(rng.integers(1, 50, n).astype(np.float32))
  ix = ops.VisibleIndexes(n, "cuda").update(vis) if indexed else None
  A = {k: {q: cu(v) for q, v in d.items()} for k, d in host.items()}
  B = {k: {q: cu(v) for q, v in d.items()} for k, d in host.items()}
  order = ("position", "log_scaling", "rotation", "alpha_logit", "sh")
  for name in order:  # reference order, one launch per group (position reads the basis before its update)
    a = A[name]
    post = posts.get(name, (0, 0.0, 0.0))
    ops.optim_step(locfg, kinds[name], 0.01, a["param"], a["grad"], a["m1"], a["m2"], vis, rv, st,
                   A["log_scaling"]["param"], A["rotation"]["param"], *post, True, indexes=ix)

  def G(name):
    b = B[name]
    post = posts.get(name, (0, 0.0, 0.0))
    basis = (B["log_scaling"]["param"], B["rotation"]["param"]) if name == "position" else (None, None)
    return ops.optim_group(kinds[name], 0.01, b["param"], b["grad"], b["m1"], b["m2"], *basis, *post)

  ops.optim_step_groups(locfg, n, [G("position"), G("alpha_logit"), G("sh")], vis, rv, st, indexes=ix)
  ops.optim_step_groups(locfg, n, [G("log_scaling"), G("rotation")], vis, rv, st, indexes=ix)
  torch.cuda.synchronize()
  for name in order:
    for q in ("param", "grad", "m1", "m2"):
      assert torch.equal(A[name][q], B[name][q]), (name, q)
  with pytest.raises(L.GsError):
    ops.optim_step_groups(locfg, n, [G("position"), G("rotation")], vis, rv, st, indexes=ix)


def test_accumulate_stats_matches_oracle():
  orc = oracle()
  rng = np.random.default_rng(7)
  N, m = 5000, 1800
  idx = np.sort(rng.choice(N, m, replace=False)).astype(np.int64)
  g2d = np.zeros((m, 6), np.float32)
  g2d[:, 2] = rng.uniform(0.01, 1, m); g2d[:, 4] = rng.uniform(0.01, 1, m)
  g2d[:, 3] = rng.uniform(-0.9, 0.9, m) * np.sqrt(g2d[:, 2] * g2d[:, 4])
  g2d[:, 5] = np.where(rng.uniform(0, 1, m) < 0.8, rng.uniform(0.01, 1, m), 0)
  vis = np.where(rng.uniform(0, 1, m) < 0.7, rng.uniform(0, 20, m), 0).astype(np.float32)
  heur = rng.uniform(0, 3, (m, 2)).astype(np.float32)
  mk = lambda: dict(prune_cost=rng.uniform(0, 1, N).astype(np.float32), split_score=rng.uniform(0, 1, N).astype(np.float32),
                    max_scale_px=rng.uniform(0, 3, N).astype(np.float32), points_in_view=rng.integers(0, 9, N).astype(np.int16),
                    visibility=rng.uniform(0, 1, N).astype(np.float32), scene_visible=rng.uniform(0, 1, N).astype(np.float32))
  ref = mk()
  dev = {k: cu(v) for k, v in ref.items()}
  for use_idx in (True, False):
    if use_idx:
      orc.accumulate_stats(idx, g2d, vis, heur, 0.01, 0.1, ref)
      ops.accumulate_stats(cu(idx), cu(g2d), cu(vis), cu(heur), 0.01, 0.1, dev, scene_visible=dev["scene_visible"])
    else:  # fused path: rows are Gaussians
      sub = {k: v[:m] for k, v in ref.items()}
      orc.accumulate_stats(None, g2d, vis, heur, 0.01, 0.1, sub)
      dsub = {k: v[:m] for k, v in dev.items()}
      ops.accumulate_stats(None, cu(g2d), cu(vis), cu(heur), 0.01, 0.1, dsub, scene_visible=dsub["scene_visible"])
    assert np.array_equal(npy(dev["points_in_view"]), ref["points_in_view"])
    for k in ("prune_cost", "split_score", "max_scale_px", "visibility", "scene_visible"):
      np.testing.assert_allclose(npy(dev[k]), ref[k], rtol=1e-5, atol=1e-6, err_msg=k)


@pytest.mark.parametrize("n,k", [(1, 1), (1000, 0), (1000, 1000), (1000, 37), (100003, 2500), (100003, 99999)])
@pytest.mark.parametrize("descending", [0, 1])
def test_select_topk_mask_bit_exact(n, k, descending):
  orc = oracle()
  rng = np.random.default_rng(n + k)
  v = rng.normal(0, 1, n).astype(np.float32)
  if n > 100:
    v[rng.choice(n, n // 4, replace=False)] = np.float32(0.25)  # heavy ties at one value
    v[rng.choice(n, 5, replace=False)] = np.nan
    v[rng.choice(n, 5, replace=False)] = -0.0
    v[rng.choice(n, 5, replace=False)] = 0.0
    v[rng.choice(n, 3, replace=False)] = np.inf
    v[rng.choice(n, 3, replace=False)] = -np.inf
  mask = npy(ops.select_topk_mask(cu(v), k, descending))
  ref = orc.select_topk_mask(v, k, descending)
  assert mask.sum() == min(k, n)
  assert np.array_equal(mask, ref)


def test_compaction_roundtrip_and_oracle():
  orc = oracle()
  rng = np.random.default_rng(9)
  n = 70001
  keep = rng.uniform(0, 1, n) < 0.63
  slots, count = ops.compact_slots(cu(keep))
  rslots, rcount = orc.compact_slots(keep)
  assert int(count.item()) == rcount == keep.sum()
  assert np.array_equal(npy(slots), rslots)
  for shape, dt in [((n, 59), np.float32), ((n,), np.int16), ((n, 3), np.uint8), ((n, 4), np.float32)]:
    src = rng.integers(0, 250, shape).astype(dt)
    dst = torch.zeros((rcount,) + shape[1:], device="cuda", dtype=cu(src).dtype)
    ops.compact_rows(cu(src), slots, dst)
    assert np.array_equal(npy(dst), src[keep])
    assert np.array_equal(npy(dst), orc.compact_rows(src, rslots, rcount))


def test_split_positions_and_gather_match_oracle():
  orc = oracle()
  rng = np.random.default_rng(11)
  n, s = 4000, 900
  pos = rng.normal(0, 1, (n, 3)).astype(np.float32)
  ls = rng.normal(-2, 1, (n, 3)).astype(np.float32)
  ls[:50] = -12.0  # below the 1e-4 clamp
  q = _quat(rng, n)
  idx = rng.choice(n, s, replace=False).astype(np.int64)
  u = rng.uniform(0, 1, s).astype(np.float32)
  for uniform in (u, None):
    rop, rols = orc.split_positions(idx, pos, ls, q, uniform, 0.7)
    op = torch.zeros((5 + 2 * s, 3), device="cuda"); ols = torch.zeros((5 + 2 * s, 3), device="cuda")
    ops.split_positions(cu(idx), cu(pos), cu(ls), cu(q), None if uniform is None else cu(uniform), 0.7, op, ols,
                        out_offset=5)
    assert np.all(npy(op)[:5] == 0)
    np.testing.assert_allclose(npy(op)[5:], rop, rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(npy(ols)[5:], rols, rtol=1e-6, atol=1e-6)
  feat = rng.normal(0, 1, (n, 48)).astype(np.float32)
  dst = torch.zeros((3 + 2 * s, 48), device="cuda")
  ops.gather_rows(cu(feat), dst, cu(idx), repeat=2, out_offset=3)
  assert np.array_equal(npy(dst)[3:], np.repeat(feat[idx], 2, axis=0))


def test_sh_eval_matches_oracle():
  orc = oracle()
  rng = np.random.default_rng(12)
  n, v = 2000, 700
  for deg in (0, 1, 2, 3):
    K = (deg + 1) ** 2
    sh = rng.normal(0, 0.5, (n, 3, K)).astype(np.float32)
    pos = rng.normal(0, 2, (n, 3)).astype(np.float32)
    idx = rng.choice(n, v, replace=False).astype(np.int64)
    cam_pos = np.array([0.3, -1.0, 2.0], np.float32)
    rgb = ops.sh_eval_fwd(cu(idx), cu(sh), deg, cu(pos), cam_pos)
    np.testing.assert_allclose(npy(rgb), orc.sh_eval_fwd(idx, sh, deg, pos, cam_pos), rtol=1e-5, atol=1e-5)
    d_rgb = rng.normal(0, 1, (v, 3)).astype(np.float32)
    g_sh, g_pos = ops.sh_eval_bwd(cu(idx), cu(sh), deg, cu(pos), cam_pos, cu(d_rgb))
    rg_sh, rg_pos = orc.sh_eval_bwd(idx, sh, deg, pos, cam_pos, d_rgb)
    np.testing.assert_allclose(npy(g_sh), rg_sh, rtol=1e-5, atol=1e-5)
    np.testing.assert_allclose(npy(g_pos), rg_pos, rtol=1e-3, atol=1e-4)


def test_cpu_tensor_is_rejected():
  with pytest.raises(L.GsError):
    ops.select_topk_mask(torch.zeros(10), 3, 0)
