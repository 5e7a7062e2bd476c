"""GPU parity of the camera-batch exchange (include/gsb200.h section 6): packing every image into a record and folding
the records in image order must equal the reference's sequential accumulation (trainer/trainer.py:500-514,
controller/point_state.py:34-50) — bit for bit where the arithmetic is identical, and vs the CPU oracle's sequential
gso_accumulate_stats within the exp/log tolerance.  Ranks are emulated in one process (one GPU): the transport (peer
loads / NCCL) only changes where a record's bytes live."""
import numpy as np
import pytest
import torch

from gsb200 import lib as L
from gsb200 import ops
from gsb200 import synthetic as syn
from gsb200.exchange import ImageExchange
from gsb200.scene import SceneConfig, SplatScene
from gsb200.trainer import Trainer
from gpu_util import cu, npy, oracle

pytestmark = pytest.mark.gpu

DIMS = (3, 3, 4, 1, 12)  # position, log_scaling, rotation, alpha_logit, sh (degree 1)


def _state(n):
  z = lambda dt=torch.float32: torch.zeros((n,), device="cuda", dtype=dt)
  return dict(prune_cost=z(), split_score=z(), max_scale_px=z(), points_in_view=z(torch.int16), visibility=z())


def _fake_image(rng, n, p_view, p_vis):
  """projected rows + raster-backward outputs of one image without rendering anything."""
  inview = rng.uniform(0, 1, n) < p_view
  visible = inview & (rng.uniform(0, 1, n) < p_vis)
  g2d = np.zeros((n, 6), np.float32)
  a = rng.uniform(0.02, 2.0, n); c = rng.uniform(0.02, 2.0, n); b = rng.uniform(-0.9, 0.9, n) * np.sqrt(a * c)
  g2d[:, 0:2] = rng.uniform(0, 300, (n, 2)); g2d[:, 2] = a; g2d[:, 3] = b; g2d[:, 4] = c
  g2d[:, 5] = np.where(inview, rng.uniform(0.05, 1.0, n), 0.0)
  vis = np.where(visible, rng.uniform(1e-3, 40.0, n), 0.0).astype(np.float32)
  heur = np.where(visible[:, None], rng.uniform(0, 5.0, (n, 2)), 0.0).astype(np.float32)
  grads = [np.where(visible[:, None], rng.normal(0, 1, (n, d)), 0.0).astype(np.float32) for d in DIMS]
  return g2d, vis, heur, grads, inview, visible


@pytest.mark.parametrize("n,images,p_view,p_vis,cap", [(5000, 3, 0.6, 0.3, 4096), (70001, 8, 0.5, 0.05, 8192),
                                                       (33, 2, 1.0, 1.0, 64), (4097, 4, 0.5, 0.0, 16)])
def test_pack_fold_equals_sequential_accumulation(n, images, p_view, p_vis, cap):
  rng = np.random.default_rng(n + images)
  orc = oracle()
  sa, pa = 0.01, 0.1
  # --- sequential path: gs_accumulate_stats per image + in-order gradient sums (what one GPU does for a batch)
  seq_state, seq_vis = _state(n), torch.zeros((n,), device="cuda")
  seq_grads = [torch.zeros((n, d), device="cuda") for d in DIMS]
  ref_state = dict(prune_cost=np.zeros(n, np.float32), split_score=np.zeros(n, np.float32), max_scale_px=np.zeros(n, np.float32),
                   points_in_view=np.zeros(n, np.int16), visibility=np.zeros(n, np.float32), scene_visible=np.zeros(n, np.float32))
  # --- exchange path: every image packed by "its rank", all records folded in image order by one replica
  ex_state, ex_vis = _state(n), torch.zeros((n,), device="cuda")
  ex_grads = [torch.zeros((n, d), device="cuda") for d in DIMS]
  groups, row_floats = ops.exchange_groups(ex_grads)
  assert row_floats == sum(DIMS)
  rec_bytes = ops.exchange_record_bytes(n, cap, row_floats)
  records = torch.zeros((images, rec_bytes), device="cuda", dtype=torch.uint8)
  ws = torch.empty((int(L.load().gs_exchange_pack_workspace_bytes(n)),), device="cuda", dtype=torch.uint8)
  rank_max = [torch.zeros((n,), device="cuda") for _ in range(images)]
  counts = []
  for j in range(images):
    g2d, vis, heur, grads, inview, visible = _fake_image(rng, n, p_view, p_vis)
    counts.append(int(visible.sum()))
    g2d_c, vis_c, heur_c = cu(g2d), cu(vis), cu(heur)
    ops.accumulate_stats(None, g2d_c, vis_c, heur_c, sa, pa, seq_state, scene_visible=seq_vis)
    for t, g in zip(seq_grads, grads):
      t += cu(g)
    orc.accumulate_stats(None, g2d, vis, heur, sa, pa, ref_state)
    # this image's rows sit alone in the (otherwise zero) gradient planes, as after project_bwd in write mode
    for t, g in zip(ex_grads, grads):
      t.copy_(cu(g))
    acc = torch.ones((n, 12), device="cuda")
    visplane = vis_c.clone()
    ops.exchange_pack(g2d_c, visplane, heur_c, rank_max[j], groups, records[j], cap, ws, clear_acc=acc, clear_vis=visplane)
    for t in ex_grads:
      assert not t.any(), "pack must leave the dense gradient planes all-zero"
    assert not visplane.any() and torch.equal(acc[:, 0] == 0, cu(visible))
    hdr = records[j][:8].view(torch.int32).cpu().tolist()
    assert hdr[0] == counts[-1] and hdr[1] == int(counts[-1] > cap)
  assert max(counts) <= cap, "test sizes keep every record within capacity"
  ops.exchange_fold(n, [records[j].data_ptr() for j in range(images)], cap, groups, sa, pa, ex_state, scene_visible=ex_vis)
  torch.cuda.synchronize()
  # identical arithmetic in identical order: bit-exact against the sequential CUDA path
  for t, s in zip(ex_grads, seq_grads):
    assert torch.equal(t, s)
  assert torch.equal(ex_vis, seq_vis)
  for k in ("prune_cost", "split_score", "points_in_view", "visibility"):
    assert torch.equal(ex_state[k], seq_state[k]), k
  # max_scale_px is kept per rank and MAX-reduced before a densification
  assert torch.equal(torch.stack(rank_max).amax(dim=0), seq_state["max_scale_px"])
  # statistics folded slice by slice (what the ranks of a DP job do) == folded whole
  if n >= 64:
    h = (n // 2) // 4 * 4
    sl_state, sl_vis = _state(n), torch.zeros((n,), device="cuda")
    scratch = [torch.zeros((n, d), device="cuda") for d in DIMS]
    sgroups, _ = ops.exchange_groups(scratch)
    ptrs = [records[j].data_ptr() for j in range(images)]
    ops.exchange_fold(n, ptrs, cap, sgroups, sa, pa, sl_state, scene_visible=sl_vis, stats_rows=(0, h))
    assert not sl_state["split_score"][h:].any() and not sl_state["points_in_view"][h:].any()
    for t in scratch:
      t.zero_()
    sl_vis.zero_()
    ops.exchange_fold(n, ptrs, cap, sgroups, sa, pa, sl_state, scene_visible=sl_vis, stats_rows=(h, n))
    for k in ("prune_cost", "split_score", "points_in_view", "visibility"):
      assert torch.equal(sl_state[k], ex_state[k]), k
    assert torch.equal(sl_vis, ex_vis)
    for t, g in zip(scratch, ex_grads):
      assert torch.equal(t, g)
  # and against the CPU oracle's sequential restatement
  for k in ("prune_cost", "split_score", "visibility"):
    np.testing.assert_allclose(npy(ex_state[k]), ref_state[k], rtol=2e-5, atol=2e-6, err_msg=k)
  assert np.array_equal(npy(ex_state["points_in_view"]), ref_state["points_in_view"])
  np.testing.assert_allclose(npy(ex_vis), ref_state["scene_visible"], rtol=1e-6)


def test_pack_reports_overflow():
  rng = np.random.default_rng(5)
  n, cap = 3000, 128
  g2d, vis, heur, grads, _, visible = _fake_image(rng, n, 0.9, 0.9)
  planes = [cu(g) for g in grads]
  groups, row_floats = ops.exchange_groups(planes)
  rec = torch.zeros((ops.exchange_record_bytes(n, cap, row_floats),), device="cuda", dtype=torch.uint8)
  ws = torch.empty((int(L.load().gs_exchange_pack_workspace_bytes(n)),), device="cuda", dtype=torch.uint8)
  ops.exchange_pack(cu(g2d), cu(vis), cu(heur), torch.zeros((n,), device="cuda"), groups, rec, cap, ws)
  hdr = rec[:8].view(torch.int32).cpu().tolist()
  assert hdr[0] == int(visible.sum()) > cap and hdr[1] == 1


def test_emulated_ranks_match_sequential_training_step():
  """three replicas, one image each, records folded by every replica == one scene stepping over the batch of three."""
  n, W, H = 30000, 320, 240
  g = syn.garden_scene(n, 2, seed=3, focal=0.5 * W / np.tan(np.radians(25.0)))
  cams = [L.make_camera(c["T"], c["fx"], c["fy"], c["cx"], c["cy"], W, H) for c in syn.ring_cameras(8, W, H)[:3]]
  gen = torch.Generator(device="cuda").manual_seed(0)
  targets = [torch.rand((H, W, 3), device="cuda", generator=gen) for _ in cams]
  cfg = lambda: SceneConfig(sh_degree=2)
  seq = SplatScene(g, cfg())
  seq_tr = Trainer(seq)
  reps = [SplatScene(g, cfg()) for _ in cams]
  exs = [ImageExchange(s, world=1, rank=0, cap=n, transport="local") for s in reps]
  for step in range(2):
    seq_tr.training_step([(c, t) for c, t in zip(cams, targets)])
    for r, s in enumerate(reps):
      s.forward_backward(cams[r], targets[r], exchange=exs[r], slot=0)
    ptrs = [e.record(0).data_ptr() for e in exs]
    counts = [e.record(0)[:8].view(torch.int32).cpu().tolist() for e in exs]
    assert all(c[0] > 100 and c[1] == 0 for c in counts)
    for r, s in enumerate(reps):
      exs[r].fold(extra_ptrs=ptrs)
      if step == 0 and r == 0:
        vis_union = int((s.visible > 0).sum().item())
        assert vis_union >= max(c[0] for c in counts)
      s.step()
    torch.cuda.synchronize()
    # replicas fold the same records in the same order: bit-identical parameters, moments and statistics
    for s in reps[1:]:
      assert torch.equal(s.params, reps[0].params) and torch.equal(s.exp_avg, reps[0].exp_avg)
      for k in ("prune_cost", "split_score", "points_in_view", "visibility"):
        assert torch.equal(s.stats[k], reps[0].stats[k]), k
    # vs the sequential step: per-image gradients come from fp32 atomics, so equal up to summation order
    a, b = reps[0], seq
    assert not a.grads.any() and not b.grads.any(), "the optimiser clears every consumed row"
    for name in ("position", "log_scaling", "rotation", "alpha_logit", "sh"):
      d = (a.p[name] - b.p[name]).abs().max().item()
      # LaProp normalises tiny gradients: a last-bit gradient change can move an update by a fraction of lr (lr 0.3 ... 0.01)
      assert d < 2e-3, (name, d)
    assert torch.equal(a.stats["points_in_view"], b.stats["points_in_view"])
    mx = torch.stack([s.stats["max_scale_px"] for s in reps]).amax(dim=0)
    if step == 0:  # identical parameters on both sides during the first step
      assert torch.equal(mx, b.stats["max_scale_px"])
    else:
      torch.testing.assert_close(mx, b.stats["max_scale_px"], rtol=1e-3, atol=1e-4)
    for k in ("prune_cost", "split_score", "visibility"):
      torch.testing.assert_close(a.stats[k], b.stats[k], rtol=1e-3, atol=1e-5)
