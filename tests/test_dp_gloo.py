"""Camera-batch data parallelism, host logic, world_size 2 over gloo on CPU.

The kernels need a GPU, so the scene is a stub with the SplatScene surface the Trainer touches (grads / visible /
forward_backward / step); what is checked is the Trainer's own logic: rank r takes batch[r::world], gradients and the
`visible` plane are SUMMED (not averaged: the reference accumulates .grad over the batch, trainer/trainer.py:500-514),
every rank ends the step with identical parameters, equal to the sequential single-process result, and the image counter
advances by len(batch) (trainer/trainer.py:532)."""
import os
import socket
import sys

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "splat-trainer_b200"))

N, D = 64, 5


class StubScene:
  """parameters (N,D); image k contributes gradient rows only where its 'camera' sees them."""
  device = torch.device("cpu")

  def __init__(self):
    self.params = torch.arange(N * D, dtype=torch.float32).reshape(N, D) / 10.0
    self.grads = torch.zeros(N * D + N)  # packed gradients + `visible` at the tail (one collective moves both)
    self.visible = self.grads[N * D:]
    self.seen = []

  def _mark(self, name):
    pass

  def forward_backward(self, cam, image, first=False, exchange=None, slot=0, target_ready=None):
    k = int(cam)
    rows = torch.arange(N)[(torch.arange(N) + k) % 3 == 0]
    g = self.grads[:N * D].view(N, D)
    if first or exchange is not None:
      assert float(self.grads.abs().sum()) == 0.0, "first image of a step must find a clean gradient buffer"
    g[rows] += (k + 1) * image.float().mean() * torch.ones(D)
    self.visible[rows] += 0.25 * (k + 1)
    self.seen.append(k)
    if exchange is not None:
      exchange.pack(slot, self)
    return torch.tensor([float(k), float(k * k)])

  def zero_grad(self):
    self.grads.zero_()

  def step(self):
    g = self.grads[:N * D].view(N, D)
    vis = self.visible > 0
    self.params[vis] -= 0.1 * g[vis] / self.visible[vis].unsqueeze(1)
    self.grads.zero_()


class StubExchange:
  """the ImageExchange protocol the Trainer drives (pack per image slot, one fold per step, records applied in image
  order j = rank + world*slot), with dense CPU records moved by a gloo all_gather."""

  def __init__(self, world, rank, k):
    self.world, self.rank, self.k = world, rank, k
    self.records = torch.zeros(k, N * D + N)
    self.folds = 0

  def stale(self):
    return False

  def pack(self, slot, scene):
    self.records[slot].copy_(scene.grads)
    scene.grads.zero_()  # the dense buffer is clean again; fold() adds every image's record in image order

  def fold(self):
    gathered = [torch.zeros_like(self.records) for _ in range(self.world)]
    dist.all_gather(gathered, self.records)
    for slot in range(self.k):
      for r in range(self.world):
        self.scene.grads += gathered[r][slot]
    self.folds += 1


def run_step(world, rank, batch, mode="dense"):
  from gsb200.trainer import Trainer
  scene = StubScene()
  if mode == "records" and world > 1:
    ex = StubExchange(world, rank, len(batch[0]) // world)
    ex.scene = scene
    tr = Trainer(scene, world, rank, exchange=ex)
  else:
    tr = Trainer(scene, world, rank, exchange="dense")
  losses = [tr.training_step(b) for b in batch]
  return scene, tr, losses


def make_batches():
  g = torch.Generator().manual_seed(0)
  images = [torch.randint(0, 255, (4, 4, 3), generator=g, dtype=torch.uint8) for _ in range(8)]
  return [[(k, images[k]) for k in range(0, 4)], [(k, images[k]) for k in range(4, 8)]]


def worker(rank, world, port, out, mode):
  os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
  dist.init_process_group("gloo", rank=rank, world_size=world)
  scene, tr, losses = run_step(world, rank, make_batches(), mode)
  if mode == "records":
    assert tr.exchange.folds == 2, "one fold per optimiser step"
  torch.save(dict(params=scene.params, seen=scene.seen, steps=tr.step_count, losses=[l.tolist() for l in losses]),
             os.path.join(out, f"rank{rank}.pt"))
  dist.destroy_process_group()


def free_port():
  with socket.socket() as s:
    s.bind(("127.0.0.1", 0))
    return s.getsockname()[1]


import pytest


@pytest.mark.parametrize("mode", ["dense", "records"])
def test_two_ranks_equal_sequential(tmp_path, mode):
  """dense: SUM all-reduce of the packed gradient buffer; records: per-image records packed by every rank and folded
  in image order (the ImageExchange protocol)."""
  world = 2
  mp.spawn(worker, args=(world, free_port(), str(tmp_path), mode), nprocs=world, join=True)
  r = [torch.load(os.path.join(tmp_path, f"rank{k}.pt")) for k in range(world)]
  seq_scene, seq_tr, _ = run_step(1, 0, make_batches())
  # sharding: rank r renders batch[r::world] of every step
  assert r[0]["seen"] == [0, 2, 4, 6] and r[1]["seen"] == [1, 3, 5, 7]
  assert sorted(r[0]["seen"] + r[1]["seen"]) == seq_scene.seen
  # SUM semantics: both ranks hold the sequential result
  assert torch.allclose(r[0]["params"], seq_scene.params, rtol=0, atol=1e-5)
  assert torch.equal(r[0]["params"], r[1]["params"])
  # the step counter counts images of the whole batch on every rank
  assert r[0]["steps"] == r[1]["steps"] == seq_tr.step_count == 8
  # per-rank losses are the sums over the rank's own images
  assert r[0]["losses"][0] == [0.0 + 2.0, 0.0 + 4.0] and r[1]["losses"][0] == [1.0 + 3.0, 1.0 + 9.0]


def test_record_pointers_follow_image_order():
  """image j of a step is batch[j]; rank r holds images r, r+world, ... in slots 0, 1, ...: the fold must visit
  (slot 0: rank 0..W-1), (slot 1: rank 0..W-1), ... for every transport."""
  from gsb200.exchange import ImageExchange
  ex = ImageExchange.__new__(ImageExchange)
  ex.world, ex.k, ex.rec_bytes, ex.parity = 3, 2, 1000, 1
  ex.transport, ex.peer_ptrs = "peer", [10_000_000, 20_000_000, 30_000_000]
  off = ex.parity * ex.k * ex.rec_bytes
  assert ex.record_ptrs() == [10_000_000 + off, 20_000_000 + off, 30_000_000 + off,
                              10_000_000 + off + 1000, 20_000_000 + off + 1000, 30_000_000 + off + 1000]
  ex.transport = "nccl"

  class G:
    def data_ptr(self):
      return 500
  ex.gathered = G()
  assert ex.record_ptrs() == [500, 500 + 2000, 500 + 4000, 500 + 1000, 500 + 3000, 500 + 5000]


def _color_grads(rank, step):
  """rank-specific gradients of the replicated colour model: dense weight gradient + one GLO row per rank."""
  g = torch.Generator().manual_seed(100 + rank + 10 * step)
  return torch.randn((10088,), generator=g) * 1e-2, rank + 1, torch.randn((32,), generator=g) * 1e-2


def color_worker(rank, world, port, out):
  os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
  dist.init_process_group("gloo", rank=rank, world_size=world)
  from gsb200.color import ColorModule
  cm = ColorModule(4, device="cpu", seed=3)
  for step in range(2):
    dense, row, glo = _color_grads(rank, step)
    cm.grad.copy_(dense)
    cm.glo_grad_row(row).copy_(glo)
    cm.all_reduce(dist)
    cm.step()
  torch.save(dict(flat=cm.flat, glo=cm.glo), os.path.join(out, f"color{rank}.pt"))
  dist.destroy_process_group()


def test_replicated_colour_model_stays_identical_under_dp(tmp_path):
  """Trainer.exchange_gradients sums the colour model's dense gradients and GLO rows over the ranks (SUM, like the
  splat gradients): every replica then takes the same Adam step as one process holding the summed gradients."""
  world = 2
  mp.spawn(color_worker, args=(world, free_port(), str(tmp_path)), nprocs=world, join=True)
  r = [torch.load(os.path.join(tmp_path, f"color{k}.pt")) for k in range(world)]
  assert torch.equal(r[0]["flat"], r[1]["flat"]) and torch.equal(r[0]["glo"], r[1]["glo"])
  from gsb200.color import ColorModule
  cm = ColorModule(4, device="cpu", seed=3)
  for step in range(2):
    for rank in range(world):
      dense, row, glo = _color_grads(rank, step)
      cm.grad.add_(dense)
      cm.glo_grad_row(row).add_(glo)
    cm.step()
  assert torch.allclose(r[0]["flat"], cm.flat, rtol=0, atol=1e-5) and torch.allclose(r[0]["glo"], cm.glo, rtol=0, atol=1e-5)
  assert r[0]["glo"][1].abs().max() > 0 and r[0]["glo"][2].abs().max() > 0 and not r[0]["glo"][0].any() and not r[0]["glo"][3].any()


# ---- row-sharded optimiser: fold + step own row slice, publish the rows, gather the state ------------
class RowsScene(StubScene):
  """StubScene + per-row optimiser state and controller statistics, stepped by row slice like SplatScene.step(rows=)."""
  n = N

  def __init__(self):
    super().__init__()
    self.m1 = dict(w=torch.zeros(N, D))
    self.m2 = dict(w=torch.zeros(N))
    self.running_vis, self.opt_step = torch.zeros(N), torch.zeros(N)
    self.stats = dict(max_scale_px=torch.zeros(N), prune_cost=torch.zeros(N), split_score=torch.zeros(N),
                      visibility=torch.zeros(N), points_in_view=torch.zeros(N, dtype=torch.int16))
    self.stats_partial = self.state_partial = None

  def step(self, rows=None):
    b, e = rows if rows is not None else (0, N)
    g = self.grads[:N * D].view(N, D)
    sel = torch.zeros(N, dtype=torch.bool)
    sel[b:e] = self.visible[b:e] > 0
    self.m1["w"][sel] = 0.9 * self.m1["w"][sel] + 0.1 * g[sel]
    self.m2["w"][sel] += 1.0
    self.opt_step[sel] += 1.0
    self.running_vis[sel] = 0.5 * self.running_vis[sel] + 0.5 * self.visible[sel]
    self.params[sel] -= 0.1 * self.m1["w"][sel]
    self.touched = sel
    self.grads.zero_()


class RowsExchange(StubExchange):
  """shard="rows" protocol: only the rank's slice of every record is folded; publish() sends the rows the rank stepped to
  every replica (here one all_gather of the slices); the statistics are folded per slice too."""
  shard, transport, group = "rows", "peer", None

  def __init__(self, world, rank, k):
    super().__init__(world, rank, k)
    self.slice_len = (N + world - 1) // world
    b = min(rank * self.slice_len, N)
    self.grad_rows = self.stats_rows = (b, min(b + self.slice_len, N))
    self.published = 0

  def fold(self):
    gathered = [torch.zeros_like(self.records) for _ in range(self.world)]
    dist.all_gather(gathered, self.records)
    b, e = self.grad_rows
    g, vis = self.scene.grads[:N * D].view(N, D), self.scene.visible
    for slot in range(self.k):
      for r in range(self.world):
        rec = gathered[r][slot]
        g[b:e] += rec[:N * D].view(N, D)[b:e]
        vis[b:e] += rec[N * D:][b:e]
        seen = rec[N * D:][b:e] > 0
        self.scene.stats["points_in_view"][b:e] += seen.to(torch.int16)
        self.scene.stats["visibility"][b:e] += rec[N * D:][b:e]
    self.folds += 1

  def publish(self):
    b, e = self.grad_rows
    send = torch.zeros(self.slice_len, D)
    send[:e - b] = self.scene.params[b:e]
    out = torch.zeros(self.world * self.slice_len, D)
    dist.all_gather_into_tensor(out, send)
    self.scene.params.copy_(out[:N])
    self.published += 1


def rows_worker(rank, world, port, out):
  os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
  dist.init_process_group("gloo", rank=rank, world_size=world)
  from gsb200 import lib as L
  from gsb200.trainer import Trainer
  scene = RowsScene()
  ex = RowsExchange(world, rank, 2)
  ex.scene = scene
  tr = Trainer(scene, world, rank, exchange=ex)
  assert tr.sharded
  for b in make_batches():
    tr.training_step(b)
  b, e = ex.grad_rows
  others = torch.ones(N, dtype=torch.bool)
  others[b:e] = False
  partial = dict(stale_moments=bool((scene.m1["w"][others] == 0).all()), flagged=bool(scene.state_partial and scene.stats_partial))
  tr.sync_statistics()
  torch.save(dict(params=scene.params, m1=scene.m1["w"], m2=scene.m2["w"], opt_step=scene.opt_step, running_vis=scene.running_vis,
                  piv=scene.stats["points_in_view"], visibility=scene.stats["visibility"], partial=partial,
                  published=ex.published, cleared=scene.state_partial is None and scene.stats_partial is None),
             os.path.join(out, f"rows{rank}.pt"))
  dist.destroy_process_group()


def test_row_sharded_optimiser_equals_sequential(tmp_path):
  """Trainer with an exchange whose shard is "rows": every rank steps only its row slice (scene.step(rows=...)), publishes
  it, and Trainer.sync_statistics() gathers the per-slice optimiser state and statistics: afterwards both replicas hold
  what one process stepping every row holds."""
  world = 2
  mp.spawn(rows_worker, args=(world, free_port(), str(tmp_path)), nprocs=world, join=True)
  r = [torch.load(os.path.join(tmp_path, f"rows{k}.pt")) for k in range(world)]
  from gsb200.trainer import Trainer
  seq = RowsScene()
  seq_tr = Trainer(seq, 1, 0, exchange="dense")
  piv, visib = torch.zeros(N, dtype=torch.int16), torch.zeros(N)
  for b in make_batches():
    seq_tr.training_step(b)
  for k in range(8):  # the statistics the sharded fold accumulated per slice: every image's visibility of every row
    rows = (torch.arange(N) + k) % 3 == 0
    piv += rows.to(torch.int16)
    visib += rows.float() * 0.25 * (k + 1)
  for k in ("params", "m1", "m2", "opt_step", "running_vis"):
    assert torch.equal(r[0][k], r[1][k]), k
  assert torch.allclose(r[0]["params"], seq.params, rtol=0, atol=1e-5)
  assert torch.allclose(r[0]["m1"], seq.m1["w"], rtol=0, atol=1e-6) and torch.equal(r[0]["opt_step"], seq.opt_step)
  assert torch.equal(r[0]["piv"], piv) and torch.equal(r[1]["piv"], piv) and torch.allclose(r[0]["visibility"], visib)
  for k in range(world):
    assert r[k]["partial"] == dict(stale_moments=True, flagged=True)  # before the gather: other slices' state untouched
    assert r[k]["published"] == 2 and r[k]["cleared"]
