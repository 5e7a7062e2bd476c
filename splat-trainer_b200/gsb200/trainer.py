"""training_step — the reference's hot loop (trainer/trainer.py:500-545) over a SplatScene.

One optimiser step over a batch of images: for each image render → loss → backward → statistics (gradients
SUM over the batch, exactly the reference's sequential accumulation), then scene.step().  Under
torch.distributed the batch is sharded by rank (rank r takes batch[r::world]) and the packed gradient buffer
(+ the visibility plane at its tail) is SUM-all-reduced with NCCL before the optimiser step — the one exchange
step of the path.
"""
import torch
import torch.distributed as dist

from . import lib as L


class StageTimer:
  """CUDA-event stage timing on the launching stream (events only; no host sync until .collect())."""

  def __init__(self):
    self.marks = []

  def mark(self, name):
    ev = torch.cuda.Event(enable_timing=True)
    ev.record(torch.cuda.current_stream())
    self.marks.append((name, ev))

  def collect(self):
    """-> dict stage -> list of durations (ms); a stage is the span from the previous mark to its own."""
    torch.cuda.synchronize()
    out = {}
    for (_, e0), (n1, e1) in zip(self.marks[:-1], self.marks[1:]):
      if n1 in ("begin", "optim_begin"):
        continue
      out.setdefault(n1, []).append(e0.elapsed_time(e1))
    self.marks = []
    return out


class Trainer:
  def __init__(self, scene, world_size=1, rank=0, exchange="auto", exchange_cap=0, images_per_rank=1, controller=None,
               check_every=50, strict_overflow=False, shard="auto"):
    """exchange: "auto" (sparse per-image records over peer memory when world_size > 1), "peer", "nccl" (records
    all-gathered with NCCL), "dense" (SUM all-reduce of the whole packed gradient buffer; statistics stay per rank:
    a bench arm only, densification raises) or an ImageExchange instance.
    shard: "rows" (each rank folds and steps its own row slice and pushes the updated rows to its peers; needs the peer
    transport), "replicated" (every rank steps every row) or "auto" (rows whenever the transport is peer memory).
    controller: a gsb200.controller.TargetController stepped after every optimiser step (reference
    trainer/trainer.py:672-689).  check_every: the two device-side overflow flags (tile list, exchange record) are read
    every that many steps and right after a densification — one 8-byte read; on overflow the buffer is regrown for the
    following steps and the event is counted in `overflow_events` (strict_overflow: raise GsError instead)."""
    self.scene = scene
    self.world_size, self.rank = world_size, rank
    self.step_count = 0
    self.steps_done = 0
    self.controller = controller
    if controller is not None and controller.trainer is None:
      controller.trainer = self
    self.check_every, self.strict_overflow = int(check_every), bool(strict_overflow)
    self.overflow_events = []
    self._check_now = True  # first step: the capacities were calibrated from one camera only
    self.dense = isinstance(exchange, str) and exchange == "dense" and world_size > 1
    self._copy_stream, self._target_flip = None, 0
    self._grads_clean = False  # unknown history of scene.grads until the first step has run
    self.exchange = None
    if isinstance(exchange, str):
      if world_size > 1 and exchange != "dense":
        from .exchange import ImageExchange
        scene.zero_grad()
        self.exchange = ImageExchange(scene, world_size, rank, cap=exchange_cap, images_per_rank=images_per_rank,
                                      transport=exchange, shard=shard)
    else:
      self.exchange = exchange

  @property
  def sharded(self):
    """True when the optimiser is sharded by row slice (ImageExchange shard="rows")."""
    return getattr(self.exchange, "shard", "replicated") == "rows"

  def _to_device(self, image):
    """host (pinned) uint8/float image -> one of two reusable device buffers, copied on a side stream so the transfer
    runs under the projection / tile map / raster forward of the same image (reference trainer/trainer.py:548-553 does
    the copy in its loader thread).  Returns (device image, event the loss must wait for | None)."""
    if image.is_cuda or torch.device(self.scene.device).type != "cuda":  # already resident (or a host-side stub scene)
      return image, None
    if self._copy_stream is None:
      self._copy_stream = torch.cuda.Stream(device=self.scene.device)
      self._targets = [None, None]
      self._consumed = [None, None]
    k = self._target_flip
    self._target_flip ^= 1
    buf = self._targets[k]
    if buf is None or buf.shape != image.shape or buf.dtype != image.dtype:
      buf = self._targets[k] = torch.empty(image.shape, dtype=image.dtype, device=self.scene.device)
    cs = self._copy_stream
    if self._consumed[k] is not None:
      cs.wait_event(self._consumed[k])  # the loss of two images ago has finished reading this buffer
    with torch.cuda.stream(cs):
      buf.copy_(image, non_blocking=True)
      ready = torch.cuda.Event()
      ready.record(cs)
    return buf, (ready, k)

  def exchange_gradients(self):
    """SUM (not mean: reference accumulates .grad over the batch, trainer/trainer.py:500-514) of the packed
    gradient buffer; the tail plane carries `visible` so one collective moves both."""
    if self.world_size > 1 and getattr(self.scene, "color", None) is not None:
      self.scene.color.all_reduce(dist)  # replicated colour model: its small dense gradients are summed
    if self.exchange is not None:
      self.exchange.fold()
      if self.world_size > 1 and getattr(self.exchange, "stats_rows", None) not in (None, (0, getattr(self.scene, "n", 0))):
        self.scene.stats_partial = "each rank holds its own row slice of the statistics: call Trainer.sync_statistics() first"
      self.scene._mark("exchange")
    elif self.world_size > 1:
      dist.all_reduce(self.scene.grads, op=dist.ReduceOp.SUM)
      self.scene.stats_partial = "exchange='dense' keeps the statistics per rank (bench arm only): densification would diverge the replicas"
      self.scene._mark("exchange")

  def calibrate_exchange(self, margin=1.5):
    """size the exchange records from the rows the last step's images blended (max over ranks) × margin.  One host
    read + one tiny collective, at start-up or after a densification; returns the measured rows per image."""
    ex = self.exchange
    rows = max(c for c, _ in ex.counts())
    if self.world_size > 1:
      t = torch.tensor([rows], device=self.scene.device, dtype=torch.int64)
      dist.all_reduce(t, op=dist.ReduceOp.MAX, group=ex.group)
      rows = int(t.item())
    from .exchange import ImageExchange
    cap = min(self.scene.n, int(rows * margin) + 4096)
    self.scene.zero_grad()
    self.exchange = ImageExchange(self.scene, ex.world, ex.rank, cap=cap, images_per_rank=ex.k, transport=ex.transport,
                                  group=ex.group, shard=getattr(ex, "shard", "auto"), reuse=ex)
    return rows

  def sync_statistics(self):
    """before a densification (or any read of scene.stats) under DP: max_scale_px is kept per rank (MAX is order
    independent); the other statistics were folded in image order by the rank that owns the row slice — all-gather the
    slices.  Afterwards scene.stats is complete and identical on every rank."""
    if self.world_size <= 1:
      return
    if self.dense:
      raise L.GsError("exchange='dense' keeps split_score / prune_cost per rank; use the record exchange to densify")
    st = self.scene.stats
    ex = self.exchange
    if ex is not None and hasattr(ex, "stale") and ex.stale():
      # the scene was resized by a densification since the last fold: its statistics are the fresh all-zero PointState on
      # every rank (reference controller/target_controller.py:122), nothing to gather
      self.scene.stats_partial = None
      self.scene.state_partial = None
      return
    dist.all_reduce(st["max_scale_px"], op=dist.ReduceOp.MAX)
    self.scene.stats_partial = None
    if ex is None or ex.stats_rows == (0, self.scene.n):
      return
    for k in ("prune_cost", "split_score", "visibility", "points_in_view"):
      self._gather_rows(st[k])
    if self.sharded and getattr(self.scene, "state_partial", None):
      # row-sharded optimiser: moments, running visibility and step counts live on the slice's owner
      sc = self.scene
      for name in sc.m1:
        self._gather_rows(sc.m1[name])
        self._gather_rows(sc.m2[name])
      self._gather_rows(sc.running_vis)
      self._gather_rows(sc.opt_step)
      sc.state_partial = None

  def _gather_rows(self, t):
    """t: (n, ...) contiguous view whose rows [b, e) of this rank's slice are current -> all rows current (all-gather of
    the equal-length, zero-padded slices, in place)."""
    ex, n = self.exchange, self.scene.n
    b, e = ex.stats_rows
    flat = t.view(n, -1)
    dt = torch.int32 if t.dtype == torch.int16 else t.dtype  # NCCL has no int16
    send = torch.zeros((ex.slice_len, flat.shape[1]), device=t.device, dtype=dt)
    send[:e - b] = flat[b:e].to(dt)
    out = torch.empty((self.world_size * ex.slice_len, flat.shape[1]), device=t.device, dtype=dt)
    dist.all_gather_into_tensor(out, send, group=ex.group)
    flat.copy_(out[:n].to(t.dtype))

  def check_overflow(self):
    """read the device-side overflow flags of the last step (tile-overlap list; exchange record) — one small host read,
    never on the per-step path.  An overflowed buffer silently truncated that step's lists / dropped rows: grow it for the
    steps to come and report."""
    self._check_now = False
    scene = self.scene
    grew = []
    if hasattr(scene, "overflowed") and scene.overflowed():
      scene.grow_overlaps(2.0)
      grew.append("tile_overlaps")
    ex = self.exchange
    if ex is not None and hasattr(ex, "counts") and not ex.stale():
      flag = torch.tensor([1.0 if any(o for _, o in ex.counts()) else 0.0], device=scene.device)
      if self.world_size > 1:
        dist.all_reduce(flag, op=dist.ReduceOp.MAX, group=ex.group)  # every rank must resize together
      if float(flag.item()) > 0:
        from .exchange import ImageExchange
        scene.zero_grad()  # rows a full record dropped are still in the dense buffer
        self.exchange = ImageExchange(scene, ex.world, ex.rank, cap=min(scene.n, 2 * ex.cap), images_per_rank=ex.k,
                                      transport=ex.transport, group=ex.group, shard=getattr(ex, "shard", "auto"), reuse=ex)
        grew.append("exchange_record")
    if grew:
      self.overflow_events.append((self.steps_done, grew))
      self._check_now = True
      if self.strict_overflow:
        raise L.GsError(f"step {self.steps_done}: capacity exceeded ({', '.join(grew)}); the step's lists were truncated")
    return grew

  def state_dict(self):
    """the reference's checkpoint layout (trainer/trainer.py:188-194): dict(step, scene, controller).  Under DP every
    rank must call it (the row slices of the statistics / optimiser state are gathered first)."""
    if self.world_size > 1 and not self.dense:
      self.sync_statistics()
    ctrl = self.controller.state_dict() if self.controller is not None else dict(
        points={k: v.clone() for k, v in self.scene.stats.items()}, start_points=self.scene.num_points)
    return dict(step=int(self.step_count), scene=self.scene.state_dict(), controller=ctrl)

  @staticmethod
  def from_state_dict(state, controller_config=None, target_points=0, total_steps=0, device="cuda", **kw):
    from .scene import SplatScene
    scene = SplatScene.from_state_dict(state["scene"], device=device)
    controller = None
    if controller_config is not None:
      from .controller import TargetController
      controller = TargetController(controller_config, scene, target_points, total_steps, step=int(state["step"]))
      controller.load_state_dict(state["controller"])
    else:
      for k, v in state["controller"]["points"].items():
        scene.stats[k].copy_(v)
    tr = Trainer(scene, controller=controller, **kw)
    tr.step_count = int(state["step"])
    return tr

  def training_step(self, batch, read_loss=False):
    """batch: list of (camera, target image) for the WHOLE step (all ranks pass the same list); this rank
    processes batch[rank::world].  Returns the summed (Σ|d|, Σd²) of this rank's images as a device tensor, or
    python floats if read_loss (one 8-byte device→host read)."""
    scene = self.scene
    mine = batch[self.rank::self.world_size]
    total = None
    if self.exchange is not None and self.exchange.stale():  # the scene was re-allocated by a densification
      from .exchange import ImageExchange
      e = self.exchange
      scene.zero_grad()
      self.exchange = ImageExchange(scene, e.world, e.rank, cap=e.cap, images_per_rank=e.k, transport=e.transport, group=e.group,
                                    shard=getattr(e, "shard", "auto"), reuse=e)
    if self.exchange is not None and len(mine) != self.exchange.k:
      raise L.GsError(f"the exchange was sized for {self.exchange.k} image(s) per rank, got {len(mine)}")
    for k, item in enumerate(mine):
      cam, image = item[0], item[1]
      extra = dict(image_idx=item[2]) if len(item) > 2 else {}  # MLP colour scenes: the image's GLO row
      target, ready = self._to_device(image)
      sums = scene.forward_backward(cam, target, first=(k == 0 and self._grads_clean), exchange=self.exchange, slot=k,
                                    target_ready=None if ready is None else ready[0], **extra)
      if ready is not None:
        done = torch.cuda.Event()
        done.record()  # after this image's backward: its loss kernels are long finished with the target
        self._consumed[ready[1]] = done
      total = sums.clone() if total is None else total + sums
    self.step_count += len(batch)  # the reference counts images, trainer/trainer.py:532
    self.exchange_gradients()
    if self.sharded:
      scene.step(rows=self.exchange.grad_rows)
      self.exchange.publish()
      scene.state_partial = "each rank holds the optimiser state of its own row slice: call Trainer.sync_statistics() first"
      scene._mark("publish")
    else:
      scene.step()
    self._grads_clean = True  # scene.step() cleared every gradient row it consumed
    self.steps_done += 1
    if self._check_now or (self.check_every > 0 and self.steps_done % self.check_every == 0):
      self.check_overflow()
    if self.controller is not None and self.controller.step(self.step_count) is not None:
      self._check_now = True  # the scene changed size: capacities are re-checked after the next step
    if read_loss == "pipelined":
      # the loss of EVERY step is read back, one step late: this step's sums start an asynchronous 8-byte copy into pinned
      # host memory and the previous step's (landed long ago) are returned — no host sync on the step's critical path
      # (the reference itself reads its losses only on logging steps, trainer/trainer.py:485-494); flush_loss() returns the
      # last one
      prev = self.flush_loss()
      if total is not None:
        if self._loss_host is None:
          self._loss_host = [torch.empty((2,), dtype=torch.float32).pin_memory() for _ in range(2)]
        slot = self._loss_host[self.steps_done & 1]
        slot.copy_(total, non_blocking=True)
        ev = torch.cuda.Event()
        ev.record()
        self._loss_pending = (slot, ev)
      return prev
    if read_loss:
      return [float(x) for x in total.cpu().tolist()] if total is not None else [0.0, 0.0]
    return total

  _loss_host, _loss_pending = None, None

  def flush_loss(self):
    """the (Σ|d|, Σd²) of the last step started with read_loss="pipelined" (waits for its copy), or None."""
    pend, self._loss_pending = self._loss_pending, None
    if pend is None:
      return None
    pend[1].synchronize()
    return [float(x) for x in pend[0].tolist()]
