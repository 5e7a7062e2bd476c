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
  def __init__(self, scene, world_size=1, rank=0, exchange="auto", exchange_cap=0, images_per_rank=1):
    """exchange: "auto" (sparse per-image records over peer memory when world_size > 1), "peer", "nccl" (records
    all-gathered with NCCL), "dense" (SUM all-reduce of the whole packed gradient buffer; statistics stay per rank)
    or an ImageExchange instance."""
    self.scene = scene
    self.world_size, self.rank = world_size, rank
    self.step_count = 0
    self._copy_stream, self._target_flip = None, 0
    self._grads_clean = False  # unknown history of scene.grads until the first step has run
    self.exchange = None
    if isinstance(exchange, str):
      if world_size > 1 and exchange != "dense":
        from .exchange import ImageExchange
        scene.zero_grad()
        self.exchange = ImageExchange(scene, world_size, rank, cap=exchange_cap, images_per_rank=images_per_rank,
                                      transport=exchange)
    else:
      self.exchange = exchange

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
      self.scene._mark("exchange")
    elif self.world_size > 1:
      dist.all_reduce(self.scene.grads, op=dist.ReduceOp.SUM)
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
                                  group=ex.group)
    return rows

  def sync_statistics(self):
    """before a densification (or any read of scene.stats) under DP: max_scale_px is kept per rank (MAX is order
    independent); the other statistics were folded in image order by the rank that owns the row slice — all-gather the
    slices.  Afterwards scene.stats is complete and identical on every rank."""
    if self.world_size <= 1:
      return
    st = self.scene.stats
    dist.all_reduce(st["max_scale_px"], op=dist.ReduceOp.MAX)
    ex = self.exchange
    if ex is None or ex.stats_rows == (0, self.scene.n):
      return
    b, e = ex.stats_rows
    for k in ("prune_cost", "split_score", "visibility", "points_in_view"):
      t = st[k]
      dt = torch.int32 if t.dtype == torch.int16 else t.dtype  # NCCL has no int16
      send = torch.zeros((ex.slice_len,), device=t.device, dtype=dt)
      send[:e - b] = t[b:e].to(dt)
      out = torch.empty((self.world_size * ex.slice_len,), device=t.device, dtype=dt)
      dist.all_gather_into_tensor(out, send, group=ex.group)
      t.copy_(out[:self.scene.n].to(t.dtype))

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
      self.exchange = ImageExchange(scene, e.world, e.rank, cap=e.cap, images_per_rank=e.k, transport=e.transport, group=e.group)
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
    scene.step()
    self._grads_clean = True  # scene.step() cleared every gradient row it consumed
    if read_loss:
      return [float(x) for x in total.cpu().tolist()] if total is not None else [0.0, 0.0]
    return total
