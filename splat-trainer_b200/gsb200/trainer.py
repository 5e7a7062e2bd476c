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
  def __init__(self, scene, world_size=1, rank=0):
    self.scene = scene
    self.world_size, self.rank = world_size, rank
    self.step_count = 0
    self._dev_target = None
    self._grads_clean = False  # unknown history of scene.grads until the first step has run

  def _to_device(self, image):
    """host (pinned) uint8/float image -> reusable device buffer, async on the current stream
    (reference trainer/trainer.py:548-553 does the same copy in its loader thread)."""
    if image.is_cuda:
      return image
    if self._dev_target is None or self._dev_target.shape != image.shape or self._dev_target.dtype != image.dtype:
      self._dev_target = torch.empty(image.shape, dtype=image.dtype, device=self.scene.device)
    self._dev_target.copy_(image, non_blocking=True)
    return self._dev_target

  def exchange_gradients(self):
    """SUM (not mean: reference accumulates .grad over the batch, trainer/trainer.py:500-514) of the packed
    gradient buffer; the tail plane carries `visible` so one collective moves both."""
    if self.world_size > 1:
      dist.all_reduce(self.scene.grads, op=dist.ReduceOp.SUM)
      self.scene._mark("allreduce")

  def training_step(self, batch, read_loss=False):
    """batch: list of (camera, target image) for the WHOLE step (all ranks pass the same list); this rank
    processes batch[rank::world].  Returns the summed (Σ|d|, Σd²) of this rank's images as a device tensor, or
    python floats if read_loss (one 8-byte device→host read)."""
    scene = self.scene
    mine = batch[self.rank::self.world_size]
    total = None
    for k, (cam, image) in enumerate(mine):
      sums = scene.forward_backward(cam, self._to_device(image), first=(k == 0 and self._grads_clean))
      total = sums.clone() if total is None else total + sums
    self.step_count += len(batch)  # the reference counts images, trainer/trainer.py:532
    self.exchange_gradients()
    scene.step()
    self._grads_clean = True  # scene.step() cleared every gradient row it consumed
    if read_loss:
      return [float(x) for x in total.cpu().tolist()] if total is not None else [0.0, 0.0]
    return total
