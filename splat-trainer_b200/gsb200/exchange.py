"""ImageExchange — the one exchange step of camera-batch data parallelism.

Every rank packs what ITS images of the step contributed (gradient rows of the splats a pixel blended, their
visibility / heuristics, in-view bit planes) into fixed-capacity records; then every rank folds the records of ALL
images of the step in image order.  The fold reproduces the reference's sequential accumulation
(trainer/trainer.py:500-514, controller/point_state.py:34-50) on every replica, so replicas stay bit-identical.

Transport (the records are written once and read in place):
  "peer"  records live in torch symmetric memory; after a stream-ordered NCCL barrier the fold kernels load the other
          ranks' records straight over NVLink (no all-gather, no staging copy, only `count` rows move).  Records are
          double-buffered by step parity so one barrier per step is enough.
  "nccl"  plain buffers + one all_gather_into_tensor of the fixed-capacity records, folded from the local copy.
  "local" world size 1 (or tests): records of several emulated ranks in one process.

Optimiser sharding (`shard`):
  "replicated"  every rank folds the summed gradients and `visible` of ALL rows and steps all of them (the replicas
                stay identical without a second exchange; cost grows with the world size).
  "rows"        (peer transport) SURVEY 8(e)(1): rank r folds and steps only rows [r·N/W, (r+1)·N/W) and then writes the
                parameter rows it touched into every peer's copy of the packed parameter buffer, which therefore lives
                in symmetric memory; a second stream-ordered barrier ends the step.  The optimiser state of a row
                (moments, running visibility, step count) exists only on its owner until Trainer.sync_statistics()
                all-gathers it (before a densification, a checkpoint or a replica comparison).
In both modes the controller statistics (split_score, prune_cost, points_in_view, visibility) are folded by each rank
for its own row slice only and all-gathered when a densification reads them (Trainer.sync_statistics).
"""
import os

import torch
import torch.distributed as dist

from . import lib as L
from . import ops

ATTRS = ("position", "log_scaling", "rotation", "alpha_logit", "sh")


class ImageExchange:
  def __init__(self, scene, world=1, rank=0, cap=0, images_per_rank=1, transport="auto", group=None, shard="auto", reuse=None):
    """reuse: the ImageExchange this one replaces (after a densification): its symmetric record buffer is taken over when it
    is large enough, so that a growing scene does not pay an allocation + rendezvous across the ranks per densification
    (the buffer is laid out for the scene's row capacity the first time)."""
    self.scene, self.world, self.rank, self.k = scene, world, rank, images_per_rank
    self.group = group
    self.n = scene.n
    dev = scene.device
    self.groups, self.row_floats = ops.exchange_groups([scene.g[a] for a in ATTRS])
    # capacity: rows one image can blend; 0 = a quarter of the scene (grown on overflow, like max_overlaps)
    self.cap = int(cap) if cap > 0 else max(4096, self.n // 4)
    self.rec_bytes = ops.exchange_record_bytes(self.n, self.cap, self.row_floats)
    self.workspace = torch.empty((int(L.load().gs_exchange_pack_workspace_bytes(self.n)),), device=dev, dtype=torch.uint8)
    self.parity = 0
    self.transport = transport
    if transport == "auto":
      self.transport = "peer" if world > 1 else "local"
    slot_bytes = self.k * self.rec_bytes
    if self.transport == "peer":
      try:
        import torch.distributed._symmetric_memory as symm
        if (reuse is not None and reuse.transport == "peer" and reuse.world == world and reuse.group is group
            and reuse.buf.numel() >= 2 * slot_bytes):
          # every rank is past the barrier that closed the last step: nobody reads these records any more
          self.buf, self.handle, self.peer_ptrs = reuse.buf, reuse.handle, reuse.peer_ptrs
          for parity in (0, 1):
            for slot in range(self.k):
              o = parity * slot_bytes + slot * self.rec_bytes
              self.buf[o:o + 8].zero_()  # count / overflow header of the (now meaningless) old records
        else:
          rows = max(self.n, int(getattr(scene, "capacity", 0) or 0))
          room = 2 * self.k * ops.exchange_record_bytes(rows, self.cap, self.row_floats)
          self.buf = symm.empty((max(room, 2 * slot_bytes),), dtype=torch.uint8, device=dev)
          self.handle = symm.rendezvous(self.buf, group or dist.group.WORLD)
          self.peer_ptrs = [int(p) for p in self.handle.buffer_ptrs]
          self.buf.zero_()
      except Exception as e:  # symmetric memory unavailable on this box: NCCL all-gather transport (still GPU-only)
        if transport == "peer":
          raise
        self.transport, self.fallback_reason = "nccl", repr(e)
    if self.transport == "nccl":
      self.buf = torch.zeros((slot_bytes,), device=dev, dtype=torch.uint8)
      self.gathered = torch.zeros((world * slot_bytes,), device=dev, dtype=torch.uint8)
    elif self.transport == "local":
      self.buf = torch.zeros((slot_bytes,), device=dev, dtype=torch.uint8)
    self._token = torch.zeros((1,), device=dev)
    # the step's two barriers: "symm" = the symmetric-memory signal-pad barrier of the record buffer's handle (one small
    # kernel, no NCCL launch: 8 GPUs, c2-sparse 4556 -> 4631 it/s) — peer transport only; "nccl" = a 4-byte all-reduce
    # (GSB200_BARRIER=nccl selects it)
    self.barrier_kind = "nccl"
    if self.transport == "peer" and hasattr(self.handle, "barrier"):
      self.barrier_kind = os.environ.get("GSB200_BARRIER", "symm")
    # controller statistics are only read at a densification: every rank folds them for its own slice of the rows
    # (Trainer.sync_statistics all-gathers the slices), so the statistics cost per rank does not grow with the world
    self.slice_len = (self.n + world - 1) // world
    self.slice_len = (self.slice_len + 127) // 128 * 128
    b = min(rank * self.slice_len, self.n)
    self.stats_rows = (b, min(b + self.slice_len, self.n)) if world > 1 else (0, self.n)
    self._grads_ptr = scene.grads.data_ptr()
    # row-sharded optimiser: needs peer memory (the updated rows are pushed into the peers' parameter planes)
    if shard == "rows" and (self.transport != "peer" or world <= 1):
      raise L.GsError(f"shard='rows' needs the peer transport and world > 1 (transport {self.transport!r}, world {world})")
    self.shard = "rows" if (shard in ("auto", "rows") and self.transport == "peer" and world > 1) else "replicated"
    self.grad_rows = self.stats_rows if self.shard == "rows" else (0, self.n)
    if self.shard == "rows":
      self._param_buf, self._param_handle = self._symmetric_params()
      self.param_groups, _ = ops.exchange_groups([scene.p[a] for a in ATTRS])
      self.param_peers = [int(p) for q, p in enumerate(self._param_handle.buffer_ptrs) if q != rank]
    self._params_ptr = scene.params.data_ptr()

  def _symmetric_params(self):
    """the scene's packed parameter buffer in symmetric memory (moved there once per buffer: start-up, and each side of
    the densification ping-pong); every rank runs this at the same point of its program."""
    import torch.distributed._symmetric_memory as symm
    scene = self.scene
    reg = scene.__dict__.setdefault("_symm_params", {})
    spare = getattr(scene, "_spare", None)
    live = {scene.params.data_ptr()} | ({spare[0].data_ptr()} if spare else set())
    for key in [k for k in reg if k not in live]:
      del reg[key]
    if scene.params.data_ptr() not in reg:
      buf = symm.empty((scene.params.numel(),), dtype=torch.float32, device=scene.device)
      handle = symm.rendezvous(buf, self.group or dist.group.WORLD)
      buf.copy_(scene.params)
      scene.params = buf
      scene._view(scene.n)
      reg[buf.data_ptr()] = (buf, handle)
    return reg[scene.params.data_ptr()]

  def publish(self):
    """rows mode, after scene.step(rows=self.grad_rows): push the slice's touched parameter rows to the peers and close
    the step with a barrier (the next forward on any rank reads all rows)."""
    scene = self.scene
    ops.exchange_publish(scene._vis_idx, self.grad_rows[0], self.param_groups, self._param_buf.data_ptr(), self.param_peers)
    self._barrier(1)

  def _barrier(self, channel=0):
    """stream-ordered barrier across the ranks: everything enqueued before it on any rank is complete (and its peer
    writes visible) before anything enqueued after it on any rank starts."""
    if self.barrier_kind == "symm":
      self.handle.barrier(channel=channel, timeout_ms=120000)  # a skewed start is not a hang
    else:
      dist.all_reduce(self._token, group=self.group)

  def stale(self):
    """True once the scene resized or swapped its packed buffers (densification): build a new ImageExchange.  The
    parameter buffer is compared too: a densification that prunes as many rows as it adds keeps n and the gradient buffer
    but swaps the ping-pong parameter buffers, and the row-sharded publish holds their addresses."""
    return (self.n != self.scene.n or self._grads_ptr != self.scene.grads.data_ptr()
            or self._params_ptr != self.scene.params.data_ptr())

  # ---- this rank's side ----------------------------------------------------------------------------------
  def _local_offset(self, slot, parity=None):
    parity = self.parity if parity is None else parity
    base = parity * self.k * self.rec_bytes if self.transport == "peer" else 0
    return base + slot * self.rec_bytes

  def record(self, slot, parity=None):
    o = self._local_offset(slot, parity)
    return self.buf[o:o + self.rec_bytes]

  def pack(self, slot, w, acc):
    """after project_bwd of one image (scene.g holds that image's rows only): write record `slot`."""
    scene = self.scene
    ops.exchange_pack(w["g2d"], w["vis"], acc[:, 9:11], scene.stats["max_scale_px"], self.groups, self.record(slot),
                      self.cap, self.workspace, clear_acc=acc, clear_vis=w["vis"])

  # ---- the exchange --------------------------------------------------------------------------------------
  def record_ptrs(self):
    """addresses of the step's records in image order: image j = rank + world·slot (batch[rank::world])."""
    out = []
    for slot in range(self.k):
      for r in range(self.world):
        if self.transport == "peer":
          out.append(self.peer_ptrs[r] + self._local_offset(slot))
        elif self.transport == "nccl":
          out.append(self.gathered.data_ptr() + r * self.k * self.rec_bytes + slot * self.rec_bytes)
        else:
          out.append(self.buf.data_ptr() + slot * self.rec_bytes)
    return out

  def fold(self, extra_ptrs=None):
    """make every rank's records visible, then fold them in image order into scene.g / scene.visible / scene.stats."""
    scene, c = self.scene, self.scene.config
    if self.transport == "peer":
      # stream-ordered barrier: every rank's pack kernels have completed before any rank's fold starts
      self._barrier(0)
    elif self.transport == "nccl":
      dist.all_gather_into_tensor(self.gathered, self.buf, group=self.group)
    scene._mark("exchange_wait")  # barrier / all-gather, including the wait for the slowest rank's image
    ptrs = extra_ptrs if extra_ptrs is not None else self.record_ptrs()
    ops.exchange_fold(self.n, ptrs, self.cap, self.groups, c.split_alpha, c.prune_alpha, scene.stats,
                      scene_visible=scene.visible, stats_rows=self.stats_rows, grad_rows=self.grad_rows)
    if self.transport == "peer":
      self.parity ^= 1

  def counts(self):
    """(visible rows, overflow flag) of this rank's records of the LAST folded step — host sync, off the hot path."""
    out = []
    for slot in range(self.k):
      h = self.record(slot, self.parity ^ 1)[:8].view(torch.int32).cpu().tolist()
      out.append((h[0], bool(h[1])))
    return out

  def bytes_per_image(self, count):
    """bytes a peer actually reads of one record with `count` visible rows (the three bit/rank planes, idx, aux, rows)."""
    words = (self.n + 31) // 32
    return 8 + 3 * 4 * words + count * (4 + 16 + 4 * self.row_floats)
