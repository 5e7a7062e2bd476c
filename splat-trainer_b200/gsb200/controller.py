"""TargetController over a SplatScene — the reference's densify / prune schedule for the fused path.

Mirrors controller/target_controller.py:20-126 (TargetConfig, find_next_densify, target_points, step) and
config/controller/target.yaml (prune_rate .025, target_count_t .8, min_views 5, max_scale_px 200, interval 50 -> 100 at
t >= .1 -> 200 at t >= .5).  The statistics live in SplatScene.stats (the reference's PointState, filled by the fused
statistics pass); decisions are radix selects (SplatScene.find_split_prune), the compaction runs in place
(SplatScene.split_and_prune).  Under camera-batch data parallelism every rank holds identical statistics after
Trainer.sync_statistics() and draws the same split axes from `generator`, so the replicas stay identical."""
import dataclasses
import math

import torch


def piecewise(start, steps, scale=1.0):
  """config/__init__.py:104-122 Piecewise: value in force at progress t."""
  def f(t):
    value = start
    for t_min, nxt in steps:
      if t < t_min:
        return value * scale
      value = nxt
    return value * scale
  return f


def smoothstep(t, a, b):
  """config/__init__.py:127-131"""
  t = min(max(t, 0.0), 1.0)
  return a + (b - a) * (3 * t ** 2 - 2 * t ** 3)


@dataclasses.dataclass
class TargetConfig:
  prune_rate: float = 0.025
  target_count_t: float = 0.8
  min_views: int = 5
  max_scale_px: float = 200.0
  min_split_px: float = 0.0
  densify_prune_interval: object = dataclasses.field(default_factory=lambda: piecewise(50, [(0.1, 100), (0.5, 200)]))


class TargetController:
  def __init__(self, config, scene, target_points, total_steps, step=0, start_points=None, seed=0, trainer=None):
    """trainer: the gsb200 Trainer (needed under data parallelism for sync_statistics), or None."""
    self.config, self.scene, self.trainer = config, scene, trainer
    self.max_points, self.total_steps = int(target_points), int(total_steps)
    self.start_points = start_points or scene.num_points
    self.generator = torch.Generator(device=scene.device).manual_seed(seed)
    self.next_densify = self.find_next_densify(step)
    self.history = []
    self.record_events, self.events = False, []  # bench.py: CUDA events around every densification

  def interval(self, step):
    v = self.config.densify_prune_interval
    return int(v(step / self.total_steps) if callable(v) else v)

  def find_next_densify(self, step):
    k = self.interval(step)
    nxt = step + k
    return nxt if nxt + k < self.total_steps else None  # the last densify leaves two intervals of plain training

  def target_points(self, step):
    t = step / (self.config.target_count_t * self.total_steps)
    return int(smoothstep(t, self.start_points, self.max_points))

  def step(self, step):
    """call once per optimiser step with the trainer's image counter; returns (old_n, new_n) when it densified."""
    if self.next_densify is None or step < self.next_densify:
      return None
    ev0 = None
    if self.record_events:
      ev0 = torch.cuda.Event(enable_timing=True)
      ev0.record()
    if self.trainer is not None:
      self.trainer.sync_statistics()
    scene, c = self.scene, self.config
    t = min(max(step / self.total_steps, 0.0), 1.0)
    n_split_draws = scene.num_points  # one uniform per row, indexed by the split rows: identical on every rank
    uniform = torch.rand((n_split_draws,), device=scene.device, generator=self.generator)
    out = scene.densify(t, self.target_points(step), uniform_rows=uniform, prune_rate=c.prune_rate, min_views=c.min_views,
                        max_scale_px=c.max_scale_px, min_split_px=c.min_split_px)
    self.next_densify = self.find_next_densify(step)
    self.history.append((step,) + tuple(out))
    if ev0 is not None:
      ev1 = torch.cuda.Event(enable_timing=True)
      ev1.record()
      self.events.append((ev0, ev1))
    return out

  def state_dict(self):
    """controller/target_controller.py:68-69: dict(points=PointState, start_points)"""
    return dict(points={k: v.clone() for k, v in self.scene.stats.items()}, start_points=self.start_points,
                next_densify=-1 if self.next_densify is None else int(self.next_densify),
                generator=self.generator.get_state())

  def load_state_dict(self, state):
    for k, v in state["points"].items():
      self.scene.stats[k].copy_(v)
    self.start_points = int(state["start_points"])
    if "next_densify" in state:
      nd = int(state["next_densify"])
      self.next_densify = None if nd < 0 else nd
    if "generator" in state:
      self.generator.set_state(state["generator"].cpu())
