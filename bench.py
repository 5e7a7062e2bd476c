#!/usr/bin/env python
"""bench.py — throughput of the Gaussian-splatting training hot path (BASELINE.json metric).

  python bench.py --gpus N --steps K --warmup W            our arm (libgsb200 on B200)
  python bench.py --impl reference --steps K --warmup W    CPU arm: the oracle restatement on the host cores

A step = one optimiser iteration of the hot path: per rank one 1297x840 image of the C2 workload
(3M splats, SH degree 3): project+SH -> tile map/sort -> raster fwd -> L1/MSE loss -> raster bwd -> project/SH
bwd -> statistics, then (N>1) NCCL all-reduce of the packed gradients, then the fused visibility-aware
optimiser.  Weak scaling: every rank renders a different camera of the batch (batch = N images).
Prints ONE JSON line on rank 0.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(ROOT, "splat-trainer_b200"))
sys.path.insert(0, ROOT)

WORKLOADS = {
  # name: (points, sh_degree, width, height)
  "c1": (100_000, 0, 640, 480),       # BASELINE.json configs[0]: the CPU-runnable case
  "c2": (3_000_000, 3, 1297, 840),    # configs[1]: the configuration the metric is quoted on
  "c4": (6_000_000, 3, 3840, 2160),   # configs[3]: high-res stress
}
VIEWS_PER_RANK = 4


def parse():
  ap = argparse.ArgumentParser()
  ap.add_argument("--gpus", type=int, default=1)
  ap.add_argument("--steps", type=int, default=20)
  ap.add_argument("--warmup", type=int, default=3)
  ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
  ap.add_argument("--workload", default="c2", choices=sorted(WORKLOADS))
  ap.add_argument("--points", type=int, default=0, help="override the splat count (testing only)")
  ap.add_argument("--no-cpu-baseline", action="store_true")
  ap.add_argument("--scene", default="sh", choices=["sh", "mlp"],
                  help="sh: SH colour per splat (BASELINE.json's configuration); mlp: the reference's default `scene: mlp` "
                       "colour path (point feature 16 + GLO 32 -> ColorModel), same geometry (SURVEY 8f-2; no CPU baseline)")
  ap.add_argument("--exchange", default="auto", choices=["auto", "peer", "nccl", "dense", "local"],
                  help="N>1 gradient/statistics exchange: per-image records over peer memory (auto/peer), records "
                       "all-gathered by NCCL, or the dense SUM all-reduce of the packed gradient buffer")
  return ap.parse_args()


def measured_traffic(stage):
  """dram__bytes_read.sum + dram__bytes_write.sum of the stage's kernels per image from the committed ncu --set full
  capture of this same command (profiles/: cold-cache, serialised launches), or None."""
  path = os.path.join(ROOT, "profiles", "r1p_traffic.json")
  try:
    with open(path) as f:
      return json.load(f)["dram_bytes_per_image"].get(stage)
  except Exception:
    return None


def peaks():
  path = os.path.join(ROOT, "MEASURED_PEAKS.json")
  if os.path.exists(path):
    with open(path) as f:
      return float(json.load(f)["hbm_gbs"]), "measured"
  return 6650.0, "fallback"


class ClockSampler:
  """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md recipe)."""
  Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
       "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

  def __init__(self, index=0):
    self.samples, self.stop, self.index = [], False, index
    self.thread = threading.Thread(target=self.run, daemon=True)

  def run(self):
    while not self.stop:
      try:
        out = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-i", str(self.index)],
                             capture_output=True, text=True, timeout=5).stdout.strip()
        if out:
          self.samples.append([x.strip() for x in out.split(",")])
      except Exception:
        pass
      time.sleep(0.1)

  def __enter__(self):
    self.thread.start()
    return self

  def __exit__(self, *a):
    self.stop = True
    self.thread.join(timeout=6)

  def summary(self):
    sm = sorted(float(s[0]) for s in self.samples if s and s[0].replace(".", "").isdigit())
    mx = [float(s[1]) for s in self.samples if len(s) > 1 and s[1].replace(".", "").isdigit()]
    names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
    reasons = [n for i, n in enumerate(names) if any(len(s) > 2 + i and s[2 + i].lower().startswith("active") for s in self.samples)]
    return dict(sm_mhz=sm[len(sm) // 2] if sm else None, sm_max_mhz=max(mx) if mx else None, reasons=reasons,
                samples=len(self.samples))


# ---- CPU arm: the oracle (test infrastructure) timed as the reported baseline ------------------------------------
class CpuIteration:
  """One full hot-path iteration of the workload in the CPU oracle (oracle/gs_oracle.c, OpenMP)."""

  def __init__(self, g, cams, targets_u8, sh_degree):
    import numpy as np
    from oracle import gs_oracle as go
    self.np, self.go = np, go
    self.orc = go.Oracle("f32")
    self.g = {k: v.copy() for k, v in g.items()}
    self.cams = [go.make_camera(c["T"], c["fx"], c["fy"], c["cx"], c["cy"], c["width"], c["height"], c["near"], c["far"])
                 for c in cams]
    self.targets, self.deg = targets_u8, sh_degree
    self.cfg = go.make_config()
    self.ocfg = go.make_optim_config()
    n = g["position"].shape[0]
    z = lambda *s: np.zeros(s, np.float32)
    self.m1 = dict(position=z(n, 3), log_scaling=z(n, 3), rotation=z(n, 4), alpha_logit=z(n), sh=np.zeros_like(g["sh"]))
    self.m2 = dict(position=z(n), log_scaling=z(n, 3), rotation=z(n), alpha_logit=z(n), sh=z(n))
    self.running, self.step = z(n), z(n)
    self.stats = dict(prune_cost=z(n), split_score=z(n), max_scale_px=z(n), points_in_view=np.zeros(n, np.int16),
                      visibility=z(n), scene_visible=z(n))

  def iteration(self, i):
    o, g, cfg, np = self.orc, self.g, self.cfg, self.np
    cam = self.cams[i % len(self.cams)]
    W, H = cam.width, cam.height
    g2d, depth, rgb = o.project_fwd(cam, cfg, g["position"], g["log_scaling"], g["rotation"], g["alpha_logit"], g["sh"], self.deg)
    o2p, ranges, total, _ = o.tile_map(cfg, g2d, depth, W, H)
    image, fT, nc, _, _ = o.raster_fwd(cfg, g2d, rgb, depth, o2p, ranges, W, H, visibility=False)
    tu8 = self.targets[i % len(self.targets)]
    dimg, sums = o.loss_l1_mse(image, None, 0.0, 10.0, target_u8=tu8)
    # multi-scale SSIM with the reference's default weights (ssim 1.0, 3 levels), like SplatScene.forward_backward
    _, dssim, _ = o.ssim_loss(np.clip(image, 0, 1), tu8.astype(np.float32) / 255.0, levels=3, weight=1.0)
    dimg = dimg + np.where((image >= 0) & (image <= 1), dssim, 0).astype(np.float32)
    acc = o.raster_bwd(cfg, g2d, rgb, o2p, ranges, W, H, fT, nc, dimg)
    # reg_loss with the reference's yaml weights, like SplatScene.forward_backward
    vis = np.ascontiguousarray(acc[:, 11]); d_op = np.ascontiguousarray(acc[:, 5])
    d_depth = np.zeros_like(vis); d_ls = np.zeros((vis.shape[0], 3), np.float32)
    o.reg_loss((0.01, 1.0, 1e-4), g2d, depth, g["log_scaling"], vis, d_op, d_depth, d_ls)
    acc[:, 5] = d_op
    grads = o.project_bwd(cam, cfg, g["position"], g["log_scaling"], g["rotation"], g["alpha_logit"], g["sh"], self.deg,
                          g2d, acc[:, 0:6], acc[:, 6:9], d_depth, row_vis=acc[:, 11])
    grads["log_scaling"] = grads["log_scaling"] + d_ls
    st = self.stats
    o.accumulate_stats(None, g2d, np.ascontiguousarray(acc[:, 11]), np.ascontiguousarray(acc[:, 9:11]), 0.01, 0.1, st)
    vis = st["scene_visible"]
    o.optim_vis_update(self.ocfg, vis, self.running, self.step)
    lr = dict(position=0.05, log_scaling=0.02, rotation=0.01, alpha_logit=0.05, sh=0.01)
    kind = dict(position=2, log_scaling=0, rotation=1, alpha_logit=0, sh=1)
    post = dict(position=(0, 0, 0), log_scaling=(2, -8.0, 8.0), rotation=(1, 0, 0), alpha_logit=(0, 0, 0), sh=(0, 0, 0))
    bls, bq = g["log_scaling"].copy(), g["rotation"].copy()
    for name in ("position", "log_scaling", "rotation", "alpha_logit", "sh"):
      o.optim_step(self.ocfg, kind[name], lr[name], g[name], grads[name], self.m1[name], self.m2[name], vis, self.running,
                   self.step, bls, bq, post[name][0], post[name][1], post[name][2], 1)
    vis[:] = 0
    return float(sums[1])


def make_workload(args):
  from gsb200 import synthetic as syn
  n, deg, W, H = WORKLOADS[args.workload]
  if args.points:
    n = args.points
  g = syn.garden_scene(n, deg, seed=0)
  cams = syn.ring_cameras(185, W, H, seed=0)  # garden has 185 views
  return n, deg, W, H, g, cams


def algorithmic_bytes(n, d, v, o, p, t, v_opt):
  """SURVEY.md §8(d) byte model per training image (+ the optimiser over the rows it touches)."""
  return dict(project_fwd=12 * n + 4 * (d - 3) * v + 44 * v,
              tile_map=28 * v + 4 * v + 12 * o + 8 * t + 6 * 24 * o + 8 * o,  # map + 6-pass 64-bit-key sort model
              raster_fwd=40 * o + 16 * p,
              raster_bwd=40 * o + 28 * p + 36 * v,
              project_bwd=36 * v + 8 * d * v,
              stats=40 * v,
              optim=7 * 4 * d * v_opt)


def run_reference(args):
  """--impl reference: the path's CPU implementation on the host cores.  The reference's own kernels live in the
  absent taichi-splatting package (unbuildable here: DESIGN.md "Oracle"), so this is the oracle port."""
  import numpy as np
  rank = int(os.environ.get("RANK", "0"))
  if rank != 0:
    return
  n, deg, W, H, g, cams = make_workload(args)
  rng = np.random.default_rng(1)
  views = cams[:VIEWS_PER_RANK]
  targets = [rng.integers(0, 256, (H, W, 3), dtype=np.uint8) for _ in views]
  cpu = CpuIteration(g, views, targets, deg)
  for i in range(args.warmup):
    cpu.iteration(i)
  t0 = time.perf_counter()
  for i in range(args.steps):
    cpu.iteration(args.warmup + i)
  dt = time.perf_counter() - t0
  cores = len(os.sched_getaffinity(0))
  value = args.steps / dt
  unit = "iters/s"
  line = {"impl": "reference", "metric": "train_iters_per_sec", "value": value, "unit": unit, "n_gpus": args.gpus,
          "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True,
          "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
          "config": {"workload": f"{args.workload}: {n} splats SH{deg} {W}x{H}, 1 image/step on the host CPU",
                     "images_per_step": 1},
          "cpu_baseline": {"value": value, "unit": unit, "cores": cores, "kind": "port",
                           "sample": f"{args.steps} full iterations of the same workload (random uint8 targets)"},
          "e2e": {"value": value, "unit": unit, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
  emit(line)


def run_b200(args):
  import numpy as np
  import torch
  import torch.distributed as dist
  from gsb200 import lib as L
  from gsb200.scene import SceneConfig, SplatScene
  from gsb200.trainer import StageTimer, Trainer

  world = int(os.environ.get("WORLD_SIZE", "1"))
  rank = int(os.environ.get("RANK", "0"))
  local = int(os.environ.get("LOCAL_RANK", "0"))
  torch.cuda.set_device(local)
  if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
  L.load()
  n, deg, W, H, g, cams = make_workload(args)
  from gsb200 import synthetic as syn
  # this rank's views of every step's batch (batch = world images, rank r renders batch[r::world])
  views = [cams[(rank + world * j) % len(cams)] for j in range(VIEWS_PER_RANK)]
  lcams = [L.make_camera(c["T"], c["fx"], c["fy"], c["cx"], c["cy"], W, H, c["near"], c["far"]) for c in views]
  # targets: renders of a perturbed copy of the scene (seed 1), stored as uint8 like the reference's dataset
  tgt_scene = SplatScene(syn.perturbed(g, seed=1), SceneConfig(sh_degree=deg))
  host_targets = []
  for cam in lcams:
    img = (tgt_scene.render(cam)["image"] * 255.0 + 0.5).to(torch.uint8)
    host_targets.append(img.cpu().pin_memory())
  del tgt_scene
  torch.cuda.empty_cache()
  dev_targets = [t.cuda() for t in host_targets]
  if args.scene == "mlp":
    gm = {k: v for k, v in g.items() if k != "sh"}
    gm["feature"] = np.random.default_rng(7).normal(0, 1, (n, 16)).astype(np.float32)
    scene = SplatScene(gm, SceneConfig(sh_degree=deg, color_mlp=True, num_images=VIEWS_PER_RANK * world))
  else:
    scene = SplatScene(g, SceneConfig(sh_degree=deg))
  if args.exchange == "local":  # N=1 only: run the pack/fold kernels of the DP exchange on one GPU (for profiling them)
    from gsb200.exchange import ImageExchange
    local_k = int(os.environ.get("GS_LOCAL_IMAGES", "1"))  # K different images per step packed + folded on one GPU
    trainer = Trainer(scene, 1, 0, exchange=ImageExchange(scene, 1, 0, transport="local", images_per_rank=local_k))
  else:
    trainer = Trainer(scene, world, rank, exchange=args.exchange)

  def batch_for(i, targets):
    # the step's full batch has `world` images; only this rank's entry is materialised (others are None-padded)
    if args.exchange == "local":
      return [(lcams[(i + q) % len(lcams)], targets[(i + q) % len(targets)]) for q in range(trainer.exchange.k)]
    b = [None] * world
    b[rank] = (lcams[i % len(lcams)], targets[i % len(targets)])
    if args.scene == "mlp":  # the image's GLO row
      b[rank] = b[rank] + (rank + world * (i % len(lcams)),)
    return b

  def step(i, targets, read_loss=False):
    return trainer.training_step(batch_for(i, targets), read_loss=read_loss)

  def barrier():
    if world > 1:
      dist.barrier()
    torch.cuda.synchronize()

  for i in range(max(args.warmup, 3)):
    step(i, dev_targets)
  torch.cuda.synchronize()
  if scene.overflowed():
    scene.grow_overlaps(2.0)
    step(0, dev_targets)
  exchange_info = None
  if trainer.exchange is not None:
    # size the records from the rows the warm-up images really blended (one host read at start-up, all ranks agree)
    rows = trainer.calibrate_exchange(margin=1.5)
    for i in range(2):
      step(i, dev_targets)
    torch.cuda.synchronize()
    ex = trainer.exchange
    exchange_info = {"transport": ex.transport, "record_capacity_rows": ex.cap, "rows_per_image": rows,
                     "bytes_read_per_peer_image": ex.bytes_per_image(rows),
                     "dense_allreduce_bytes": int(scene.grads.numel() * 4)}
    if getattr(ex, "fallback_reason", None):
      exchange_info["peer_unavailable"] = ex.fallback_reason[:200]
  elif world > 1:
    exchange_info = {"transport": "dense-allreduce", "dense_allreduce_bytes": int(scene.grads.numel() * 4)}

  # ---- device-resident throughput (value) with per-stage CUDA events ----------------------------------------------
  timer = StageTimer()
  scene.timer = timer
  barrier()
  e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
  with ClockSampler(local) as clocks:
    e0.record()
    h0 = time.perf_counter()
    for i in range(args.steps):
      step(args.warmup + i, dev_targets)
    host_enqueue_ms = 1e3 * (time.perf_counter() - h0) / args.steps  # CPU time to enqueue one step (no sync inside)
    e1.record()
    barrier()
  ms = e0.elapsed_time(e1)
  stages = {k: float(np.mean(v)) for k, v in timer.collect().items()}
  scene.timer = None
  overflow = scene.overflowed()

  # ---- end to end through the public API with HOST buffers --------------------------------------------------------
  for i in range(3):  # untimed: first use of the upload stream and its two device buffers
    step(i, host_targets, read_loss=True)
  barrier()
  t0 = time.perf_counter()
  per_step = []
  for i in range(args.steps):
    loss = step(args.warmup + i, host_targets, read_loss=True)
    per_step.append(time.perf_counter())
  torch.cuda.synchronize()
  if world > 1:
    dist.barrier()
  e2e_s = time.perf_counter() - t0

  if os.environ.get("GS_BENCH_DEBUG"):
    print("e2e per-step ms:", [round(1e3 * (b - a), 3) for a, b in zip([t0] + per_step[:-1], per_step)], file=sys.stderr)
  if world > 1:
    t = torch.tensor([ms, e2e_s], device="cuda", dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms, e2e_s = float(t[0].item()), float(t[1].item())

  w = scene._work
  counters = w["tile"].counters.cpu().tolist()
  O = int(counters[0])
  V = int((w["g2d"][:, 5] > 0).sum().item())
  v_opt = int(scene._vis_idx.count.item())  # rows the last optimiser step touched (device-side visible.nonzero())
  P, T, D = W * H, ((W + 15) // 16) * ((H + 15) // 16), 11 + (16 if args.scene == "mlp" else 3 * (deg + 1) ** 2)
  bytes_ = algorithmic_bytes(n, D, V, O, P, T, v_opt)
  hbm, which = peaks()
  dom = max(("raster_fwd", "raster_bwd", "tile_map", "project_fwd", "project_bwd", "optim"), key=lambda k: stages.get(k, 0.0))
  dom_ms = stages[dom]
  achieved = bytes_[dom] / (dom_ms * 1e-3) / 1e9
  iter_bytes = sum(bytes_.values())
  ms_step = ms / args.steps
  # kernels of libgsb200 per image: project_fwd, tile map (live compaction, depth sort = histogram + 4 passes,
  # super-tile pairs, their sort = histogram + passes, segment tables, expand count, chunk prefix, expand write),
  # raster_fwd, loss, raster_bwd, project_bwd, statistics; per step: optimiser begin + 5 parameter groups
  n_super = ((W + 127) // 128) * ((H + 127) // 128)
  super_passes = (max(1, (n_super - 1).bit_length()) + 7) // 8
  launches_image = 1 + (1 + 5 + 1 + 1 + super_passes + 4) + 5
  launches = launches_image + 6
  if trainer.exchange is not None:
    launches += 1 + 2  # pack = 2 kernels in place of the statistics pass; fold = 2 kernels (NCCL's own not counted)

  if rank == 0:
    line = {
      # the reference's step counter advances once per IMAGE (trainer/trainer.py:532, step_rate :639-646), so the
      # whole-job value is images/s; optimiser steps/s is reported beside it
      "metric": "train_iters_per_sec", "value": world * args.steps / (ms * 1e-3), "unit": "iters/s",
      "optimizer_steps_per_sec": args.steps / (ms * 1e-3),
      "raster_fwd_bwd_mpix_per_sec": world * P / ((stages["raster_fwd"] + stages["raster_bwd"]) * 1e-3) / 1e6,
      "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms_step,
      "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
      "config": {"workload": f"{args.workload}: {n} splats {'MLP colour (feature 16 + GLO 32)' if args.scene == 'mlp' else f'SH{deg}'} {W}x{H}, camera batch = {world} (1 image per GPU per step), "
                             "fwd + loss (MSE 10 + 3-level SSIM 1.0 + reg) + bwd + fused visibility-aware LaProp step", "images_per_step": world,
                 "l2": f"inputs larger than L2 (params+grads+moments {4 * 4 * D * n / 1e9:.1f} GB, overlap list {4 * O / 1e6:.0f} MB)",
                 "parallelism": f"dp{world}", "V_in_view": V, "O_overlaps": O, "V_optimised": v_opt, "overflow": overflow},
      "stages_ms": stages, "host_enqueue_ms_per_step": host_enqueue_ms,
      "roofline": {"bound": "hbm", "kernel": dom, "achieved": achieved, "peak": hbm, "unit": "GB/s", "frac": achieved / hbm,
                   "traffic": measured_traffic(dom) if args.workload == "c2" and not args.points and args.scene == "sh" else None,
                   "traffic_source": "profiles/r1p_traffic.json (ncu --set full of the c2 command)",
                   "peak_source": which, "algorithmic_bytes": bytes_[dom],
                   "iteration": {"algorithmic_bytes": iter_bytes, "achieved": iter_bytes / (ms_step * 1e-3) / 1e9,
                                 "frac": iter_bytes / (ms_step * 1e-3) / 1e9 / hbm}},
      "e2e": {"value": world * args.steps / e2e_s, "unit": "iters/s", "h2d_bytes_per_step": int(host_targets[0].numel()),
              "d2h_bytes_per_step": 8},
      "gpu_launches": launches * args.steps,
      "clocks": clocks.summary(),
      "last_loss_mse": loss[1] / (P * 3),
    }
    if exchange_info:
      if trainer.exchange is not None:
        exchange_info["overflow"] = any(o for _, o in trainer.exchange.counts())
      line["exchange"] = exchange_info
    if world == 1 and not args.no_cpu_baseline and args.scene == "sh":
      cpu = CpuIteration(g, views[:1], [host_targets[0].numpy()], deg)
      cpu.iteration(0)  # warm-up (page faults, OpenMP pool)
      t0 = time.perf_counter()
      reps = 2
      for i in range(reps):
        cpu.iteration(i)
      dt = (time.perf_counter() - t0) / reps
      line["cpu_baseline"] = {"value": 1.0 / dt, "unit": "iters/s", "cores": len(os.sched_getaffinity(0)), "kind": "port",
                              "sample": f"{reps} full iterations of the same workload (after 1 warm-up) in the OpenMP oracle"}
    emit(line)
  if world > 1:
    dist.destroy_process_group()


_REAL_STDOUT = sys.stdout


def emit(line):
  """the ONE JSON line goes to the real stdout; everything else any library prints (NCCL's version banner, torchrun
  notices) was sent to stderr by the redirect in __main__."""
  _REAL_STDOUT.write(json.dumps(line) + "\n")
  _REAL_STDOUT.flush()


if __name__ == "__main__":
  _REAL_STDOUT = os.fdopen(os.dup(1), "w")
  os.dup2(2, 1)
  a = parse()
  if a.impl == "reference":
    run_reference(a)
  else:
    run_b200(a)
