#!/usr/bin/env python
"""bench.py — throughput of the Gaussian-splatting training hot path (BASELINE.json metric).

  python bench.py --gpus N --steps K --warmup W            our arm (libgsb200 on B200)
  python bench.py --impl reference --steps K --warmup W    CPU arm: the oracle restatement on the host cores
  python bench.py --workload c5 [--gpus N]                 configs[4]: 500k -> 5M growth with densification in the loop

A step = one optimiser iteration of the hot path: per rank ONE image of the workload — project + SH -> tile map / sort ->
raster forward -> loss (MSE 10 + 3-level SSIM 1.0 + reg_loss, the reference's defaults) -> raster backward -> projection /
SH backward -> statistics — then, for N > 1, the exchange of the per-image gradient records over NVLink peer memory
(folded in image order on every rank), then the fused visibility-aware LaProp step.  Weak scaling: every rank renders a
different camera of the batch (batch = N images).  The timed window of K steps is repeated `--windows` times inside the
run, every window starting from the same training state (restored outside the timed region); `value` is the median
window.  Prints ONE JSON line on rank 0.

Workloads (gsb200.synthetic.WORKLOADS): c2 = BASELINE.json configs[1] generated exactly as SURVEY.md 8(d) writes it (the
default); c2-sparse = round 1's generator (opaque small splats: ~3 % of the in-view rows blend); c2-trained = translucent
splats (most in-view rows blend); c1, c4 = configs[0], configs[3]; c5 = configs[4].
"""
import argparse
import hashlib
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(ROOT, "splat-trainer_b200"))
sys.path.insert(0, ROOT)

VIEWS_PER_RANK = 4
LOSS = "fwd + loss (MSE 10 + 3-level SSIM 1.0 + reg) + bwd + fused visibility-aware LaProp step"


def parse():
  ap = argparse.ArgumentParser()
  ap.add_argument("--gpus", type=int, default=1)
  ap.add_argument("--steps", type=int, default=20)
  ap.add_argument("--warmup", type=int, default=3)
  ap.add_argument("--windows", type=int, default=5, help="the --steps window is timed this many times; value = the median window")
  ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
  ap.add_argument("--workload", default="c2", choices=["c1", "c2", "c2-sparse", "c2-trained", "c4", "c5"])
  ap.add_argument("--points", type=int, default=0, help="override the splat count (testing only)")
  ap.add_argument("--no-cpu-baseline", action="store_true")
  ap.add_argument("--scene", default="sh", choices=["sh", "mlp"],
                  help="sh: SH colour per splat (BASELINE.json's configuration); mlp: the reference's default `scene: mlp` "
                       "colour path (point feature 16 + GLO 32 -> ColorModel), same geometry (SURVEY 8f-2; no CPU baseline)")
  ap.add_argument("--exchange", default="auto", choices=["auto", "peer", "nccl", "dense", "local"],
                  help="N>1 gradient/statistics exchange: per-image records over peer memory (auto/peer), records "
                       "all-gathered by NCCL, or the dense SUM all-reduce of the packed gradient buffer")
  ap.add_argument("--shard", default="auto", choices=["auto", "rows", "replicated"],
                  help="N>1: optimiser sharded by row slice (each rank steps N/W rows and pushes them to its peers) or replicated")
  ap.add_argument("--raster", default="async", choices=["async", "gather"])
  ap.add_argument("--exchange-cap", type=int, default=0, help="testing: initial rows per exchange record (0 = a quarter of the scene)")
  ap.add_argument("--c5-steps", type=int, default=2000, help="c5: optimiser steps of the whole 500k -> 5M run")
  return ap.parse_args()


def build_id():
  """hash of the CUDA sources: ties a committed ncu capture to the build it was taken from"""
  h = hashlib.sha1()
  d = os.path.join(ROOT, "splat-trainer_b200", "csrc")
  for f in sorted(os.listdir(d)):
    if f.endswith((".cu", ".cuh")):
      h.update(open(os.path.join(d, f), "rb").read())
  return h.hexdigest()[:12]


def measured_traffic(workload):
  """dram__bytes_read.sum + dram__bytes_write.sum per image and per stage from the committed `ncu --set full` capture
  of this command (profiles/r2_traffic_<workload>.json), used only when it was taken from THIS build."""
  path = os.path.join(ROOT, "profiles", f"r2_traffic_{workload}.json")
  try:
    with open(path) as f:
      t = json.load(f)
    return t if t.get("build_id") == build_id() else None
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
      time.sleep(0.05)

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


def workload_string(args, wl, world):
  colour = "MLP colour (feature 16 + GLO 32)" if args.scene == "mlp" else f"SH{wl['sh_degree']}"
  return f"{wl['description']}; {colour}; 1 image per step per arm unit; {LOSS}"


# ---- CPU arm: the oracle (test infrastructure) timed as the reported baseline ------------------------------------
def host_threads():
  return len(os.sched_getaffinity(0))


def make_cpu_scene(wl):
  """the workload in the CPU oracle (oracle/gs_train.py, OpenMP over all the host cores this process may use —
  set explicitly: torchrun exports OMP_NUM_THREADS=1)."""
  from gsb200.scene import SceneConfig
  from oracle import gs_oracle as go
  from oracle import gs_train as gt
  cfg = SceneConfig(sh_degree=wl["sh_degree"])
  lrs = dict(position=cfg.lr_position, log_scaling=cfg.lr_log_scaling, rotation=cfg.lr_rotation,
             alpha_logit=cfg.lr_alpha_logit, sh=cfg.lr_sh)
  sc = gt.OracleScene(wl["gaussians"], wl["sh_degree"], lrs, w_l1=cfg.w_l1, w_mse=cfg.w_mse, w_ssim=cfg.w_ssim,
                      ssim_levels=cfg.ssim_levels, reg_weights=cfg.reg_weights, blur_cov=cfg.blur_cov)
  sc.orc.set_threads(host_threads())
  return sc, go, gt


def oracle_cameras(go, cams, W, H):
  return [go.make_camera(c["T"], c["fx"], c["fy"], c["cx"], c["cy"], W, H, c["near"], c["far"]) for c in cams]


def run_reference(args):
  """--impl reference: the path's CPU implementation on the host cores.  The reference's own kernels live in the
  absent taichi-splatting package (unbuildable here: DESIGN.md "Oracle"), so this is the oracle port — same workload,
  same targets (rendered from the workload's target scene, stored as uint8), visibility on, every host core."""
  import numpy as np
  rank = int(os.environ.get("RANK", "0"))
  if rank != 0:
    return
  from gsb200 import synthetic as syn
  wl = syn.workload("c2" if args.workload == "c5" else args.workload, points=args.points)
  W, H = wl["width"], wl["height"]
  sc, go, gt = make_cpu_scene(wl)
  views = wl["cameras"][:VIEWS_PER_RANK]
  ocams = oracle_cameras(go, views, W, H)
  tgt = gt.OracleScene(wl["target_gaussians"], wl["sh_degree"], sc.lrs)
  tgt.orc.set_threads(host_threads())
  targets = [np.clip(tgt.render(oc) * 255.0 + 0.5, 0, 255).astype(np.uint8) for oc in ocams]
  del tgt
  for i in range(args.warmup):
    sc.training_step([(ocams[i % len(ocams)], targets[i % len(ocams)])])
  t0 = time.perf_counter()
  for i in range(args.steps):
    k = (args.warmup + i) % len(ocams)
    sc.training_step([(ocams[k], targets[k])])
  dt = time.perf_counter() - t0
  cores = host_threads()
  value = args.steps / dt
  unit = "iters/s"
  world = int(os.environ.get("WORLD_SIZE", "1"))
  line = {"impl": "reference", "metric": "train_iters_per_sec", "value": value, "unit": unit, "n_gpus": args.gpus,
          "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True,
          "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
          "config": {"workload": workload_string(args, wl, world), "images_per_step": 1},
          "cpu_baseline": {"value": value, "unit": unit, "cores": cores, "kind": "port",
                           "sample": f"{args.steps} full iterations of the same workload, same targets, after {args.warmup} warm-up, "
                                     f"OpenMP oracle on {sc.orc.get_threads()} threads"},
          "e2e": {"value": value, "unit": unit, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
  emit(line)


def algorithmic_bytes(n, d, v, o, p, t, v_blend):
  """SURVEY.md §8(d) byte model per training image.  The backward / statistics / optimiser terms are charged for the
  rows a pixel actually blended (v_blend), not for every in-view row: the kernels legitimately never touch the others."""
  return dict(project_fwd=12 * n + 4 * (d - 3) * v + 44 * v,
              tile_map=28 * v + 4 * v + 12 * o + 8 * t + 6 * 24 * o + 8 * o,  # map + 6-pass 64-bit-key sort model
              raster_fwd=40 * o + 16 * p,
              raster_bwd=40 * o + 28 * p + 36 * v_blend,
              project_bwd=36 * v_blend + 8 * d * v_blend,
              stats=40 * v_blend,
              optim=7 * 4 * d * v_blend)


def median(xs):
  s = sorted(xs)
  return s[len(s) // 2]


def setup_b200(args, workload_name):
  """process group, library, workload, scene, targets, trainer — shared by the steady-state bench and the c5 run"""
  import numpy as np
  import torch
  import torch.distributed as dist
  from gsb200 import lib as L
  from gsb200 import ops
  from gsb200 import synthetic as syn
  from gsb200.scene import SceneConfig, SplatScene

  world = int(os.environ.get("WORLD_SIZE", "1"))
  rank = int(os.environ.get("RANK", "0"))
  local = int(os.environ.get("LOCAL_RANK", "0"))
  torch.cuda.set_device(local)
  if world > 1 and not dist.is_initialized():
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
  L.load()
  # the knn3-derived scales of the SURVEY 8(d) scene come from the library's own k-NN kernel (gs_knn_scale)
  knn = lambda p: ops.estimate_scale(torch.as_tensor(p, device="cuda"), 3).cpu().numpy()
  wl = syn.workload(workload_name, points=args.points, knn=knn)
  W, H = wl["width"], wl["height"]
  cams = wl["cameras"]
  views = [cams[(rank + world * j) % len(cams)] for j in range(VIEWS_PER_RANK)]
  lcams = [L.make_camera(c["T"], c["fx"], c["fy"], c["cx"], c["cy"], W, H, c["near"], c["far"]) for c in views]
  # targets: renders of the workload's target scene, stored as uint8 like the reference's dataset
  tgt_scene = SplatScene(wl["target_gaussians"], SceneConfig(sh_degree=wl["sh_degree"]))
  host_targets = []
  for cam in lcams:
    for _ in range(4):
      img = tgt_scene.render(cam)["image"]
      if not tgt_scene.overflowed():
        break
      tgt_scene.grow_overlaps(2.0)
    host_targets.append((img * 255.0 + 0.5).to(torch.uint8).cpu().pin_memory())
  del tgt_scene
  torch.cuda.empty_cache()
  return dict(world=world, rank=rank, local=local, wl=wl, W=W, H=H, views=views, lcams=lcams, host_targets=host_targets,
              dev_targets=[t.cuda() for t in host_targets])


def replicas_identical(scene, world, trainer=None):
  """outside the timed region: every rank holds bit-identical parameters and moments (MIN == MAX of a 64-bit digest).
  Row-sharded optimiser: the moments of a row live on its owner, so the slices are gathered first — the parameters
  were pushed to every replica by the steps themselves."""
  import torch
  import torch.distributed as dist
  if world <= 1:
    return None
  if trainer is not None and trainer.sharded:
    trainer.sync_statistics()
  digest = []
  for t in (scene.params[:scene._total], scene.exp_avg[:scene._total], scene.exp_avg_sq_flat, scene.running_vis):
    digest.append(t.view(torch.int32).to(torch.int64).sum())
  lo = torch.stack(digest).clone()
  hi = lo.clone()
  dist.all_reduce(lo, op=dist.ReduceOp.MIN)
  dist.all_reduce(hi, op=dist.ReduceOp.MAX)
  return bool(torch.equal(lo, hi))


def replica_diagnosis(scene, trainer, rank, step):
  """GS_C5_DIAG=1 (debugging aid, not a bench mode): after every step gather the row slices and compare every state tensor
  across the ranks; prints the first divergence (tensor, differing elements, first row) and returns True."""
  import torch
  import torch.distributed as dist
  if trainer.exchange is not None and trainer.exchange.stale():
    return False  # right after a densification: compared after the next step
  trainer.sync_statistics()
  named = dict(params=scene.params[:scene._total], exp_avg=scene.exp_avg[:scene._total], exp_avg_sq=scene.exp_avg_sq_flat,
               running_vis=scene.running_vis, opt_step=scene.opt_step, grads=scene.grads,
               **{"stats." + k: v.float() for k, v in scene.stats.items()})
  for k in ATTR_NAMES:
    named["p." + k] = scene.p[k].reshape(scene.n, -1)
    named["m1." + k] = scene.m1[k].reshape(scene.n, -1)
    named["m2." + k] = scene.m2[k].reshape(scene.n, -1)
  bad = []
  for name, t in named.items():
    lo, hi = t.clone(), t.clone()
    dist.all_reduce(lo, op=dist.ReduceOp.MIN); dist.all_reduce(hi, op=dist.ReduceOp.MAX)
    diff = (lo != hi)
    cnt = int(diff.sum().item())
    if cnt:
      first = int(diff.reshape(diff.shape[0], -1).any(dim=1).nonzero()[0].item()) if diff.dim() > 1 else int(diff.nonzero()[0].item())
      bad.append((name, cnt, first, tuple(t.shape)))
  if bad and rank == 0:
    ex = trainer.exchange
    print(f"[diag] step {step}: n={scene.n} slice_len={ex.slice_len} rows={ex.grad_rows} densified={len(trainer.controller.history) if trainer.controller else 0} "
          f"overflow_events={trainer.overflow_events} diverged: {bad}", file=sys.stderr, flush=True)
  return bool(bad)


ATTR_NAMES = ("position", "log_scaling", "rotation", "alpha_logit", "sh")


def run_b200(args):
  import numpy as np
  import torch
  import torch.distributed as dist
  from gsb200 import lib as L
  from gsb200.scene import SceneConfig, SplatScene
  from gsb200.trainer import StageTimer, Trainer

  S = setup_b200(args, args.workload)
  world, rank, local, wl, W, H = S["world"], S["rank"], S["local"], S["wl"], S["W"], S["H"]
  lcams, host_targets, dev_targets = S["lcams"], S["host_targets"], S["dev_targets"]
  n, deg, g = wl["n"], wl["sh_degree"], wl["gaussians"]
  if args.scene == "mlp":
    gm = {k: v for k, v in g.items() if k != "sh"}
    gm["feature"] = np.random.default_rng(7).normal(0, 1, (n, 16)).astype(np.float32)
    scene = SplatScene(gm, SceneConfig(sh_degree=deg, color_mlp=True, num_images=VIEWS_PER_RANK * world, raster=args.raster))
  else:
    scene = SplatScene(g, SceneConfig(sh_degree=deg, raster=args.raster))
  if args.exchange == "local":  # N=1 only: run the pack/fold kernels of the DP exchange on one GPU (for profiling them)
    from gsb200.exchange import ImageExchange
    local_k = int(os.environ.get("GS_LOCAL_IMAGES", "1"))  # K different images per step packed + folded on one GPU
    trainer = Trainer(scene, 1, 0, exchange=ImageExchange(scene, 1, 0, transport="local", images_per_rank=local_k), check_every=0)
  else:
    trainer = Trainer(scene, world, rank, exchange=args.exchange, check_every=0, shard=args.shard)  # overflow is checked around the timed region

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

  warm = max(args.warmup, 3)
  for i in range(warm + len(lcams)):  # every view of the rotation once: capacities see the largest of them
    step(i, dev_targets)
    torch.cuda.synchronize()
    trainer.check_overflow()
  exchange_info = None
  if trainer.exchange is not None:
    # size the records from the rows the warm-up images really blended (one host read at start-up, all ranks agree)
    rows = trainer.calibrate_exchange(margin=1.5)
    for i in range(2):
      step(i, dev_targets)
    torch.cuda.synchronize()
    ex = trainer.exchange
    exchange_info = {"transport": ex.transport, "optimizer_shard": getattr(ex, "shard", "replicated"),
                     "record_capacity_rows": ex.cap, "rows_per_image": rows,
                     "bytes_read_per_peer_image": ex.bytes_per_image(rows),
                     "dense_allreduce_bytes": int(scene.grads.numel() * 4)}
    if getattr(ex, "fallback_reason", None):
      exchange_info["peer_unavailable"] = ex.fallback_reason[:200]
  elif world > 1:
    exchange_info = {"transport": "dense-allreduce", "dense_allreduce_bytes": int(scene.grads.numel() * 4)}

  # every timed window starts from the SAME training state (parameters, moments, statistics restored outside the timed
  # region): the scene trains while it is timed, so without this later windows measure a different scene
  def snapshot():
    snap = [t.clone() for t in (scene.params, scene.exp_avg, scene.exp_avg_sq_flat, scene.running_vis, scene.opt_step)]
    snap += [v.clone() for v in scene.stats.values()]
    if scene.color is not None:
      snap.append({k: (v.clone() if torch.is_tensor(v) else v) for k, v in scene.color.state_dict().items()})
    return snap

  def restore(snap):
    live = [scene.params, scene.exp_avg, scene.exp_avg_sq_flat, scene.running_vis, scene.opt_step] + list(scene.stats.values())
    for dst, src in zip(live, snap):
      dst.copy_(src)
    if scene.color is not None:
      scene.color.load_state_dict(snap[-1])
    torch.cuda.synchronize()

  state0 = snapshot()

  # ---- device-resident throughput (value): R windows of K steps, no stage events inside ---------------------------
  scene.timer = None
  lib = L.load()
  window_ms, launches_per_step = [], None
  with ClockSampler(local) as clocks:
    for r in range(args.windows):
      restore(state0)
      barrier()
      e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
      l0 = lib.gs_launch_count()
      e0.record()
      h0 = time.perf_counter()
      for i in range(args.steps):
        step(warm + i, dev_targets)
      host_enqueue_ms = 1e3 * (time.perf_counter() - h0) / args.steps  # CPU time to enqueue one step (no sync inside)
      e1.record()
      barrier()
      launches_per_step = (lib.gs_launch_count() - l0) / args.steps
      t = torch.tensor([e0.elapsed_time(e1)], device="cuda", dtype=torch.float64)
      if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
      window_ms.append(float(t.item()))
  ms = median(window_ms)
  overflow = scene.overflowed()

  # ---- per-stage CUDA events (a separate loop: one event per stage perturbs the enqueue) --------------------------
  timer = StageTimer()
  restore(state0)
  scene.timer = timer
  barrier()
  for i in range(args.steps):
    step(warm + i, dev_targets)
  stages = {k: float(np.mean(v)) for k, v in timer.collect().items()}
  scene.timer = None

  # ---- end to end through the public API with HOST buffers --------------------------------------------------------
  for i in range(3):  # untimed: first use of the upload stream and its two device buffers
    step(i, host_targets, read_loss=True)
  e2e_windows = []
  loss = None
  for r in range(args.windows):
    restore(state0)
    barrier()
    t0 = time.perf_counter()
    for i in range(args.steps):
      got = step(warm + i, host_targets, read_loss="pipelined")  # every step's loss is read back, one step late
      loss = got if got is not None else loss
    loss = trainer.flush_loss() or loss  # ... and the last one before the window closes
    torch.cuda.synchronize()
    if world > 1:
      dist.barrier()
    t = torch.tensor([time.perf_counter() - t0], device="cuda", dtype=torch.float64)
    if world > 1:
      dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_windows.append(float(t.item()))
  e2e_s = median(e2e_windows)
  identical = replicas_identical(scene, world, trainer)
  del state0

  # ---- what the last image did (outside every timed region) -----------------------------------------------------------
  w = scene._work
  counters = w["tile"].counters.cpu().tolist()
  O = int(counters[0])
  V = int((w["g2d"][:, 5] > 0).sum().item())
  v_opt = int(scene._vis_idx.count.item())  # rows the last optimiser step touched (device-side visible.nonzero())
  blend = None
  if w["aux"] is not None:  # instrumented forward of the same image: pairs actually evaluated / blended
    from gsb200 import ops
    cnt = torch.zeros((8,), device="cuda", dtype=torch.int64)
    tb = w["tile"]
    ops.raster_fwd(scene.raster_cfg, w["g2d"], w["rgb"], w["depth"], tb.overlap_to_point, tb.tile_ranges, W, H,
                   out=(w["image"], w["final_T"], w["n_contrib"]), aux=w["aux"], counters=cnt)
    c = cnt.cpu().tolist()
    blend = dict(records_staged=c[3], warp_splat_iterations=c[0], iterations_with_a_blend=c[1], blend_pairs=c[2])
  P, T = W * H, ((W + 15) // 16) * ((H + 15) // 16)
  D = 11 + (16 if args.scene == "mlp" else 3 * (deg + 1) ** 2)
  # under DP the optimiser visits the union of the batch's rows; the per-image terms use this rank's own image
  rows_image = int(trainer.exchange.counts()[0][0]) if (trainer.exchange is not None and world > 1) else v_opt
  bytes_ = algorithmic_bytes(n, D, V, O, P, T, rows_image)
  bytes_["optim"] = 7 * 4 * D * v_opt
  hbm, which = peaks()
  # dominant KERNEL: the tile map is a stage of nine kernels (its share is in stages_ms; the 8(d) byte model charges it a
  # 6-pass 64-bit-key sort it does not do, so a fraction for it would be meaningless) — the candidates are the single kernels
  dom = max(("raster_fwd", "raster_bwd", "project_fwd", "project_bwd"), key=lambda k: stages.get(k, 0.0))
  dom_ms = stages[dom]
  achieved = bytes_[dom] / (dom_ms * 1e-3) / 1e9
  iter_bytes = sum(bytes_.values())
  ms_step = ms / args.steps
  traffic = measured_traffic(args.workload) if (not args.points and args.scene == "sh" and world == 1) else None
  dram_step = traffic["dram_bytes_per_step"] if traffic else None

  if rank == 0:
    line = {
      # the reference's step counter advances once per IMAGE (trainer/trainer.py:532, step_rate :639-646), so the
      # whole-job value is images/s; optimiser steps/s is reported beside it
      "metric": "train_iters_per_sec", "value": world * args.steps / (ms * 1e-3), "unit": "iters/s",
      "optimizer_steps_per_sec": args.steps / (ms * 1e-3),
      "raster_fwd_bwd_mpix_per_sec": world * P / ((stages["raster_fwd"] + stages["raster_bwd"]) * 1e-3) / 1e6,
      "n_gpus": world, "steps": args.steps, "warmup": warm, "ms_per_step": ms_step,
      "windows": {"count": args.windows, "each_from_the_same_training_state": True, "ms_per_step": [x / args.steps for x in window_ms],
                  "min": min(window_ms) / args.steps, "median": ms_step, "max": max(window_ms) / args.steps},
      "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
      "config": {"workload": workload_string(args, wl, world), "images_per_step": world,
                 "l2": f"inputs larger than L2 (params+grads+moments {4 * 4 * D * n / 1e9:.1f} GB, overlap list {4 * O / 1e6:.0f} MB)",
                 "parallelism": f"dp{world}", "raster": args.raster, "V_in_view": V, "O_overlaps": O, "rows_blended": rows_image,
                 "V_optimised": v_opt, "blend": blend, "overflow": overflow, "build_id": build_id()},
      "stages_ms": stages, "host_enqueue_ms_per_step": host_enqueue_ms,
      "roofline": {"bound": "hbm", "kernel": dom, "achieved": achieved, "peak": hbm, "unit": "GB/s", "frac": achieved / hbm,
                   "traffic": traffic["dram_bytes_per_image"].get(dom) if traffic else None,
                   "traffic_source": f"profiles/r2_traffic_{args.workload}.json (ncu --set full of this command, build {build_id()})" if traffic else None,
                   "peak_source": which, "algorithmic_bytes": bytes_[dom],
                   "tile_map_stage": {"ms": stages.get("tile_map"), "dram_bytes": traffic["dram_bytes_per_image"].get("tile_map") if traffic else None,
                                      "note": "nine kernels; measured DRAM traffic only (the 8(d) model's 6-pass 64-bit sort is not what runs)"},
                   "iteration": {"algorithmic_bytes": iter_bytes, "achieved": iter_bytes / (ms_step * 1e-3) / 1e9,
                                 "frac": iter_bytes / (ms_step * 1e-3) / 1e9 / hbm, "dram_bytes_per_step": dram_step,
                                 "iteration_measured_frac": dram_step / (ms_step * 1e-3) / 1e9 / hbm if dram_step else None}},
      "e2e": {"value": world * args.steps / e2e_s, "unit": "iters/s", "h2d_bytes_per_step": int(host_targets[0].numel()),
              "d2h_bytes_per_step": 8, "windows_s": e2e_windows},
      "gpu_launches": int(round(launches_per_step * args.steps)), "gpu_launches_per_step": launches_per_step,
      "clocks": clocks.summary(),
      "last_loss_mse": loss[1] / (P * 3),
    }
    if identical is not None:
      line["replicas_identical"] = identical
    if exchange_info:
      if trainer.exchange is not None:
        exchange_info["overflow"] = any(o for _, o in trainer.exchange.counts())
      line["exchange"] = exchange_info
    if world == 1 and not args.no_cpu_baseline and args.scene == "sh":
      sc, go, gt = make_cpu_scene(wl)
      ocam = oracle_cameras(go, S["views"][:1], W, H)[0]
      tu8 = host_targets[0].numpy()
      sc.training_step([(ocam, tu8)])  # warm-up (page faults, OpenMP pool)
      t0 = time.perf_counter()
      reps = 2
      for i in range(reps):
        sc.training_step([(ocam, tu8)])
      dt = (time.perf_counter() - t0) / reps
      line["cpu_baseline"] = {"value": 1.0 / dt, "unit": "iters/s", "cores": host_threads(), "kind": "port",
                              "sample": f"{reps} full iterations of the same workload and target (after 1 warm-up) in the "
                                        f"OpenMP oracle on {sc.orc.get_threads()} threads"}
    emit(line)
  if world > 1:
    dist.destroy_process_group()


def run_c5(args):
  """BASELINE.json configs[4]: 500k -> 5M growth with split / clone / prune compaction in the training loop, following
  the reference's TargetController schedule (controller/target_controller.py:73-126, config/controller/target.yaml) over
  `--c5-steps` optimiser steps.  Under N > 1 the statistics are synchronised and every rank densifies redundantly."""
  import numpy as np
  import torch
  import torch.distributed as dist
  from gsb200.controller import TargetConfig, TargetController
  from gsb200.scene import SceneConfig, SplatScene
  from gsb200.trainer import Trainer

  args.points = args.points or 500_000
  S = setup_b200(args, "c2")
  world, rank, local, wl, W, H = S["world"], S["rank"], S["local"], S["wl"], S["W"], S["H"]
  lcams, dev_targets = S["lcams"], S["dev_targets"]
  target_points = 10 * wl["n"]
  total_images = args.c5_steps * world  # the reference's schedule counts images (trainer/trainer.py:532)
  scene = SplatScene(wl["gaussians"], SceneConfig(sh_degree=wl["sh_degree"], raster=args.raster, capacity=int(1.02 * target_points)))
  trainer = Trainer(scene, world, rank, exchange=args.exchange, check_every=25, shard=args.shard, exchange_cap=args.exchange_cap)
  controller = TargetController(TargetConfig(), scene, target_points, total_images, trainer=trainer, seed=0)
  controller.record_events = True
  trainer.controller = controller

  def batch_for(i):
    b = [None] * world
    b[rank] = (lcams[i % len(lcams)], dev_targets[i % len(dev_targets)])
    return b

  for i in range(max(args.warmup, 3)):  # untimed: capacities, exchange, first-use allocations (no densify before step 50)
    trainer.training_step(batch_for(i))
  torch.cuda.synchronize()
  if world > 1:
    dist.barrier()
  torch.cuda.reset_peak_memory_stats()
  e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
  with ClockSampler(local) as clocks:
    e0.record()
    t0 = time.perf_counter()
    for i in range(args.c5_steps):
      trainer.training_step(batch_for(i))
      if world > 1 and os.environ.get("GS_C5_DIAG") and replica_diagnosis(scene, trainer, rank, i):
        break
    e1.record()
    torch.cuda.synchronize()
    if world > 1:
      dist.barrier()
    wall = time.perf_counter() - t0
  t = torch.tensor([e0.elapsed_time(e1)], device="cuda", dtype=torch.float64)
  if world > 1:
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
  ms = float(t.item())
  dens = [a.elapsed_time(b) for a, b in controller.events]
  identical = replicas_identical(scene, world, trainer)
  if rank == 0:
    line = {"metric": "train_iters_per_sec", "value": world * args.c5_steps / (ms * 1e-3), "unit": "iters/s",
            "n_gpus": world, "steps": args.c5_steps, "warmup": max(args.warmup, 3), "ms_per_step": ms / args.c5_steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"c5: {wl['n']} -> {target_points} splats SH{wl['sh_degree']} {W}x{H}, TargetController schedule over "
                                   f"{total_images} images (densify every 50/100/200 images x1), camera batch = {world}; {LOSS}",
                       "parallelism": f"dp{world}", "final_points": scene.num_points, "capacity": scene.capacity},
            "densify": {"count": len(dens), "ms_mean": float(np.mean(dens)) if dens else None, "ms_max": max(dens) if dens else None,
                        "ms_total": float(np.sum(dens)) if dens else 0.0, "points": [h[2] for h in controller.history]},
            "wall_s": wall, "peak_memory_gb": torch.cuda.max_memory_allocated() / 1e9,
            "overflow_events": trainer.overflow_events, "clocks": clocks.summary()}
    if identical is not None:
      line["replicas_identical"] = identical
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
  elif a.workload == "c5":
    run_c5(a)
  else:
    run_b200(a)
