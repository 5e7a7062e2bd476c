#!/usr/bin/env python
"""psnr_parity.py — north_star's training-parity gate: train the SAME scene for K steps with the full default loss
(MSE 10 + 3-level SSIM 1.0 + reg_loss, LaProp with the reference's yaml learning rates) on the GPU (libgsb200 through
SplatScene / Trainer) and in the CPU oracle (oracle/gs_train.py) from the same seed, and compare the PSNR
(reference trainer/evaluation.py:29-31) of both against the target.  Gate: |ΔPSNR| <= 0.05 dB at the end.

  python tools/psnr_parity.py [--steps 300] [--points 100000] [--sh 0] [--width 640 --height 480] [--views 8]
                              [--scene sparse|spec] [--target perturbed|independent] [--lr-scale S] [--floor]

Default = the shape of BASELINE.json configs[0] (C1: 100k splats, SH degree 0, 640x480) as a teacher-student run: the
targets are renders of a perturbed copy of the scene, seen from 8 ring cameras in turn, so the training converges.
(A random scene fitted to an unrelated random image with the reference's constant learning rates is a chaotic system: the
oracle run against a copy of itself perturbed by 1e-7 drifts apart by 0.2-1.3 dB within 200 steps — `--floor` measures
exactly that — so there |dPSNR| says nothing about the implementation.)  Prints one JSON line.
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "splat-trainer_b200"))
sys.path.insert(0, ROOT)

import numpy as np


def run(steps=300, points=100_000, sh=0, width=640, height=480, views=8, every=50, seed=0, scene="sparse", target="perturbed",
        lr_scale=1.0, floor=False, log=None):
  """scene: "sparse" (gsb200.synthetic.garden_scene) or "spec" (the SURVEY 8(d) scene); target: "perturbed" (render of a
  perturbed copy: a teacher the student can reach, so training converges) or "independent" (seed + 1 scene).
  lr_scale multiplies the reference's learning rates.  floor: also train the oracle from a copy of the scene whose
  positions are perturbed by 1e-7 relative — |dPSNR| between the two oracle runs is the sensitivity of the training
  dynamics themselves, the floor under which no two implementations can be told apart."""
  import torch
  from gsb200 import lib as L
  from gsb200 import synthetic as syn
  from gsb200.scene import SceneConfig, SplatScene
  from gsb200.trainer import Trainer
  from oracle import gs_oracle as go
  from oracle import gs_train as gt

  W, H = width, height
  focal = 0.5 * W / np.tan(np.radians(25.0))
  make = (lambda s: syn.spec_scene(points, sh, seed=s)) if scene == "spec" else (lambda s: syn.garden_scene(points, sh, seed=s, focal=focal))
  g = make(seed)
  gt_scene = syn.perturbed(g, seed=seed + 1) if target == "perturbed" else make(seed + 1)
  cams = syn.ring_cameras(max(views, 8), W, H, seed=seed)[:views]
  lcams = [L.make_camera(c["T"], c["fx"], c["fy"], c["cx"], c["cy"], W, H, c["near"], c["far"]) for c in cams]
  ocams = [go.make_camera(c["T"], c["fx"], c["fy"], c["cx"], c["cy"], W, H, c["near"], c["far"]) for c in cams]
  base = SceneConfig(sh_degree=sh)
  cfg = SceneConfig(sh_degree=sh, lr_position=base.lr_position * lr_scale, lr_log_scaling=base.lr_log_scaling * lr_scale,
                    lr_rotation=base.lr_rotation * lr_scale, lr_alpha_logit=base.lr_alpha_logit * lr_scale, lr_sh=base.lr_sh * lr_scale)
  lrs = dict(position=cfg.lr_position, log_scaling=cfg.lr_log_scaling, rotation=cfg.lr_rotation,
             alpha_logit=cfg.lr_alpha_logit, sh=cfg.lr_sh)
  # targets: rendered by the oracle, stored as uint8 like the reference's dataset
  tgt = gt.OracleScene(gt_scene, sh, lrs)
  targets = [np.clip(tgt.render(oc) * 255.0 + 0.5, 0, 255).astype(np.uint8) for oc in ocams]
  tf = [t.astype(np.float32) / np.float32(255.0) for t in targets]
  del tgt

  def oracle_scene(gaussians):
    return gt.OracleScene(gaussians, sh, lrs, w_l1=cfg.w_l1, w_mse=cfg.w_mse, w_ssim=cfg.w_ssim, ssim_levels=cfg.ssim_levels,
                          reg_weights=cfg.reg_weights, split_alpha=cfg.split_alpha, prune_alpha=cfg.prune_alpha,
                          blur_cov=cfg.blur_cov)

  gpu_scene = SplatScene(g, cfg)
  trainer = Trainer(gpu_scene)
  oscene = oracle_scene(g)
  twin = None
  if floor:
    g2 = {k: v.copy() for k, v in g.items()}
    g2["position"] = (g2["position"] * (1 + 1e-7 * np.random.default_rng(5).normal(size=g2["position"].shape))).astype(np.float32)
    twin = oracle_scene(g2)
  dev_targets = [torch.from_numpy(t).cuda() for t in targets]

  def psnr_gpu():
    return float(np.mean([gt.psnr(gpu_scene.render(lc)["image"].cpu().numpy(), t) for lc, t in zip(lcams, tf)]))

  def psnr_of(s):
    return float(np.mean([gt.psnr(np.clip(s.render(oc), 0, 1), t) for oc, t in zip(ocams, tf)]))

  def point(step):
    c = dict(step=step, gpu=psnr_gpu(), cpu=psnr_of(oscene))
    if twin is not None:
      c["cpu_twin"] = psnr_of(twin)
    return c

  curve = [point(0)]
  t_gpu = t_cpu = 0.0
  for i in range(steps):
    v = i % views
    t0 = time.perf_counter()
    trainer.training_step([(lcams[v], dev_targets[v])])
    torch.cuda.synchronize()
    t1 = time.perf_counter()
    oscene.training_step([(ocams[v], targets[v])])
    t2 = time.perf_counter()
    if twin is not None:
      twin.training_step([(ocams[v], targets[v])])
    t_gpu += t1 - t0; t_cpu += t2 - t1
    if (i + 1) % every == 0 or i + 1 == steps:
      curve.append(point(i + 1))
      if log:
        log(curve[-1])
  last = curve[-1]
  out = dict(workload=f"{points} splats SH{sh} {W}x{H}, {scene} scene, {target} target, {views} camera(s), {steps} steps, "
                      f"MSE 10 + 3-level SSIM 1.0 + reg, LaProp, reference learning rates x {lr_scale:g}",
             learning_rates=lrs, curve=curve, final_gpu_psnr=last["gpu"], final_cpu_psnr=last["cpu"],
             delta_db=abs(last["gpu"] - last["cpu"]), max_delta_db=max(abs(c["gpu"] - c["cpu"]) for c in curve),
             seconds=dict(gpu=t_gpu, cpu_oracle=t_cpu))
  if twin is not None:
    out["floor_db"] = abs(last["cpu"] - last["cpu_twin"])
    out["max_floor_db"] = max(abs(c["cpu"] - c["cpu_twin"]) for c in curve)
  return out


if __name__ == "__main__":
  ap = argparse.ArgumentParser()
  ap.add_argument("--steps", type=int, default=300)
  ap.add_argument("--points", type=int, default=100_000)
  ap.add_argument("--sh", type=int, default=0)
  ap.add_argument("--width", type=int, default=640)
  ap.add_argument("--height", type=int, default=480)
  ap.add_argument("--views", type=int, default=8)
  ap.add_argument("--scene", default="sparse", choices=["sparse", "spec"])
  ap.add_argument("--target", default="perturbed", choices=["perturbed", "independent"])
  ap.add_argument("--lr-scale", type=float, default=1.0)
  ap.add_argument("--floor", action="store_true", help="also measure the dynamics' own sensitivity (oracle vs perturbed oracle)")
  a = ap.parse_args()
  out = run(a.steps, a.points, a.sh, a.width, a.height, a.views, scene=a.scene, target=a.target, lr_scale=a.lr_scale, floor=a.floor,
            log=lambda c: print(json.dumps(c), file=sys.stderr, flush=True))
  print(json.dumps(out))
  sys.exit(0 if out["delta_db"] <= max(0.05, 3 * out.get("floor_db", 0.0)) else 1)
