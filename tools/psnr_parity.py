#!/usr/bin/env python
"""psnr_parity.py — north_star's training-parity gate: train the SAME scene for K steps with the full default loss
(MSE 10 + 3-level SSIM 1.0 + reg_loss, LaProp with the reference's yaml learning rates) on the GPU (libgsb200 through
SplatScene / Trainer) and in the CPU oracle (oracle/gs_train.py) from the same seed, and compare the PSNR
(reference trainer/evaluation.py:29-31) of both against the target.  Gate: |ΔPSNR| <= 0.05 dB at the end.

  python tools/psnr_parity.py [--steps 300] [--points 100000] [--sh 0] [--width 640 --height 480] [--views 1]

Default = BASELINE.json configs[0] (C1): 100k splats, SH degree 0, one 640x480 camera.  Prints one JSON line.
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


def run(steps=300, points=100_000, sh=0, width=640, height=480, views=1, every=50, seed=0, log=None):
  import torch
  from gsb200 import lib as L
  from gsb200 import synthetic as syn
  from gsb200.scene import SceneConfig, SplatScene
  from gsb200.trainer import Trainer
  from oracle import gs_oracle as go
  from oracle import gs_train as gt

  W, H = width, height
  g = syn.spec_scene(points, sh, seed=seed)
  cams = syn.ring_cameras(max(views, 8), W, H, seed=seed)[:views]
  lcams = [L.make_camera(c["T"], c["fx"], c["fy"], c["cx"], c["cy"], W, H, c["near"], c["far"]) for c in cams]
  ocams = [go.make_camera(c["T"], c["fx"], c["fy"], c["cx"], c["cy"], W, H, c["near"], c["far"]) for c in cams]
  cfg = SceneConfig(sh_degree=sh)
  lrs = dict(position=cfg.lr_position, log_scaling=cfg.lr_log_scaling, rotation=cfg.lr_rotation,
             alpha_logit=cfg.lr_alpha_logit, sh=cfg.lr_sh)
  # targets: the render of an independent scene drawn with seed + 1, stored as uint8, made by the oracle
  tgt = gt.OracleScene(syn.spec_scene(points, sh, seed=seed + 1), sh, lrs)
  targets = [np.clip(tgt.render(oc) * 255.0 + 0.5, 0, 255).astype(np.uint8) for oc in ocams]
  tf = [t.astype(np.float32) / np.float32(255.0) for t in targets]
  del tgt

  scene = SplatScene(g, cfg)
  trainer = Trainer(scene)
  oscene = gt.OracleScene(g, sh, lrs, w_l1=cfg.w_l1, w_mse=cfg.w_mse, w_ssim=cfg.w_ssim, ssim_levels=cfg.ssim_levels,
                          reg_weights=cfg.reg_weights, split_alpha=cfg.split_alpha, prune_alpha=cfg.prune_alpha,
                          blur_cov=cfg.blur_cov)
  dev_targets = [torch.from_numpy(t).cuda() for t in targets]

  def psnr_gpu():
    return float(np.mean([gt.psnr(scene.render(lc)["image"].cpu().numpy(), t) for lc, t in zip(lcams, tf)]))

  def psnr_cpu():
    return float(np.mean([gt.psnr(np.clip(oscene.render(oc), 0, 1), t) for oc, t in zip(ocams, tf)]))

  curve = [dict(step=0, gpu=psnr_gpu(), cpu=psnr_cpu())]
  t_gpu = t_cpu = 0.0
  for i in range(steps):
    v = i % views
    t0 = time.perf_counter()
    trainer.training_step([(lcams[v], dev_targets[v])])
    torch.cuda.synchronize()
    t1 = time.perf_counter()
    oscene.training_step([(ocams[v], targets[v])])
    t2 = time.perf_counter()
    t_gpu += t1 - t0; t_cpu += t2 - t1
    if (i + 1) % every == 0 or i + 1 == steps:
      curve.append(dict(step=i + 1, gpu=psnr_gpu(), cpu=psnr_cpu()))
      if log:
        log(curve[-1])
  dp = {k: float(np.abs(scene.p[k].cpu().numpy().reshape(oscene.g[k].shape) - oscene.g[k]).max()) for k in gt.ATTRS}
  return dict(workload=f"{points} splats SH{sh} {W}x{H}, {views} camera(s), {steps} steps, MSE 10 + 3-level SSIM 1.0 + reg, LaProp",
              learning_rates=lrs, curve=curve, final_gpu_psnr=curve[-1]["gpu"], final_cpu_psnr=curve[-1]["cpu"],
              delta_db=abs(curve[-1]["gpu"] - curve[-1]["cpu"]), max_abs_param_diff=dp,
              seconds=dict(gpu=t_gpu, cpu_oracle=t_cpu))


if __name__ == "__main__":
  ap = argparse.ArgumentParser()
  ap.add_argument("--steps", type=int, default=300)
  ap.add_argument("--points", type=int, default=100_000)
  ap.add_argument("--sh", type=int, default=0)
  ap.add_argument("--width", type=int, default=640)
  ap.add_argument("--height", type=int, default=480)
  ap.add_argument("--views", type=int, default=1)
  a = ap.parse_args()
  out = run(a.steps, a.points, a.sh, a.width, a.height, a.views, log=lambda c: print(json.dumps(c), file=sys.stderr, flush=True))
  print(json.dumps(out))
  sys.exit(0 if out["delta_db"] <= 0.05 else 1)
