#!/usr/bin/env python
"""raster_lab.py — times the rasteriser kernels alone on a BASELINE-shaped image (CUDA events, L2 flushed between
repetitions) and counts the work they do: gather-staged kernels (raster.cu) against the record-staged asynchronous
kernels (raster_rec.cu), with and without the heaviest-first persistent tile queue.  Also checks the two paths agree.

  python tools/raster_lab.py [--scene sparse|spec|trained] [--points N] [--reps R]
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "splat-trainer_b200"))
sys.path.insert(0, ROOT)

import numpy as np
import torch

from gsb200 import lib as L
from gsb200 import ops
from gsb200 import synthetic as syn


def timed(fn, reps, flush):
  ts = []
  for _ in range(reps):
    flush.zero_()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); fn(); e1.record()
    torch.cuda.synchronize()
    ts.append(e0.elapsed_time(e1))
  ts.sort()
  return ts[len(ts) // 2], ts[0], ts[-1]


def main():
  ap = argparse.ArgumentParser()
  ap.add_argument("--scene", default="sparse")
  ap.add_argument("--points", type=int, default=3_000_000)
  ap.add_argument("--width", type=int, default=1297)
  ap.add_argument("--height", type=int, default=840)
  ap.add_argument("--reps", type=int, default=9)
  ap.add_argument("--camera", type=int, default=0)
  a = ap.parse_args()
  dev = torch.device("cuda")
  L.load()
  W, H, n = a.width, a.height, a.points
  knn = lambda p: ops.estimate_scale(torch.as_tensor(p, device=dev), 3).cpu().numpy()
  g = dict(sparse=lambda: syn.garden_scene(n, 3, seed=0), spec=lambda: syn.spec_scene(n, 3, seed=0, knn=knn),
           trained=lambda: syn.trained_scene(n, 3, seed=0))[a.scene]()
  c = syn.ring_cameras(185, W, H, seed=0)[a.camera]
  cam = L.make_camera(c["T"], c["fx"], c["fy"], c["cx"], c["cy"], W, H)
  cfg = L.default_raster_config(blur_cov=0.3, compute_visibility=1, compute_point_heuristic=1)
  t = {k: torch.as_tensor(v, device=dev) for k, v in g.items()}
  g2d, depth, rgb = ops.project_fwd(cam, cfg, t["position"], t["log_scaling"], t["rotation"], t["alpha_logit"], t["sh"], 3)
  o2p, ranges, counters, _ = ops.tile_map(cfg, g2d, depth, W, H)
  torch.cuda.synchronize()
  O = int(counters[0].item())
  V = int((g2d[:, 5] > 0).sum().item())
  flush = torch.empty((256 << 20,), device=dev, dtype=torch.uint8)
  target = torch.rand((H, W, 3), device=dev)
  class Out(dict):
    def __setitem__(self, k, v):  # partial results go to stderr at once: a hang later on must not lose them
      dict.__setitem__(self, k, v)
      print(json.dumps({k: v}), file=sys.stderr, flush=True)

  out = Out()
  for k, v in dict(scene=a.scene, n=n, W=W, H=H, V=V, O=O).items():
    out[k] = v

  # ---- gather-staged kernels (raster.cu) ------------------------------------------------------------------------------
  image, fT, nc, _, _ = ops.raster_fwd(cfg, g2d, rgb, depth, o2p, ranges, W, H)
  d_image = (image.clamp(0, 1) - target) * (2.0 / (H * W * 3))
  acc0 = ops.raster_bwd(cfg, g2d, rgb, o2p, ranges, W, H, fT, nc, d_image)
  torch.cuda.synchronize()
  lens = (ranges[:, 1] - ranges[:, 0]).float()
  ncf = nc.float()
  out["list_len"] = dict(mean=float(lens.mean()), max=float(lens.max()))
  out["n_contrib"] = dict(mean=float(ncf.mean()), max=float(ncf.max()))
  out["rows_blended"] = int((acc0[:, 11] > 0).sum().item())
  out["fwd_gather_ms"] = timed(lambda: ops.raster_fwd(cfg, g2d, rgb, depth, o2p, ranges, W, H, out=(image, fT, nc)), a.reps, flush)
  acc = torch.zeros_like(acc0)

  def bwd_gather():
    acc.zero_()
    ops.raster_bwd(cfg, g2d, rgb, o2p, ranges, W, H, fT, nc, d_image, acc=acc)

  out["bwd_gather_ms"] = timed(bwd_gather, a.reps, flush)

  # ---- record-staged kernels (raster_rec.cu) -------------------------------------------------------------------------
  for persistent in (False, True):
    tag = "queue" if persistent else "grid"
    aux = ops.RasterAux(n, W, H, dev, persistent=persistent)
    out["records_ms"] = timed(lambda: ops.raster_records(cfg, g2d, rgb, aux.records), a.reps, flush)
    image2, fT2, nc2 = torch.empty_like(image), torch.empty_like(fT), torch.empty_like(nc)
    ops.raster_fwd(cfg, g2d, rgb, depth, o2p, ranges, W, H, out=(image2, fT2, nc2), aux=aux)
    torch.cuda.synchronize()
    out[f"fwd_rec_{tag}_equal"] = bool(torch.equal(image, image2) and torch.equal(fT, fT2) and torch.equal(nc, nc2))
    tile_max = torch.zeros_like(aux.tile_work)
    TX = (W + 15) // 16
    ys, xs = torch.meshgrid(torch.arange(H, device=dev), torch.arange(W, device=dev), indexing="ij")
    tile_max.scatter_reduce_(0, ((ys // 16) * TX + xs // 16).reshape(-1), nc.reshape(-1), "amax")
    out[f"tile_work_{tag}_ok"] = bool(torch.equal(tile_max, aux.tile_work))
    out[f"fwd_rec_{tag}_ms"] = timed(lambda: ops.raster_fwd(cfg, g2d, rgb, depth, o2p, ranges, W, H, out=(image2, fT2, nc2), aux=aux),
                                     a.reps, flush)
    acc2 = torch.zeros_like(acc0)
    ops.raster_bwd(cfg, g2d, rgb, o2p, ranges, W, H, fT, nc, d_image, acc=acc2, aux=aux)
    torch.cuda.synchronize()
    scale = acc0.abs().amax(dim=0).clamp_min(1e-30)
    out[f"bwd_rec_{tag}_max_rel_err"] = float(((acc2 - acc0).abs() / scale).max().item())

    def bwd_rec():
      acc2.zero_()
      ops.raster_bwd(cfg, g2d, rgb, o2p, ranges, W, H, fT, nc, d_image, acc=acc2, aux=aux)

    out[f"bwd_rec_{tag}_ms"] = timed(bwd_rec, a.reps, flush)
    if persistent:
      counters = torch.zeros((8,), device=dev, dtype=torch.int64)
      ops.raster_fwd(cfg, g2d, rgb, depth, o2p, ranges, W, H, out=(image2, fT2, nc2), aux=aux, counters=counters)
      torch.cuda.synchronize()
      cs = counters.tolist()
      out["fwd_counters"] = dict(warp_splat_iterations=cs[0], iterations_with_a_blend=cs[1], pixel_splat_blends=cs[2],
                                 records_staged=cs[3])
  acc_zero_ms = timed(lambda: acc.zero_(), a.reps, flush)
  out["acc_zero_ms"] = acc_zero_ms
  print(json.dumps(out))


if __name__ == "__main__":
  main()
