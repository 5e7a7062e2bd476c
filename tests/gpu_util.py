"""Helpers for the -m gpu parity tests: numpy <-> CUDA tensors, oracle <-> libgsb200 struct conversion."""
import numpy as np
import torch

from gsb200 import lib as L
from oracle import gs_oracle as go


def cu(a, dtype=None):
  t = torch.from_numpy(np.ascontiguousarray(a))
  if dtype is not None:
    t = t.to(dtype)
  return t.cuda().contiguous()


def npy(t):
  return t.detach().cpu().numpy()


def to_lib_camera(cam):
  """oracle Camera -> libgsb200 Camera (identical layout)."""
  return L.make_camera(list(cam.T), cam.fx, cam.fy, cam.cx, cam.cy, cam.width, cam.height, cam.near_plane,
                       cam.far_plane)


def to_lib_config(cfg):
  out = L.RasterConfig()
  for name, _ in L.RasterConfig._fields_:
    setattr(out, name, getattr(cfg, name))
  return out


def to_lib_optim(ocfg):
  out = L.OptimConfig()
  for name, _ in L.OptimConfig._fields_:
    setattr(out, name, getattr(ocfg, name))
  return out


def gaussians_to_cuda(g):
  return {k: cu(v) for k, v in g.items()}


def oracle():
  return go.Oracle("f32")
