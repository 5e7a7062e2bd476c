"""random_3d_gaussians / random_camera (reference scripts/test_split.py:1,24-27)."""
import math

import torch

from ..data_types import Gaussians3D
from ..perspective import CameraParams


def random_camera(pos_scale=1.0, image_size=None, device="cuda", generator=None):
  g = generator
  if image_size is None:
    w = int(torch.randint(256, 1024, (1,), generator=g).item())
    image_size = (w, int(w * 0.75))
  W, H = image_size
  f = 0.5 * W / math.tan(math.radians(25.0))
  eye = torch.randn(3, generator=g) * pos_scale
  eye = eye / eye.norm().clamp_min(1e-6) * 4.0 * pos_scale
  fwd = -eye / eye.norm()
  up = torch.tensor([0.0, 0.0, 1.0])
  right = torch.linalg.cross(fwd, up); right = right / right.norm().clamp_min(1e-6)
  down = torch.linalg.cross(fwd, right)
  R = torch.stack([right, down, fwd])
  T = torch.eye(4)
  T[:3, :3] = R
  T[:3, 3] = -R @ eye
  return CameraParams(T_camera_world=T.to(device), projection=torch.tensor([f, f, W / 2, H / 2], device=device),
                      image_size=(W, H), near_plane=0.1, far_plane=100.0)


def random_3d_gaussians(n, camera_params=None, scale_factor=1.0, alpha_range=(0.1, 0.9), device="cuda", generator=None):
  g = generator
  pos = torch.randn((n, 3), generator=g) * 1.0
  ls = torch.log(torch.rand((n, 3), generator=g) * 0.2 * scale_factor + 0.02)
  q = torch.randn((n, 4), generator=g); q = q / q.norm(dim=1, keepdim=True)
  a = torch.rand((n, 1), generator=g) * (alpha_range[1] - alpha_range[0]) + alpha_range[0]
  return Gaussians3D(position=pos.to(device), log_scaling=ls.to(device), rotation=q.to(device),
                     alpha_logit=torch.logit(a).to(device), feature=torch.rand((n, 3), generator=g).to(device))
