"""Seeded synthetic inputs shared by the oracle tests and the GPU parity tests."""
import numpy as np

from oracle import gs_oracle as go


def look_at(eye, target=(0, 0, 0), up=(0, 0, 1)):
  """4x4 camera-from-world with +z forward, +x right, +y down (COLMAP / reference convention)."""
  eye = np.asarray(eye, np.float64); target = np.asarray(target, np.float64); up = np.asarray(up, np.float64)
  f = target - eye; f /= np.linalg.norm(f)
  r = np.cross(f, up); r /= np.linalg.norm(r)
  d = np.cross(f, r)
  R = np.stack([r, d, f])
  T = np.eye(4)
  T[:3, :3] = R
  T[:3, 3] = -R @ eye
  return T


def random_camera(W=96, H=64, fov_deg=50.0, eye=(3.0, 0.4, 0.8)):
  fx = 0.5 * W / np.tan(np.radians(fov_deg) / 2)
  return go.make_camera(look_at(eye), fx, fx, W / 2, H / 2, W, H, 0.1, 100.0)


def random_gaussians(n, seed=0, sh_degree=3, scale=0.08, spread=1.0):
  rng = np.random.default_rng(seed)
  K = (sh_degree + 1) ** 2
  pos = rng.normal(0, spread, (n, 3)) * np.array([1.0, 1.0, 0.5])
  ls = np.log(scale) + rng.normal(0, 0.5, (n, 3))
  q = rng.normal(0, 1, (n, 4))
  al = rng.uniform(-2.0, 3.0, n)
  sh = rng.normal(0, 0.3, (n, 3, K))
  sh[:, :, 0] = rng.uniform(-1.5, 1.5, (n, 3))
  f32 = lambda a: np.ascontiguousarray(a, np.float32)
  return dict(position=f32(pos), log_scaling=f32(ls), rotation=f32(q), alpha_logit=f32(al), sh=f32(sh))
