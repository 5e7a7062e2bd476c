"""Seeded synthetic scenes / cameras of the shapes BASELINE.json names.

Garden-shaped: 70 % of the splats in a central blob N(0, diag(1.5,1.5,0.6)²), 30 % in a far shell of radius
U(4,20); log-scale ~ N(log(0.7·d3), 0.5²) where d3 is the expected distance to the 3rd nearest neighbour derived
from the analytic local density of the generating distribution (a KNN over 3M points is init-time work the
reference does with pykeops, gaussians/loading.py:45-61 — out of scope); unit quaternions xyzw; opacity
U(0.05,0.95); SH DC = rgb_to_sh(U(0,1)) (reference util/misc.py:41-49), higher bands N(0, 0.05²).
Cameras: ring of radius 4, height U(0.5,2), looking at the origin, horizontal FOV 50°, near 0.1, far 100
(reference config/trainer/default.yaml:83-84).  Everything is generated with numpy's PCG64 from the seed so the
CPU oracle and every rank see identical inputs.
"""
import math

import numpy as np

SH0 = 0.282094791773878


def look_at(eye, target=(0.0, 0.0, 0.0), up=(0.0, 0.0, 1.0)):
  """4x4 camera-from-world, +z forward, +x right, +y down (COLMAP convention used by the reference)."""
  eye = np.asarray(eye, np.float64); target = np.asarray(target, np.float64); up = np.asarray(up, np.float64)
  f = target - eye; f /= np.linalg.norm(f)
  r = np.cross(f, up); r /= np.linalg.norm(r)
  d = np.cross(f, r)
  R = np.stack([r, d, f])
  T = np.eye(4)
  T[:3, :3] = R
  T[:3, 3] = -R @ eye
  return T


def ring_cameras(n_views, width, height, seed=0, radius=4.0, fov_deg=50.0):
  """list of dict(T (4,4) float64, fx, fy, cx, cy, width, height, near, far)."""
  rng = np.random.default_rng(seed + 1000)
  fx = 0.5 * width / math.tan(math.radians(fov_deg) / 2)
  cams = []
  for i in range(n_views):
    ang = 2 * math.pi * i / n_views
    h = rng.uniform(0.5, 2.0)
    eye = (radius * math.cos(ang), radius * math.sin(ang), h)
    cams.append(dict(T=look_at(eye), fx=fx, fy=fx, cx=width / 2, cy=height / 2, width=width, height=height,
                     near=0.1, far=100.0))
  return cams


def garden_scene(n, sh_degree=3, seed=0, scale_mult=0.7, shell_min=6.0, max_px=10.0, focal=1390.7, ring_radius=4.0):
  """dict of float32 numpy arrays: position (n,3), log_scaling (n,3), rotation (n,4), alpha_logit (n,), sh (n,3,K)."""
  rng = np.random.default_rng(seed)
  K = (sh_degree + 1) ** 2
  n_blob = int(0.7 * n)
  n_shell = n - n_blob
  sig = np.array([1.5, 1.5, 0.6])
  blob = rng.normal(0, 1, (n_blob, 3)) * sig
  # keep the blob clear of the camera ring (radius 4): the reference's controller prunes any splat seen larger than
  # 200 px (config/controller/target.yaml:9-26, controller/target_controller.py:77), so a trained scene has none
  rxy = np.hypot(blob[:, 0], blob[:, 1])
  far = rxy > 3.0
  blob[far, :2] *= (3.0 * np.sqrt(rng.uniform(0, 1, far.sum())) / rxy[far])[:, None]
  # local density of the blob at each point -> expected 3rd-NN distance
  z = blob / sig
  pdf = np.exp(-0.5 * (z * z).sum(1)) / ((2 * math.pi) ** 1.5 * sig.prod())
  rho_blob = n_blob * pdf
  r = rng.uniform(shell_min, 20.0, n_shell)
  dirs = rng.normal(0, 1, (n_shell, 3)); dirs /= np.linalg.norm(dirs, axis=1, keepdims=True)
  shell = dirs * r[:, None]
  rho_shell = n_shell / ((20.0 - shell_min) * 4 * math.pi * r * r)  # uniform in radius
  pos = np.concatenate([blob, shell])
  rho = np.concatenate([rho_blob, rho_shell])
  d3 = (3.0 * 3.0 / (4 * math.pi * np.maximum(rho, 1e-9))) ** (1.0 / 3.0)  # radius of the ball holding 3 neighbours
  d3 = np.minimum(d3, 0.15)
  ls = np.log(scale_mult * d3)[:, None] + rng.normal(0, 0.5, (n, 3))
  # no splat projects larger than max_px (1 sigma) from the nearest point of the camera ring: reconstruction density
  # follows image resolution, and the reference's controller prunes oversized splats (max_scale_px, see above)
  d_ring = np.sqrt((np.hypot(pos[:, 0], pos[:, 1]) - ring_radius) ** 2 + (pos[:, 2] - 1.25) ** 2)
  ls = np.minimum(ls, np.log(max_px * np.maximum(d_ring, 0.05) / focal)[:, None])
  ls = np.clip(ls, -8, 8)
  q = rng.normal(0, 1, (n, 4)); q /= np.linalg.norm(q, axis=1, keepdims=True)
  op = rng.uniform(0.05, 0.95, n)
  al = np.log(op / (1 - op))
  sh = rng.normal(0, 0.05, (n, 3, K))
  sh[:, :, 0] = (rng.uniform(0, 1, (n, 3)) - 0.5) / SH0
  perm = rng.permutation(n)  # interleave blob and shell so any row slice is a fair sample
  f32 = lambda a: np.ascontiguousarray(a[perm], np.float32)
  return dict(position=f32(pos), log_scaling=f32(ls), rotation=f32(q), alpha_logit=f32(al), sh=f32(sh))


def knn_mean_distance_cpu(points, k=3):
  """mean distance to the k nearest neighbours by a k-d tree on the host (scipy) — the same quantity as the library's
  gs_knn_scale / the reference's estimate_scale (gaussians/loading.py:45-61); used where no GPU may be touched (the CPU
  arm of bench.py, CPU tests)."""
  from scipy.spatial import cKDTree
  pts = np.asarray(points, np.float64)
  d, _ = cKDTree(pts).query(pts, k=k + 1, workers=-1)
  return d[:, 1:].mean(axis=1).astype(np.float32)


def spec_scene(n, sh_degree=3, seed=0, knn=None, opacity=(0.05, 0.95)):
  """The scene exactly as SURVEY.md §8(d) writes it: 70 % of the splats in a central blob N(0, diag(1.5,1.5,0.6)²), 30 %
  in a far shell of radius U(4,20); log_scaling ~ N(log(0.7·knn3 distance), 0.5²) clamped to [-8, 8]; quaternion =
  normalize(N(0,I4)) xyzw; alpha_logit = logit(U(0.05,0.95)); SH DC = rgb_to_sh(U(0,1)), higher bands N(0, 0.05²).
  knn: callable (n,3) float32 -> (n,) mean distance to the 3 nearest neighbours (default: the host k-d tree; bench.py's
  GPU arm passes the library's own gs_knn_scale)."""
  rng = np.random.default_rng(seed)
  K = (sh_degree + 1) ** 2
  n_blob = int(0.7 * n)
  n_shell = n - n_blob
  blob = rng.normal(0, 1, (n_blob, 3)) * np.array([1.5, 1.5, 0.6])
  r = rng.uniform(4.0, 20.0, n_shell)
  dirs = rng.normal(0, 1, (n_shell, 3)); dirs /= np.linalg.norm(dirs, axis=1, keepdims=True)
  pos = np.concatenate([blob, dirs * r[:, None]]).astype(np.float32)
  perm = rng.permutation(n)
  pos = np.ascontiguousarray(pos[perm])
  d3 = np.asarray((knn or knn_mean_distance_cpu)(pos), np.float64)
  ls = np.clip(np.log(0.7 * np.maximum(d3, 1e-12))[:, None] + rng.normal(0, 0.5, (n, 3)), -8, 8)
  q = rng.normal(0, 1, (n, 4)); q /= np.linalg.norm(q, axis=1, keepdims=True)
  op = rng.uniform(opacity[0], opacity[1], n)
  al = np.log(op / (1 - op))
  sh = rng.normal(0, 0.05, (n, 3, K))
  sh[:, :, 0] = (rng.uniform(0, 1, (n, 3)) - 0.5) / SH0
  f32 = lambda a: np.ascontiguousarray(a, np.float32)
  return dict(position=pos, log_scaling=f32(ls), rotation=f32(q), alpha_logit=f32(al), sh=f32(sh))


def trained_scene(n, sh_degree=3, seed=0, focal=1390.7):
  """A scene shaped like one the reference's controller has trained for a while: the sparse generator's geometry with
  small translucent splats (sigma <= 5 px from the camera ring, opacity U(0.02, 0.15)), so a pixel blends hundreds of
  layers before it saturates and most in-view rows receive gradient (measured at C2: 70 % of the in-view rows blend,
  mean n_contrib ~1000) — the opposite corner from the sparse scene's ~3 %."""
  g = garden_scene(n, sh_degree, seed=seed, max_px=5.0, focal=focal)
  rng = np.random.default_rng(seed + 77)
  op = rng.uniform(0.02, 0.15, n)
  g["alpha_logit"] = np.ascontiguousarray(np.log(op / (1 - op)), np.float32)
  return g


def perturbed(g, seed=1, pos_sigma=0.01, sh_sigma=0.2, alpha_sigma=0.5):
  """a second scene for the training TARGET images: same splats, perturbed colour/opacity/position."""
  rng = np.random.default_rng(seed)
  out = {k: v.copy() for k, v in g.items()}
  out["position"] += rng.normal(0, pos_sigma, g["position"].shape).astype(np.float32)
  out["sh"] += (rng.normal(0, sh_sigma, g["sh"].shape) * (np.arange(g["sh"].shape[2]) == 0)).astype(np.float32)
  out["alpha_logit"] += rng.normal(0, alpha_sigma, g["alpha_logit"].shape).astype(np.float32)
  return out


# ---- the BASELINE.json workloads ---------------------------------------------------------------------------------------
WORKLOADS = {
  # name: (points, sh_degree, width, height, scene generator, target)
  "c1": (100_000, 0, 640, 480, "spec", "independent"),            # configs[0]: the CPU-runnable case
  "c2": (3_000_000, 3, 1297, 840, "spec", "independent"),         # configs[1], generated exactly as SURVEY 8(d) writes it
  "c2-sparse": (3_000_000, 3, 1297, 840, "sparse", "perturbed"),  # round 1's generator: opaque small splats, ~3 % of rows blend
  "c2-trained": (3_000_000, 3, 1297, 840, "trained", "perturbed"),  # translucent splats: most in-view rows blend
  "c4": (6_000_000, 3, 3840, 2160, "sparse", "perturbed"),        # configs[3]: high-res stress
}


def workload(name, points=0, knn=None, seed=0):
  """-> dict(n, sh_degree, width, height, gaussians, target_gaussians, cameras (185 ring views), description)."""
  n, deg, W, H, kind, tgt = WORKLOADS[name]
  if points:
    n = points
  focal = 0.5 * W / math.tan(math.radians(25.0))
  make = dict(spec=lambda s: spec_scene(n, deg, seed=s, knn=knn),
              sparse=lambda s: garden_scene(n, deg, seed=s, focal=focal),
              trained=lambda s: trained_scene(n, deg, seed=s, focal=focal))[kind]
  g = make(seed)
  target = make(seed + 1) if tgt == "independent" else perturbed(g, seed=seed + 1)
  what = dict(spec="scene as SURVEY 8(d): blob + shell U(4,20), knn3 scales, opacity U(.05,.95)",
              sparse="opaque small splats (10 px cap): few rows blend",
              trained="small translucent splats (opacity U(.02,.15), 5 px cap): ~70 % of the in-view rows blend")[kind]
  return dict(n=n, sh_degree=deg, width=W, height=H, gaussians=g, target_gaussians=target,
              cameras=ring_cameras(185, W, H, seed=seed),
              description=f"{name}: {n} splats SH{deg} {W}x{H}, {what}; target = "
                          f"{'independent seed+1 scene' if tgt == 'independent' else 'perturbed copy'}")
