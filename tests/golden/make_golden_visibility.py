"""Golden vectors for the per-view cluster features, the frustum visibility counts and the evaluation metrics, made by
RUNNING the reference's own code (build container only: `python tests/golden/make_golden_visibility.py`):
  * PointClusters.view_features           visibility/cluster.py:40-55
  * point_visibility(cameras, points)     visibility/query_points.py:62-92 over the reference's own Cameras / Projections
  * compute_psnr / l1                     util/colors.py, trainer/evaluation.py:29-37
The reference modules are imported unchanged from /root/reference over gsb200's import shims (gsb200.compat: the
operator shim + the third-party stand-ins; `tensorclass` here is the stand-in, the arithmetic is the reference's)."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "splat-trainer_b200"))
OUT = os.path.dirname(os.path.abspath(__file__))

from gsb200 import compat, synthetic as syn  # noqa: E402

compat.enable("/root/reference")
import torch._dynamo  # noqa: E402

torch._dynamo.config.disable = True
import splat_trainer.scene.scene  # noqa: E402,F401  (the reference's import order)
from splat_trainer.camera_table.camera_table import Cameras, Projections  # noqa: E402
from splat_trainer.util.colors import compute_psnr  # noqa: E402
from splat_trainer.visibility.cluster import PointClusters  # noqa: E402
from splat_trainer.visibility.query_points import point_visibility  # noqa: E402


def main():
  rng = np.random.default_rng(77)
  n, k = 5000, 64
  labels = torch.from_numpy(rng.integers(0, k, n).astype(np.int64))
  pc = PointClusters(labels, torch.zeros((k, 3)))
  v = 3000
  idx = torch.from_numpy(np.sort(rng.choice(n, v, replace=False)).astype(np.int64))
  vis = torch.from_numpy((rng.uniform(0, 1, v) ** 3).astype(np.float32))
  feats = pc.view_features(idx, vis)  # default threshold 0.01
  feats5 = pc.view_features(idx, vis, vis_threshold=0.5)

  W, H = 320, 240
  cams = syn.ring_cameras(12, W, H, seed=3)
  T = torch.tensor(np.stack([c["T"] for c in cams]), dtype=torch.float32)
  intr = torch.tensor([[c["fx"], c["fy"], c["cx"], c["cy"]] for c in cams], dtype=torch.float32)
  m = len(cams)
  proj = Projections(intrinsics=intr, image_size=torch.tensor([[W, H]] * m, dtype=torch.int32),
                     depth_range=torch.tensor([[0.1, 100.0]] * m, dtype=torch.float32), batch_size=(m,))
  cameras = Cameras(camera_t_world=T, projection=proj, camera_idx=torch.arange(m).unsqueeze(1), frame_idx=torch.arange(m).unsqueeze(1),
                    labels=torch.ones((m, 1), dtype=torch.int64), image_names=[f"im{i}" for i in range(m)], batch_size=(m,))
  pts = torch.from_numpy((rng.normal(0, 1, (4000, 3)) * np.array([3.0, 3.0, 1.5])).astype(np.float32))
  counts = point_visibility(cameras, pts)

  a = torch.from_numpy(rng.uniform(0, 1, (48, 64, 3)).astype(np.float32))
  b = torch.from_numpy(np.clip(a.numpy() + rng.normal(0, 0.05, a.shape), 0, 1).astype(np.float32))
  np.savez(os.path.join(OUT, "visibility.npz"), labels=labels.numpy().astype(np.int32), k=k, idx=idx.numpy(), vis=vis.numpy(),
           features=feats.numpy(), features_thr05=feats5.numpy(),
           cam_T=T.numpy(), cam_intrinsics=intr.numpy(), W=W, H=H, near=0.1, far=100.0, points=pts.numpy(),
           counts=counts.numpy().astype(np.int32),
           img_a=a.numpy(), img_b=b.numpy(), psnr=float(compute_psnr(a, b)), l1=float(torch.nn.functional.l1_loss(a, b)))
  print("visibility.npz:", int((feats > 0).sum()), "clusters hit,", int((counts > 0).sum()), "points seen, psnr", float(compute_psnr(a, b)))


if __name__ == "__main__":
  main()
