"""roma stand-in: the two calls the reference makes (gaussians/split.py:11,17; camera_table/pose_table.py:8)."""
import torch


def unitquat_to_rotmat(q):
  """xyzw unit quaternion (..., 4) -> rotation matrix (..., 3, 3) (roma's documented convention)."""
  x, y, z, w = q.unbind(-1)
  R = torch.stack([1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y),
                   2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x),
                   2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)], dim=-1)
  return R.reshape(q.shape[:-1] + (3, 3))


def rotmat_to_unitquat(R):
  """rotation matrix (..., 3, 3) -> xyzw unit quaternion, w >= 0 branch-free form (Shepperd's method per element)."""
  m = R.reshape(-1, 3, 3)
  t = m[:, 0, 0] + m[:, 1, 1] + m[:, 2, 2]
  q = torch.zeros((m.shape[0], 4), dtype=R.dtype, device=R.device)
  for i in range(m.shape[0]):
    a = m[i]
    if t[i] > 0:
      s = torch.sqrt(t[i] + 1.0) * 2
      q[i] = torch.stack([(a[2, 1] - a[1, 2]) / s, (a[0, 2] - a[2, 0]) / s, (a[1, 0] - a[0, 1]) / s, 0.25 * s])
    elif a[0, 0] > a[1, 1] and a[0, 0] > a[2, 2]:
      s = torch.sqrt(1.0 + a[0, 0] - a[1, 1] - a[2, 2]) * 2
      q[i] = torch.stack([0.25 * s, (a[0, 1] + a[1, 0]) / s, (a[0, 2] + a[2, 0]) / s, (a[2, 1] - a[1, 2]) / s])
    elif a[1, 1] > a[2, 2]:
      s = torch.sqrt(1.0 + a[1, 1] - a[0, 0] - a[2, 2]) * 2
      q[i] = torch.stack([(a[0, 1] + a[1, 0]) / s, 0.25 * s, (a[1, 2] + a[2, 1]) / s, (a[0, 2] - a[2, 0]) / s])
    else:
      s = torch.sqrt(1.0 + a[2, 2] - a[0, 0] - a[1, 1]) * 2
      q[i] = torch.stack([(a[0, 2] + a[2, 0]) / s, (a[1, 2] + a[2, 1]) / s, 0.25 * s, (a[1, 0] - a[0, 1]) / s])
  return q.reshape(R.shape[:-2] + (4,))


def quat_product(p, q):
  px, py, pz, pw = p.unbind(-1)
  qx, qy, qz, qw = q.unbind(-1)
  return torch.stack([pw * qx + px * qw + py * qz - pz * qy, pw * qy - px * qz + py * qw + pz * qx,
                      pw * qz + px * qy - py * qx + pz * qw, pw * qw - px * qx - py * qy - pz * qz], dim=-1)
