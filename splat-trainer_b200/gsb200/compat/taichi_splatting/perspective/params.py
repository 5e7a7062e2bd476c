"""CameraParams (reference trainer/trainer.py:291-301, camera_table/camera_table.py:167-175)."""
from dataclasses import dataclass, replace

import torch


@dataclass
class CameraParams:
  T_camera_world: torch.Tensor      # (4,4) x_cam = R x_world + t
  projection: torch.Tensor          # (4,) fx fy cx cy
  image_size: tuple                 # (W, H)
  near_plane: float = 0.1
  far_plane: float = 100.0

  @property
  def device(self):
    return self.T_camera_world.device

  @property
  def T_world_camera(self):
    return torch.linalg.inv(self.T_camera_world)

  @property
  def camera_position(self):
    return self.T_world_camera[:3, 3]

  @property
  def focal_length(self):
    return self.projection[0:2]

  @property
  def principal_point(self):
    return self.projection[2:4]

  def to(self, device=None, dtype=None):
    return replace(self, T_camera_world=self.T_camera_world.to(device=device, dtype=dtype),
                   projection=self.projection.to(device=device, dtype=dtype))

  def scale_size(self, s):
    return replace(self, projection=self.projection * s,
                   image_size=(int(round(self.image_size[0] * s)), int(round(self.image_size[1] * s))))
