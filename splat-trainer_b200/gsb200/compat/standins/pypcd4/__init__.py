"""pypcd4 stand-in: import stub only (the calls behind it are replaced by gsb200: io.py / ops.kmeans / ops.estimate_scale)."""


class PointCloud:
  """named in annotations only (util/pointcloud.py:108); .pcd files are outside the hot path"""

  @staticmethod
  def from_path(filename):
    raise RuntimeError("pypcd4 is absent: .pcd point clouds are not supported by the stand-in")

  from_points = from_path
