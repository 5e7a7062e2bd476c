"""pycolmap stand-in: import stub only (the calls behind it are replaced by gsb200: io.py / ops.kmeans / ops.estimate_scale)."""


class _Absent:
  """placeholder type for annotations (dataset/colmap/dataset.py:34); COLMAP dataset IO is out of scope"""

  def __init__(self, *a, **k):
    raise RuntimeError("pycolmap is absent: the COLMAP dataset reader is outside this library's scope")


def __getattr__(name):
  if name.startswith("__"):
    raise AttributeError(name)
  return _Absent
