"""pykeops.torch stand-in: importing LazyTensor succeeds; using it raises (gsb200.ops.kmeans / estimate_scale replace the
two call sites, visibility/cluster.py:139-180 and gaussians/loading.py:45-61)."""


class LazyTensor:
  def __init__(self, *args, **kwargs):
    raise RuntimeError("pykeops is absent: use gsb200.ops.kmeans / assign_clusters / estimate_scale")
