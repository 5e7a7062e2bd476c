"""plyfile stand-in: import stub only (the calls behind it are replaced by gsb200: io.py / ops.kmeans / ops.estimate_scale)."""


class PlyElement:
  @staticmethod
  def describe(data, name):
    e = PlyElement()
    e.data, e.name = data, name
    return e


class PlyData:
  """vertex-table container; reading / writing the bytes is gsb200.io (own PLY reader / writer)."""

  def __init__(self, elements=(), text=False, byte_order="<"):
    self.elements = list(elements)

  def __getitem__(self, name):
    for e in self.elements:
      if e.name == name:
        return e
    raise KeyError(name)

  def write(self, filename):
    from gsb200 import io
    io.write_ply(str(filename), self["vertex"].data)

  @staticmethod
  def read(filename):
    from gsb200 import io
    return PlyData([PlyElement.describe(io.read_ply(str(filename)), "vertex")])
