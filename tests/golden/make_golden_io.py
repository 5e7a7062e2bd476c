"""Golden vectors of the Gaussian PLY conversion made by RUNNING the reference's own to_plydata /
from_plydata (scene/io.py) — build container only.  plyfile and taichi_splatting are absent: plyfile is replaced by a
stand-in that keeps the structured vertex array (PlyElement.describe / PlyData([el])['vertex']), Gaussians3D by an
attribute bag.  What is pinned is the reference's field order, names and value conversions; the byte layout of the file
itself is plyfile's (PLY specification, binary_little_endian) and is unpinned here."""
import importlib.util
import os
import sys
import types

import numpy as np
import torch

OUT = os.path.dirname(os.path.abspath(__file__))


class Element:
  def __init__(self, data):
    self.data = data

  def __getitem__(self, k):
    return self.data[k]


class PlyData:
  def __init__(self, els):
    self.el = els[0]

  def __getitem__(self, k):
    assert k == "vertex"
    return self.el


class Bag:
  def __init__(self, **kw):
    self.__dict__.update(kw)

  def to(self, _):
    return self


sys.modules["plyfile"] = types.SimpleNamespace(PlyData=PlyData, PlyElement=types.SimpleNamespace(describe=lambda v, name: Element(v)))
sys.modules["taichi_splatting"] = types.SimpleNamespace(Gaussians3D=Bag)
spec = importlib.util.spec_from_file_location("ref_io", "/root/reference/splat_trainer/scene/io.py")
ref_io = importlib.util.module_from_spec(spec)
spec.loader.exec_module(ref_io)


def main():
  gen = torch.Generator().manual_seed(3)
  n = 57
  g = dict(position=torch.randn((n, 3), generator=gen), log_scaling=torch.randn((n, 3), generator=gen),
           rotation=torch.randn((n, 4), generator=gen) * 3, alpha_logit=torch.randn((n, 1), generator=gen))
  out = {k: v.numpy() for k, v in g.items()}
  for name, feat, with_sh in (("sh", torch.randn((n, 3, 16), generator=gen), True), ("feature", torch.randn((n, 16), generator=gen), False)):
    bag = Bag(feature=feat, batch_size=(n,), **g)
    ply = ref_io.to_plydata(bag, with_sh=with_sh)
    v = ply["vertex"].data
    out[name] = feat.numpy()
    out[f"{name}_names"] = np.array(v.dtype.names)
    out[f"{name}_table"] = np.stack([v[k] for k in v.dtype.names], axis=1)
    back = ref_io.from_plydata(ply, with_sh=with_sh)
    for k in ("position", "rotation", "alpha_logit", "log_scaling", "feature"):
      out[f"{name}_back_{k}"] = getattr(back, k).numpy()
  np.savez_compressed(os.path.join(OUT, "ply_io.npz"), **out)
  print({k: getattr(v, "shape", v) for k, v in out.items()})


if __name__ == "__main__":
  main()
