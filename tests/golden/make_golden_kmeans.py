"""Golden vectors of the point clustering made by RUNNING the reference's own kmeans_iter /
assign_clusters (visibility/cluster.py:139-171) — build container only: `python tests/golden/make_golden_kmeans.py`.

pykeops is absent here; its LazyTensor is replaced by a dense stand-in that evaluates the same expression
((x_i - c_j) ** 2).sum(-1).argmin(dim=1) with plain torch broadcasting (it keeps a VIEW of the centroid tensor, like
KeOps does, so the in-place centroid updates of the loop are seen).  The two functions themselves are executed
unmodified from the reference's source."""
import ast
import os
import sys

import numpy as np
import torch

REF = "/root/reference/splat_trainer"
OUT = os.path.dirname(os.path.abspath(__file__))


class LazyTensor:
  def __init__(self, t):
    self.f = lambda: t

  def _wrap(self, f):
    out = LazyTensor.__new__(LazyTensor)
    out.f = f
    return out

  def __sub__(self, other):
    return self._wrap(lambda: self.f() - other.f())

  def __pow__(self, p):
    return self._wrap(lambda: self.f() ** p)

  def sum(self, dim):
    return self._wrap(lambda: self.f().sum(dim))

  def argmin(self, dim):
    return self.f().argmin(dim=dim)

  def Kmin(self, k, dim):
    return torch.topk(self.f(), k, dim=dim, largest=False, sorted=True).values


def extract(path, names, env):
  tree = ast.parse(open(os.path.join(REF, path)).read())
  for node in tree.body:
    if isinstance(node, ast.FunctionDef) and node.name in names:
      node.decorator_list = []
      node.returns = None
      for a in node.args.args:
        a.annotation = None
      exec(compile(ast.Module([node], []), path, "exec"), env)
  return env


def main():
  env = extract("visibility/cluster.py", {"assign_clusters", "kmeans_iter"}, dict(torch=torch, LazyTensor=LazyTensor))
  gen = torch.Generator().manual_seed(5)
  n, k = 3000, 37
  blobs = torch.randn((12, 3), generator=gen) * 4
  x = blobs[torch.randint(0, 12, (n,), generator=gen)] + torch.randn((n, 3), generator=gen) * 0.6
  x[100] = x[7]  # duplicate point
  init = x[torch.randperm(n, generator=gen)[:k]].clone()
  # (no duplicate centroids: an emptied cluster becomes 0/0 = NaN, and what argmin does with a NaN distance differs
  # between this dense stand-in (torch picks the NaN) and KeOps' comparison-based reduction (never picks it) — that
  # corner is unpinned; the oracle and the kernel follow the comparison semantics)
  out = dict(x=x.numpy(), init=init.numpy())
  for iters in (1, 2, 10):
    labels, cent = env["kmeans_iter"](x, init.clone(), iters)
    out[f"labels_{iters}"] = labels.numpy().astype(np.int32)
    out[f"centroids_{iters}"] = cent.numpy()
  probe = torch.randn((500, 3), generator=gen) * 4
  good = torch.from_numpy(out["centroids_10"])
  out.update(probe=probe.numpy(), probe_labels=env["assign_clusters"](probe, good).numpy().astype(np.int32), probe_centroids=good.numpy())
  # estimate_scale (gaussians/loading.py:45-61) on a clumpy cloud with a duplicate and an outlier
  env2 = extract("gaussians/loading.py", {"estimate_scale"}, dict(torch=torch, LazyTensor=LazyTensor))
  pts = x[:1500].clone()
  pts[1499] = torch.tensor([60.0, -45.0, 30.0])

  class Cloud:
    points = pts
  for nn in (3, 5):
    out[f"knn_scale_{nn}"] = env2["estimate_scale"](Cloud, nn).numpy()
  out["knn_points"] = pts.numpy()
  np.savez_compressed(os.path.join(OUT, "kmeans.npz"), **out)
  print({k_: v.shape for k_, v in out.items()}, "empty clusters:", int(np.isnan(out["centroids_10"]).any(axis=1).sum()))


if __name__ == "__main__":
  main()
