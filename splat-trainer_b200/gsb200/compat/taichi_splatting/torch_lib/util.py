"""count_nonfinite / check_finite (reference trainer/trainer.py:54,581-586; scene/mlp_scene.py:27)."""
import torch


def count_nonfinite(t, name="tensor"):
  """-> {qualified name: number of non-finite entries} for every tensor with at least one, walking dicts."""
  out = {}
  if isinstance(t, torch.Tensor):
    if t.is_floating_point():
      n = int((~torch.isfinite(t)).sum().item())
      if n > 0:
        out[name] = n
  elif isinstance(t, dict):
    for k, v in t.items():
      out.update(count_nonfinite(v, f"{name}.{k}"))
  elif isinstance(t, (list, tuple)):
    for i, v in enumerate(t):
      out.update(count_nonfinite(v, f"{name}[{i}]"))
  return out


def check_finite(t, name="tensor", warn=False):
  bad = count_nonfinite(t, name)
  if bad:
    msg = ", ".join(f"{k}: {v} non-finite" for k, v in bad.items())
    if warn:
      import warnings
      warnings.warn(msg)
    else:
      raise ValueError(msg)
