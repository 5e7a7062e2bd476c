"""tensordict stand-in: the sliver of `TensorDict`, `@tensorclass` and `TensorDictParams` the reference uses
(controller/point_state.py:8-32,107-110; scene/mlp_scene.py:74-78,291-298,395; camera_table/camera_table.py:10-11;
gaussians/split.py:8,39-48; scene/color_model.py:78).  Plain Python over torch tensors; not on the hot path."""
import dataclasses

import torch
from torch import nn

from . import tensorclass as _tc_module  # so `from tensordict.tensorclass import tensorclass` works (util/pointcloud.py:1)


def _first_dim(values):
  for v in values:
    if isinstance(v, torch.Tensor) and v.ndim > 0:
      return [v.shape[0]]
    if isinstance(v, TensorDict):
      return list(v.batch_size)
  return []


class TensorDict(dict):
  """dict of tensors (and nested TensorDicts) sharing leading batch dimensions."""

  def __init__(self, source=None, batch_size=None, device=None, **kwargs):
    items = dict(source or {})
    items.update(kwargs)
    super().__init__({k: (TensorDict(v) if isinstance(v, dict) and not isinstance(v, TensorDict) else v) for k, v in items.items()})
    self.batch_size = torch.Size(batch_size if batch_size is not None else _first_dim(self.values()))
    if device is not None:
      for k, v in list(self.items()):
        dict.__setitem__(self, k, v.to(device))

  # ---- functional helpers -------------------------------------------------------------------------------
  def _like(self, items, batch_size=None):
    return TensorDict(items, batch_size=self.batch_size if batch_size is None else batch_size)

  def apply(self, fn, *others, batch_size=None, **kwargs):
    out = {}
    for k, v in self.items():
      rest = [o[k] for o in others]
      out[k] = v.apply(fn, *rest) if isinstance(v, TensorDict) else fn(v, *rest)
    return TensorDict(out, batch_size=batch_size if batch_size is not None else None)

  def replace(self, *args, **kwargs):
    items = dict(self)
    for a in args:
      items.update(a)
    items.update(kwargs)
    return self._like(items)

  def update(self, other=None, **kwargs):
    for k, v in dict(other or {}, **kwargs).items():
      dict.__setitem__(self, k, v)
    return self

  def select(self, *keys):
    return self._like({k: self[k] for k in keys})

  def exclude(self, *keys):
    return self._like({k: v for k, v in self.items() if k not in keys})

  def to(self, *args, **kwargs):
    return self._like({k: v.to(*args, **kwargs) for k, v in self.items()})

  def to_dict(self):
    return {k: (v.to_dict() if isinstance(v, TensorDict) else v) for k, v in self.items()}

  def to_tensordict(self):
    return self

  def detach(self):
    return self.apply(lambda t: t.detach())

  def clone(self):
    return self.apply(lambda t: t.clone())

  def cpu(self):
    return self.to("cpu")

  @property
  def device(self):
    for v in self.values():
      return v.device
    return None

  @property
  def batch_dims(self):
    return len(self.batch_size)

  @property
  def shape(self):
    return self.batch_size

  def __getitem__(self, key):
    if isinstance(key, str):
      return dict.__getitem__(self, key)
    if isinstance(key, tuple) and key and all(isinstance(k, str) for k in key):
      out = self
      for k in key:
        out = out[k]
      return out
    return TensorDict({k: v[key] for k, v in self.items()})

  def __setitem__(self, key, value):
    if isinstance(key, str):
      dict.__setitem__(self, key, value)
    else:  # masked / indexed assignment of a TensorDict or dict of rows
      for k, v in self.items():
        v[key] = value[k]

  def numel(self):
    n = 1
    for b in self.batch_size:
      n *= b
    return n

  def state_dict(self):
    return self.to_dict()

  def load_state_dict(self, state):
    for k, v in state.items():
      if k in self:
        self[k].copy_(v) if isinstance(self[k], torch.Tensor) else self[k].load_state_dict(v)

  @classmethod
  def __torch_function__(cls, func, types, args=(), kwargs=None):
    kwargs = kwargs or {}
    if func in (torch.cat, torch.stack):
      tds = args[0]
      dim = kwargs.get("dim", args[1] if len(args) > 1 else 0)
      return TensorDict({k: func([t[k] for t in tds], dim=dim) for k in tds[0].keys()})
    return NotImplemented


def cat(tds, dim=0):
  return TensorDict({k: torch.cat([t[k] for t in tds], dim=dim) for k in tds[0].keys()})


tensorclass = _tc_module.tensorclass
TensorClass = object


class TensorDictParams(nn.Module):
  """nn.Module holding a TensorDict's tensors as parameters (camera_table/camera_table.py:11)."""

  def __init__(self, td, no_convert=False):
    super().__init__()
    self._keys = list(td.keys())
    for k, v in td.items():
      self.register_parameter(k, v if isinstance(v, nn.Parameter) else nn.Parameter(v, requires_grad=v.is_floating_point()))

  def __getitem__(self, k):
    return getattr(self, k)

  def keys(self):
    return list(self._keys)

  def items(self):
    return [(k, getattr(self, k)) for k in self._keys]
