"""`@tensorclass` stand-in: a record of named tensors sharing leading batch dimensions (controller/point_state.py:10-32,
scene/point_statistics.py:8, scene/color_model.py:78, camera_table/camera_table.py:44-120, util/pointcloud.py:12).
Supports what the reference does with such records: keyword construction with `batch_size=`, attribute access,
indexing by mask / index / slice, `.to`, `.detach`, `.clone`, `.replace`, `.apply`, `.to_tensordict`, `.state_dict` /
`.load_state_dict`, `torch.cat` / `torch.stack`, `.unsqueeze`, `.item`, `.batch_size`, `.device`."""
import typing

import torch


def _is_batched(v):
  return isinstance(v, torch.Tensor) or hasattr(v, "_tc_fields")


def _make(cls):
  fields = [k for k in typing.get_type_hints(cls, include_extras=False).keys() if not k.startswith("_")] \
      if _hints_ok(cls) else list(getattr(cls, "__annotations__", {}).keys())
  defaults = {k: getattr(cls, k) for k in fields if hasattr(cls, k) and not isinstance(getattr(cls, k), property)}

  def __init__(self, *args, batch_size=None, device=None, **kwargs):
    values = dict(zip(fields, args))
    values.update(kwargs)
    for k in fields:
      if k not in values:
        if k in defaults:
          values[k] = defaults[k]
        else:
          raise TypeError(f"{cls.__name__}: missing field {k}")
    for k, v in values.items():
      if device is not None and _is_batched(v):
        v = v.to(device)
      object.__setattr__(self, k, v)
    if batch_size is None:
      batch_size = ()
      for k in fields:
        v = values[k]
        if isinstance(v, torch.Tensor) and v.ndim > 0:
          batch_size = (v.shape[0],)
          break
    object.__setattr__(self, "batch_size", torch.Size(batch_size))

  def _rebuild(self, fn):
    nb = len(self.batch_size)
    vals = {k: (fn(getattr(self, k)) if _is_batched(getattr(self, k)) else getattr(self, k)) for k in fields}
    out = cls.__new__(cls)
    for k, v in vals.items():
      object.__setattr__(out, k, v)
    bs = ()
    for v in vals.values():
      if isinstance(v, torch.Tensor):
        bs = tuple(v.shape[:nb]) if v.ndim >= nb else tuple(v.shape)
        break
    object.__setattr__(out, "batch_size", torch.Size(bs))
    return out

  def __getitem__(self, idx):
    nb = len(self.batch_size)
    vals = {k: (getattr(self, k)[idx] if _is_batched(getattr(self, k)) else getattr(self, k)) for k in fields}
    out = cls.__new__(cls)
    for k, v in vals.items():
      object.__setattr__(out, k, v)
    probe = torch.empty(tuple(self.batch_size) + (0,))[idx]
    object.__setattr__(out, "batch_size", torch.Size(probe.shape[:-1]))
    return out

  def __setitem__(self, idx, value):
    for k in fields:
      v = getattr(self, k)
      if _is_batched(v):
        v[idx] = getattr(value, k) if hasattr(value, k) else value[k]

  def __len__(self):
    return self.batch_size[0] if len(self.batch_size) else 0

  def to(self, *a, **kw):
    return _rebuild(self, lambda t: t.to(*a, **kw))

  def apply(self, fn, batch_size=None):
    out = _rebuild(self, fn)
    if batch_size is not None:
      object.__setattr__(out, "batch_size", torch.Size(batch_size))
    return out

  def replace(self, **kw):
    out = _rebuild(self, lambda t: t)
    for k, v in kw.items():
      object.__setattr__(out, k, v)
    return out

  def items(self):
    return [(k, getattr(self, k)) for k in fields]

  def to_dict(self):
    return {k: (v.to_dict() if hasattr(v, "_tc_fields") else v) for k, v in items(self)}

  def to_tensordict(self):
    from . import TensorDict
    return TensorDict({k: v for k, v in items(self) if _is_batched(v)}, batch_size=self.batch_size)

  def state_dict(self):
    return to_dict(self)

  def load_state_dict(self, state):
    for k, v in state.items():
      cur = getattr(self, k)
      if isinstance(cur, torch.Tensor):
        cur.copy_(v)
      elif hasattr(cur, "_tc_fields"):
        cur.load_state_dict(v)
      else:
        object.__setattr__(self, k, v)
    return self

  def item(self):
    return _rebuild(self, lambda t: t)

  def __repr__(self):
    body = ", ".join(f"{k}={tuple(v.shape) if isinstance(v, torch.Tensor) else v!r}" for k, v in items(self))
    return f"{cls.__name__}({body}, batch_size={tuple(self.batch_size)})"

  @classmethod
  def __torch_function__(klass, func, types, args=(), kwargs=None):
    kwargs = kwargs or {}
    if func in (torch.cat, torch.stack):
      objs = args[0]
      dim = kwargs.get("dim", args[1] if len(args) > 1 else 0)
      first = objs[0]
      vals = {k: (func([getattr(o, k) for o in objs], dim=dim) if _is_batched(getattr(first, k)) else getattr(first, k))
              for k in fields}
      out = cls.__new__(cls)
      for k, v in vals.items():
        object.__setattr__(out, k, v)
      probe = func([torch.empty(tuple(o.batch_size) + (0,)) for o in objs], dim=dim)
      object.__setattr__(out, "batch_size", torch.Size(probe.shape[:-1]))
      return out
    return NotImplemented

  ns = dict(__init__=__init__, __getitem__=__getitem__, __setitem__=__setitem__, __len__=__len__, to=to, apply=apply,
            replace=replace, items=items, to_dict=to_dict, to_tensordict=to_tensordict, state_dict=state_dict,
            load_state_dict=load_state_dict, item=item, __repr__=__repr__, __torch_function__=__torch_function__,
            detach=lambda self: _rebuild(self, lambda t: t.detach()), clone=lambda self: _rebuild(self, lambda t: t.clone()),
            cpu=lambda self: _rebuild(self, lambda t: t.cpu()), cuda=lambda self: _rebuild(self, lambda t: t.cuda()),
            unsqueeze=lambda self, d: _rebuild(self, lambda t: t.unsqueeze(d)),
            contiguous=lambda self: _rebuild(self, lambda t: t.contiguous()),
            keys=lambda self: list(fields), values=lambda self: [getattr(self, k) for k in fields],
            _tc_fields=tuple(fields))
  for k, v in ns.items():
    if k not in cls.__dict__ or k in ("__init__", "__repr__"):
      setattr(cls, k, v)
  if "device" not in cls.__dict__:
    cls.device = property(lambda self: next((getattr(self, k).device for k in fields if _is_batched(getattr(self, k))), None))
  if "shape" not in cls.__dict__:
    cls.shape = property(lambda self: self.batch_size)
  return cls


def _hints_ok(cls):
  try:
    typing.get_type_hints(cls)
    return True
  except Exception:
    return False


def tensorclass(cls=None, **options):
  if cls is None:
    return lambda c: _make(c)
  return _make(cls)
