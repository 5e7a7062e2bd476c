"""ParameterClass + visibility-aware per-point optimisers over gs_optim_step (reference scene/mlp_scene.py:21,
56-85,214-239; scene/util.py:4,28-48; controller/mcmc_controller.py:71-74,95).

ParameterClass holds named per-point tensors as nn.Parameters plus ONE optimiser with per-group options
(`parameter_groups = {name: dict(lr=..., type='scalar'|'vector'|'local_vector')}`, reference config/scene/mlp.yaml:8-14).
`step(visibility=, indexes=, basis=)` runs one fused kernel per group over the listed rows only."""
import torch

from gsb200 import lib as L
from gsb200 import ops

from ..data_types import TensorBag

_KIND = dict(scalar=L.GS_GROUP_SCALAR, vector=L.GS_GROUP_VECTOR, local_vector=L.GS_GROUP_LOCAL_VECTOR)


class VisibilityOptimizer(torch.optim.Optimizer):
  """Base of the visibility-aware optimisers.  param_groups entries: params=[p], name, lr, type, point_lr, mask_lr."""
  MODE = 0

  def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, vis_beta=0.999, vis_smooth=0.01, grad_clip=None,
               bias_correction=True):
    defaults = dict(lr=lr, betas=tuple(betas), eps=eps, vis_beta=vis_beta, vis_smooth=vis_smooth, grad_clip=grad_clip,
                    bias_correction=bias_correction, type="scalar", point_lr=None, mask_lr=None, name=None)
    super().__init__(params, defaults)
    self.shared = {}  # per-point state shared by all groups: running_vis, step

  def _shared_state(self, n, device):
    s = self.shared
    if "step" not in s or s["step"].shape[0] != n:
      s["step"] = torch.zeros((n,), device=device)
      s["running_vis"] = torch.zeros((n,), device=device)
    return s

  def _config(self, group):
    b1, b2 = group["betas"]
    clip = group["grad_clip"]
    return L.make_optim_config(mode=self.MODE, beta1=b1, beta2=b2, eps=group["eps"], vis_beta=group["vis_beta"],
                               vis_smooth=group["vis_smooth"], grad_clip=0.0 if clip is None else float(clip),
                               bias_correction=int(group["bias_correction"]))

  @torch.no_grad()
  def step(self, indexes, visibility=None, basis=None, basis_params=None):
    """indexes (V,) rows to update; visibility (V,) their weights (None = 1); basis is accepted for signature
    compatibility — the kernel rebuilds R(q)·diag(exp s) (gaussians/split.py:15-19) from basis_params =
    (log_scaling, rotation) taken BEFORE their own update."""
    groups = [g for g in self.param_groups if g["params"][0].grad is not None]
    if not groups or indexes.numel() == 0:
      return
    p0 = groups[0]["params"][0]
    n, dev = p0.shape[0], p0.device
    s = self._shared_state(n, dev)
    idx64 = indexes.to(torch.int64)
    visible = torch.zeros((n,), device=dev)
    visible[idx64] = 1.0 if visibility is None else visibility.to(torch.float32)
    vi = ops.VisibleIndexes(n, dev)
    vi.begin(self._config(groups[0]), visible, s["running_vis"], s["step"])
    # local_vector groups first: their basis uses log_scaling / rotation before those are stepped
    order = sorted(groups, key=lambda g: 0 if g["type"] == "local_vector" else 1)
    for g in order:
      p = g["params"][0]
      st = self.state[p]
      if "exp_avg" not in st:
        st["exp_avg"] = torch.zeros_like(p.data)
        st["exp_avg_sq"] = torch.zeros_like(p.data) if g["type"] == "scalar" else torch.zeros((n,), device=dev)
      kind = _KIND[g["type"]]
      bls, bq = (basis_params if kind == L.GS_GROUP_LOCAL_VECTOR else (None, None))
      if kind == L.GS_GROUP_LOCAL_VECTOR and basis_params is None:
        raise L.GsError("local_vector group needs basis_params=(log_scaling, rotation)")
      ops.optim_step(self._config(g), kind, float(g["lr"]), p.data, p.grad, st["exp_avg"], st["exp_avg_sq"], visible,
                     s["running_vis"], s["step"], basis_ls=bls, basis_q=bq, zero_grad=False, indexes=vi)


class VisibilityAwareAdam(VisibilityOptimizer):
  MODE = 0


class VisibilityAwareLaProp(VisibilityOptimizer):
  MODE = 1


class SparseAdam(VisibilityAwareAdam):
  """rows listed in `indexes` are updated with weight 1 (no visibility weighting)."""


class ParameterClass:
  def __init__(self, tensors, parameter_groups, optimizer=VisibilityAwareLaProp, **optim_options):
    self._params = {k: torch.nn.Parameter(v.detach().clone().contiguous(), requires_grad=k in parameter_groups)
                    for k, v in tensors.items()}
    self.parameter_groups = {k: dict(v) for k, v in parameter_groups.items()}
    self.optimizer_cls, self.optim_options = optimizer, dict(optim_options)
    groups = [dict(params=[self._params[k]], name=k, **{a: b for a, b in g.items()}) for k, g in self.parameter_groups.items()]
    self.optimizer = optimizer(groups, **optim_options)

  # ---- access --------------------------------------------------------------------------------------------
  def __getattr__(self, name):
    params = self.__dict__.get("_params", {})
    if name in params:
      return params[name]
    raise AttributeError(name)

  def keys(self):
    return self._params.keys()

  @property
  def tensors(self):
    return TensorBag({k: v for k, v in self._params.items()})

  @property
  def batch_size(self):
    return (next(iter(self._params.values())).shape[0],)

  @property
  def tensor_state(self):
    """per-tensor optimiser state (moments), keyed by tensor name"""
    return {k: dict(self.optimizer.state[p]) for k, p in self._params.items() if p in self.optimizer.state}

  # ---- optimiser -------------------------------------------------------------------------------------------
  def update_group(self, name, **kw):
    for g in self.optimizer.param_groups:
      if g["name"] == name:
        g.update(kw)

  def update_groups(self, **groups):
    for name, g in groups.items():
      self.update_group(name, **g)
    return {g["name"]: g["lr"] for g in self.optimizer.param_groups}

  def zero_grad(self):
    for p in self._params.values():
      p.grad = None

  def step(self, indexes, visibility=None, basis=None):
    bp = None
    if "log_scaling" in self._params and "rotation" in self._params:
      bp = (self._params["log_scaling"].data, self._params["rotation"].data)
    self.optimizer.step(indexes=indexes, visibility=visibility, basis=basis, basis_params=bp)

  # ---- row selection / growth (params AND optimiser state, reference scene/mlp_scene.py:306-310) ---------------------
  def _rebuild(self, tensors, states, shared):
    new = ParameterClass(tensors, self.parameter_groups, optimizer=self.optimizer_cls, **self.optim_options)
    for g_old, g_new in zip(self.optimizer.param_groups, new.optimizer.param_groups):
      for k, v in g_old.items():
        if k != "params":
          g_new[k] = v
    for k, st in states.items():
      new.optimizer.state[new._params[k]] = st
    new.optimizer.shared = shared
    return new

  def __getitem__(self, sel):
    tensors = {k: v.data[sel] for k, v in self._params.items()}
    states = {k: {a: b[sel] for a, b in st.items()} for k, st in self.tensor_state.items()}
    shared = {a: b[sel] for a, b in self.optimizer.shared.items()}
    return self._rebuild(tensors, states, shared)

  def append_tensors(self, tensors):
    """new rows get zero optimiser state"""
    m = next(iter(tensors.values())).shape[0]
    cat = {k: torch.cat([v.data, tensors[k].to(v.device, v.dtype)]) for k, v in self._params.items()}
    pad = lambda b: torch.cat([b, torch.zeros((m,) + tuple(b.shape[1:]), device=b.device, dtype=b.dtype)])
    states = {k: {a: pad(b) for a, b in st.items()} for k, st in self.tensor_state.items()}
    shared = {a: pad(b) for a, b in self.optimizer.shared.items()}
    return self._rebuild(cat, states, shared)

  def detach(self):
    return TensorBag({k: v.detach() for k, v in self._params.items()})

  # ---- checkpointing: plain tensors / scalars only (reference loads with weights_only=True) ----------------------------
  def state_dict(self):
    return dict(tensors={k: v.data for k, v in self._params.items()},
                parameter_groups={k: dict(v) for k, v in self.parameter_groups.items()},
                group_options=[{k: v for k, v in g.items() if k != "params"} for g in self.optimizer.param_groups],
                state={k: dict(st) for k, st in self.tensor_state.items()}, shared=dict(self.optimizer.shared))

  @staticmethod
  def from_state_dict(state, optimizer=VisibilityAwareLaProp, **optim_options):
    pc = ParameterClass(state["tensors"], state["parameter_groups"], optimizer=optimizer, **optim_options)
    for g, saved in zip(pc.optimizer.param_groups, state.get("group_options", [])):
      g.update({k: v for k, v in saved.items() if k in ("lr", "point_lr", "mask_lr")})
    for k, st in state.get("state", {}).items():
      pc.optimizer.state[pc._params[k]] = {a: b.to(pc._params[k].device) for a, b in st.items()}
    pc.optimizer.shared = {a: b for a, b in state.get("shared", {}).items()}
    return pc
