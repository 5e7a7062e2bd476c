"""Generate golden vectors by RUNNING the reference's own Python code (run in the build container only:
`python tests/golden/make_golden.py`; /root/reference does not exist on the GPU box, the .npz files travel).

The reference modules import packages that are absent here (tensordict, roma, taichi_splatting, ...), so the
functions under test are loaded from their source files with minimal stand-ins injected for those imports:
  * roma.unitquat_to_rotmat  — xyzw unit quaternion -> rotation matrix (roma's documented convention)
  * tensordict.TensorDict    — a dict with .apply(fn, batch_size=), .update(dict), .batch_size
  * torch.compile            — identity (the functions are plain eager torch)
  * torch.multinomial        — inverse-CDF draw from RECORDED uniforms, so the random split axis can be replayed by
                               the oracle and the CUDA kernel (same distribution as the reference's multinomial)
Everything else executes the reference's code unmodified, from where it lies under /root/reference.
"""
import ast
import importlib.util
import math
import os
import sys
import types

import numpy as np
import torch

REF = "/root/reference/splat_trainer"
OUT = os.path.dirname(os.path.abspath(__file__))

torch.compile = lambda *a, **k: (a[0] if a and callable(a[0]) else (lambda f: f))


# ---- stand-ins ------------------------------------------------------------------------------------------
def unitquat_to_rotmat(q):
  x, y, z, w = q.unbind(-1)
  R = torch.stack([1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y),
                   2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x),
                   2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)], dim=-1)
  return R.reshape(q.shape[:-1] + (3, 3))


class TensorDict(dict):
  def __init__(self, d=None, batch_size=None):
    super().__init__(d or {})
    self.batch_size = list(batch_size) if batch_size is not None else [next(iter(self.values())).shape[0]]

  def apply(self, fn, batch_size=None):
    return TensorDict({k: fn(v) for k, v in self.items()}, batch_size)

  def update(self, d):
    out = TensorDict(dict(self), self.batch_size)
    dict.update(out, d)
    return out


sys.modules["roma"] = types.SimpleNamespace(unitquat_to_rotmat=unitquat_to_rotmat)
sys.modules["tensordict"] = types.SimpleNamespace(TensorDict=TensorDict)


def load(path, name):
  spec = importlib.util.spec_from_file_location(name, os.path.join(REF, path))
  mod = importlib.util.module_from_spec(spec)
  spec.loader.exec_module(mod)
  return mod


def extract(path, names, env):
  """exec only the named top-level functions / class methods of a reference file (its module imports fail here)."""
  src = open(os.path.join(REF, path)).read()
  tree = ast.parse(src)
  for node in ast.walk(tree):
    if isinstance(node, ast.FunctionDef) and node.name in names:
      node.decorator_list = []
      node.returns = None  # annotations name types of the absent packages (Rendering, ...)
      for a in node.args.args + node.args.kwonlyargs:
        a.annotation = None
      code = compile(ast.Module([node], []), os.path.join(REF, path), "exec")
      exec(code, env)
  return env


def main():
  rng = np.random.default_rng(2026)
  f32 = lambda a: torch.from_numpy(np.ascontiguousarray(a, np.float32))

  # ---- SH basis: scene/mlp/rsh.py:110-157 ------------------------------------------------------------------
  rsh = load("scene/mlp/rsh.py", "ref_rsh")
  dirs = rng.normal(0, 1, (256, 3)); dirs /= np.linalg.norm(dirs, axis=1, keepdims=True)
  dirs[:6] = np.array([[1, 0, 0], [0, 1, 0], [0, 0, 1], [-1, 0, 0], [0, -1, 0], [0, 0, -1]])
  Y = rsh.rsh_cart_3(f32(dirs)).numpy()
  misc = load("util/misc.py", "ref_misc")
  rgb = rng.uniform(0, 1, (64, 3)).astype(np.float32)
  np.savez(os.path.join(OUT, "sh_basis.npz"), dirs=dirs.astype(np.float32), Y=Y, rgb=rgb,
           rgb_to_sh=misc.rgb_to_sh(f32(rgb)).numpy(), sh0=np.float32(misc.sh0))

  # ---- exp_lerp: util/misc.py:56-59 ---------------------------------------------------------------------------
  a = rng.normal(0, 3, 512).astype(np.float32); b = rng.normal(0, 3, 512).astype(np.float32)
  a[:4] = b[:4]  # exp_lerp(t, a, a) = a
  outs = {f"t{t}": misc.exp_lerp(t, f32(a), f32(b)).numpy() for t in (0.0, 0.01, 0.1, 0.5, 1.0)}
  np.savez(os.path.join(OUT, "exp_lerp.npz"), a=a, b=b, **outs)

  # ---- take_n: controller/target_controller.py:150-160 -------------------------------------------------------
  env = extract("controller/target_controller.py", {"take_n"}, dict(torch=torch))
  vals = rng.normal(0, 1, 4000).astype(np.float32)  # distinct values: argsort ties are unspecified in torch
  cases = {}
  for k in (0, 1, 37, 2000, 4000):
    for desc in (False, True):
      cases[f"k{k}_d{int(desc)}"] = env["take_n"](f32(vals), k, descending=desc).numpy()
  np.savez(os.path.join(OUT, "take_n.npz"), values=vals, **cases)

  # ---- split: gaussians/split.py:15-19, 87-113 --------------------------------------------------------------------
  split = load("gaussians/split.py", "ref_split")
  n = 300
  pts = dict(position=rng.normal(0, 1, (n, 3)), log_scaling=rng.normal(-2, 1, (n, 3)), rotation=rng.normal(0, 1, (n, 4)),
             alpha_logit=rng.normal(0, 1, (n, 1)), feature=rng.normal(0, 1, (n, 48)))
  pts["log_scaling"][:20] = -11.0  # below the eps clamp
  td = TensorDict({k: f32(v) for k, v in pts.items()}, [n])
  basis = split.point_basis(td["log_scaling"], td["rotation"]).numpy()
  out_arg = split.split_gaussians_uniform(td, k=2, random_axis=False)
  uniform = rng.uniform(0, 1, n).astype(np.float32)

  def multinomial_from_uniform(probs, num_samples=1, **kw):
    assert num_samples == 1
    cdf = torch.cumsum(probs / probs.sum(dim=1, keepdim=True), dim=1)
    u = f32(uniform).unsqueeze(1)
    return (u >= cdf[:, :2]).sum(dim=1, keepdim=True)  # inverse CDF over 3 classes

  real = torch.multinomial
  torch.multinomial = multinomial_from_uniform
  out_rand = split.split_gaussians_uniform(td, k=2, random_axis=True)
  torch.multinomial = real
  np.savez(os.path.join(OUT, "split.npz"), uniform=uniform, basis=basis,
           **{f"in_{k}": v.astype(np.float32) for k, v in pts.items()},
           **{f"argmax_{k}": v.numpy() for k, v in out_arg.items()},
           **{f"random_{k}": v.numpy() for k, v in out_rand.items()})

  # ---- PointState.add_rendering / masked_heuristics: controller/point_state.py:34-57 ------------------------
  env = extract("controller/point_state.py", {"add_rendering", "masked_heuristics"}, dict(torch=torch, exp_lerp=misc.exp_lerp))
  N, m = 3000, 1100
  state = types.SimpleNamespace(prune_cost=f32(rng.uniform(0, 1, N)), split_score=f32(rng.uniform(0, 1, N)),
                                max_scale_px=f32(rng.uniform(0, 3, N)),
                                points_in_view=torch.from_numpy(rng.integers(0, 9, N).astype(np.int16)),
                                visibility=f32(rng.uniform(0, 1, N)))
  before = {k: getattr(state, k).clone().numpy() for k in ("prune_cost", "split_score", "max_scale_px", "points_in_view", "visibility")}
  idx = torch.from_numpy(np.sort(rng.choice(N, m, replace=False)))
  vis = np.where(rng.uniform(0, 1, m) < 0.7, rng.uniform(0, 20, m), 0).astype(np.float32)
  scale = rng.uniform(0.3, 40, (m, 2)).astype(np.float32)
  pc, ss = rng.uniform(0, 3, m).astype(np.float32), rng.uniform(0, 3, m).astype(np.float32)
  vmask = torch.from_numpy(vis > 0)
  points = types.SimpleNamespace(idx=idx, visibility=f32(vis), screen_scale=f32(scale), split_score=f32(ss), prune_cost=f32(pc),
                                 visible=types.SimpleNamespace(idx=idx[vmask]))
  env["add_rendering"](state, types.SimpleNamespace(points=points))
  after = {k: getattr(state, k).numpy() for k in before}
  mp, ms = env["masked_heuristics"](state, 5)
  np.savez(os.path.join(OUT, "point_state.npz"), idx=idx.numpy(), vis=vis, screen_scale=scale, prune_cost=pc, split_score=ss,
           masked_prune=mp.numpy(), masked_split=ms.numpy(),
           **{f"before_{k}": v for k, v in before.items()}, **{f"after_{k}": v for k, v in after.items()})

  # ---- TargetController.find_split_prune_indexes: controller/target_controller.py:73-96 ----------------------
  env = extract("controller/target_controller.py", {"take_n", "find_split_prune_indexes"}, dict(torch=torch, math=math))
  ps = types.SimpleNamespace(**{k: torch.from_numpy(v.copy()) for k, v in after.items()})
  ps.batch_size = [N]
  ps.masked_heuristics = lambda mv: extract("controller/point_state.py", {"masked_heuristics"}, dict(torch=torch))["masked_heuristics"](ps, mv)
  ps.max_scale_px[:7] = 500.0
  stats_in = {k: getattr(ps, k).clone().numpy() for k in after}
  cfg = types.SimpleNamespace(max_scale_px=200.0, prune_rate=0.025, min_views=5, min_split_px=0.0)
  ctrl = types.SimpleNamespace(config=cfg, points=ps)
  split_mask, prune_mask = env["find_split_prune_indexes"](ctrl, 0.3, 3300)
  np.savez(os.path.join(OUT, "find_split_prune.npz"), t=np.float32(0.3), target_points=np.int64(3300),
           split_mask=split_mask.numpy(), prune_mask=prune_mask.numpy(), **{f"in_{k}": v for k, v in stats_in.items()})
  # ---- MCMC controller: controller/mcmc_controller.py:69-100 -----------------------------------------------------
  # noise branch (:91-100), executed with the reference's own soft_lt and sample_gaussians on recorded normals
  n = 500
  al = rng.normal(-1.0, 2.0, (n, 1)).astype(np.float32)
  ls = rng.normal(-2, 1, (n, 3)).astype(np.float32); q = rng.normal(0, 1, (n, 4)).astype(np.float32)
  pos = rng.normal(0, 1, (n, 3)).astype(np.float32)
  piv = rng.integers(0, 12, n).astype(np.int16)
  randn = rng.normal(0, 1, (n, 3)).astype(np.float32)
  thr, min_views, level = 0.1, 5, 0.37
  enough = torch.from_numpy(piv.astype(np.int64)) > min_views
  opacity = torch.sigmoid(f32(al)).squeeze(1)
  target = misc.soft_lt(opacity, thr / 2, margin=16.0)
  sub = TensorDict(dict(log_scaling=f32(ls)[enough], rotation=f32(q)[enough]), [int(enough.sum())])
  noise = f32(randn)[enough] * (target[enough] * level).unsqueeze(1)
  out_pos = f32(pos).clone()
  out_pos[enough] += split.sample_gaussians(sub, noise)
  # prune / respawn branch (:76-89) with the reference's take_n
  tn = extract("controller/target_controller.py", {"take_n"}, dict(torch=torch))["take_n"]
  max_scale = f32(rng.uniform(0, 260, n)); sscore = f32(rng.uniform(0, 3, n))
  prune_mask = (max_scale > 200.0) | (opacity < thr)
  k = int(prune_mask.sum().item())
  too_small = max_scale < 1.5
  masked = torch.where(prune_mask | too_small, 0, sscore)
  split_mask = tn(masked, k, descending=True)
  np.savez(os.path.join(OUT, "mcmc.npz"), alpha_logit=al, log_scaling=ls, rotation=q, position=pos, points_in_view=piv,
           randn=randn, opacity_threshold=np.float32(thr), min_views=np.int32(min_views), noise_level=np.float32(level),
           out_position=out_pos.numpy(), max_scale_px=max_scale.numpy(), split_score=sscore.numpy(),
           min_split_px=np.float32(1.5), prune_mask=prune_mask.numpy(), split_mask=split_mask.numpy())

  # ---- regularisation: scene/mlp_scene.py:246-265 (compute_reg) + util/misc.py:68-69 (saturate) ------------------
  env = extract("scene/mlp_scene.py", {"compute_reg"}, dict(torch=torch, saturate=misc.saturate))
  m = 700
  op = torch.from_numpy(rng.uniform(0.01, 0.99, m).astype(np.float32)).requires_grad_()
  lsr = torch.from_numpy(rng.normal(-2, 1, (m, 3)).astype(np.float32)).requires_grad_()
  dep = torch.from_numpy(rng.uniform(0.3, 20, (m, 1)).astype(np.float32)).requires_grad_()
  wv = f32(rng.uniform(1e-3, 30, m))
  regs = env["compute_reg"](None, op, lsr, dep, torch.zeros((m, 1)), wv)
  weights = dict(scale=0.01, opacity=1.0, aspect=1e-4)  # config/scene/mlp.yaml:16-19 at t = 0
  total = sum(regs[k] * w for k, w in weights.items())
  total.backward()
  np.savez(os.path.join(OUT, "reg_loss.npz"), opacity=op.detach().numpy(), log_scaling=lsr.detach().numpy(),
           depths=dep.detach().numpy(), weight=wv.numpy(), w_scale=np.float32(0.01), w_opacity=np.float32(1.0),
           w_aspect=np.float32(1e-4), reg_scale=regs["scale"].item(), reg_opacity=regs["opacity"].item(),
           reg_aspect=regs["aspect"].item(), total=total.item(), d_opacity=op.grad.numpy(), d_log_scaling=lsr.grad.numpy(),
           d_depths=dep.grad.numpy())
  print("golden vectors written to", OUT)


if __name__ == "__main__":
  main()
