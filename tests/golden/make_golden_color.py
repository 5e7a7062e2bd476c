"""Golden vectors of the colour MLP path made by RUNNING the reference's own ColorModel
(scene/color_model.py, scene/mlp/torch_mlp.py, scene/mlp/rsh.py) — build container only:
`python tests/golden/make_golden_color.py` -> tests/golden/color_mlp.npz.

Stand-ins for packages that are absent here: beartype (identity decorator), tensordict.tensorclass (a plain attribute
bag), splat_trainer.config (VaryingFloat helpers that ColorModel.forward never calls).  The model code itself executes
unmodified from /root/reference.  Sizes follow config/scene/mlp.yaml:33-43 (point_features 16, image_features 32,
hidden 32, hidden_layers 1, sh_degree 5).  Three outputs are recorded: plain fp32, the same under
torch.autocast(float16) as MLPScene.eval_colors runs it (scene/mlp_scene.py:362), and fp64 autograd gradients."""
import importlib.util
import os
import sys
import types

import numpy as np
import torch

REF = "/root/reference/splat_trainer"
OUT = os.path.dirname(os.path.abspath(__file__))


def load(path, name):
  spec = importlib.util.spec_from_file_location(name, os.path.join(REF, path))
  mod = importlib.util.module_from_spec(spec)
  sys.modules[name] = mod
  spec.loader.exec_module(mod)
  return mod


def tensorclass(cls):
  class Bag:
    def __init__(self, **kw):
      self.__dict__.update(kw)
  for k, v in cls.__dict__.items():
    if callable(v) and not k.startswith("__"):
      setattr(Bag, k, v)
  Bag.__name__ = cls.__name__
  return Bag


sys.modules["beartype"] = types.SimpleNamespace(beartype=lambda f: f)
sys.modules["tensordict"] = types.SimpleNamespace(tensorclass=tensorclass)
for pkg in ("splat_trainer", "splat_trainer.scene", "splat_trainer.scene.mlp"):
  sys.modules[pkg] = types.ModuleType(pkg)
sys.modules["splat_trainer.config"] = types.SimpleNamespace(VaryingFloat=float, eval_varying=lambda v, t: v,
                                                            schedule_groups=lambda *a, **k: None)
rsh = load("scene/mlp/rsh.py", "splat_trainer.scene.mlp.rsh")
sys.modules["splat_trainer.scene.mlp"].rsh = rsh
torch_mlp = load("scene/mlp/torch_mlp.py", "splat_trainer.scene.mlp.torch_mlp")
color_model = load("scene/color_model.py", "splat_trainer.scene.color_model")

NAMES = dict(base_w1="base_model.layers.0.m.weight", base_b1="base_model.layers.0.m.bias",
             base_w2="base_model.layers.1.weight", base_b2="base_model.layers.1.bias",
             dir_wsh="directional_model.encode_dir.mlp.layers.0.weight", dir_bsh="directional_model.encode_dir.mlp.layers.0.bias",
             dir_w1="directional_model.mlp.layers.0.m.weight", dir_b1="directional_model.mlp.layers.0.m.bias",
             dir_w2="directional_model.mlp.layers.1.weight", dir_b2="directional_model.mlp.layers.1.bias")


def main():
  torch.manual_seed(7)
  cfg = color_model.ColorModelConfig(hidden_features=32, hidden_layers=1, sh_degree=5)
  model = cfg.create_model(glo_features=32, point_features=16)
  sd = model.state_dict()
  assert set(NAMES.values()) == set(sd.keys()), sorted(sd.keys())
  # the freshly initialised model is nearly linear; scale the weights up so the GLU gates and the exp/sigmoid
  # activations leave their linear range and the test has teeth
  with torch.no_grad():
    for k, v in sd.items():
      v.mul_(2.5 if k.endswith("weight") else 1.0)
      if k.endswith("bias"):
        v.add_(0.3 * torch.randn_like(v))
  n = 301
  gen = torch.Generator().manual_seed(11)
  feature = torch.randn((n, 16), generator=gen) * 1.5
  position = torch.randn((n, 3), generator=gen) * 4.0
  cam_pos = torch.tensor([0.3, -1.2, 5.0])
  position[5] = cam_pos  # zero-length direction: F.normalize's eps branch
  glo = torch.randn((1, 32), generator=gen) * 0.7
  dirs = torch.nn.functional.normalize(torch.randn((64, 3), generator=gen), dim=1)
  out = {k: sd[v].numpy().copy() for k, v in NAMES.items()}
  out.update(feature=feature.numpy(), position=position.numpy(), cam_pos=cam_pos.numpy(), glo=glo.numpy()[0],
             dirs=dirs.numpy(), rsh5=rsh.rsh_cart_5(dirs.double()).numpy(), rsh3=rsh.rsh_cart_3(dirs.double()).numpy())
  with torch.no_grad():
    c = model(feature, position, cam_pos, glo)
    out.update(diffuse_f32=c.diffuse.numpy(), specular_f32=c.specular.numpy(), total_f32=c.total(1.0).numpy())
    with torch.autocast(device_type="cpu", dtype=torch.float16):
      c = model(feature, position, cam_pos, glo)
      out.update(diffuse_amp=c.diffuse.float().numpy(), specular_amp=c.specular.float().numpy(),
                 total_amp=c.total(1.0).float().numpy(), amp_dtype=str(c.diffuse.dtype))
  # gradients in float64 of  Σ cd·diffuse + cs·specular
  m64 = cfg.create_model(glo_features=32, point_features=16).double()
  m64.load_state_dict({k: v.double() for k, v in sd.items()})
  f64, g64 = feature.double().requires_grad_(True), glo.double().requires_grad_(True)
  cd, cs = torch.randn((n, 3), generator=gen).double(), torch.randn((n, 3), generator=gen).double()
  c = m64(f64, position.double(), cam_pos.double(), g64)
  ((c.diffuse * cd).sum() + (c.specular * cs).sum()).backward()
  out.update(cd=cd.numpy(), cs=cs.numpy(), d_feature=f64.grad.numpy(), d_glo=g64.grad.numpy()[0],
             diffuse_f64=c.diffuse.detach().numpy(), specular_f64=c.specular.detach().numpy())
  params = dict(m64.named_parameters())
  for k, v in NAMES.items():
    out["d_" + k] = params[v].grad.numpy()
  np.savez_compressed(os.path.join(OUT, "color_mlp.npz"), **out)
  print("color_mlp.npz:", {k: getattr(v, "shape", v) for k, v in out.items()})


if __name__ == "__main__":
  main()
