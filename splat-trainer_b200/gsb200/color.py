"""ColorModule — the colour model of the reference's default scene (`scene: mlp`): GLO embedding table + the two MLPs
of ColorModel (reference scene/color_model.py:11-46,88-170, sizes config/scene/mlp.yaml:33-43).

The per-splat work (LayerNorm, SH, both MLPs forward and backward) is gs_color_mlp_fwd / gs_color_mlp_bwd
(csrc/color_mlp.cu).  What lives here is the small dense state around it: ten weight arrays (10k floats), their Adam
(reference ColorModel.optimizer, betas (0.9, 0.999)) and the per-image GLO rows with their sparse Adam
(GLOTable.optimizer, betas (0.8, 0.95)) — a handful of elementwise torch ops on one flat tensor each."""
import math

import torch

from . import ops

SHAPES = dict(base_w1=(64, 48), base_b1=(64,), base_w2=(4, 32), base_b2=(4,), dir_wsh=(96, 36), dir_bsh=(96,),
              dir_w1=(64, 48), dir_b1=(64,), dir_w2=(4, 32), dir_b2=(4,))
POINT_FEATURES, GLO_FEATURES = 16, 32


class ColorModule:
  def __init__(self, num_images, device="cuda", seed=0, lr=1e-2, lr_glo=2.0, weights=None, specular_reg=1e-5):
    """weights: optional dict of arrays shaped like SHAPES (e.g. a reference checkpoint's ColorModel state);
    default = nn.Linear's initialisation (U(-1/sqrt(in), 1/sqrt(in)) for weight and bias)."""
    self.device = torch.device(device)
    total = sum(math.prod(s) for s in SHAPES.values())
    rows = max(num_images, 1)
    self.flat = torch.zeros((total,), device=self.device)
    # everything a DP step has to sum over the ranks sits in ONE buffer: weight gradients, GLO gradients, touched flags
    self.exchange = torch.zeros((total + rows * GLO_FEATURES + rows,), device=self.device)
    self.grad = self.exchange[:total]
    self.exp_avg, self.exp_avg_sq = torch.zeros_like(self.flat), torch.zeros_like(self.flat)
    self.w, self.g = {}, {}
    o = 0
    gen = torch.Generator(device="cpu").manual_seed(seed)
    for k in ops.COLOR_WEIGHTS:
      shape = SHAPES[k]
      cnt = math.prod(shape)
      self.w[k] = self.flat[o:o + cnt].view(shape)
      self.g[k] = self.grad[o:o + cnt].view(shape)
      if weights is not None:
        self.w[k].copy_(torch.as_tensor(weights[k], dtype=torch.float32))
      else:
        fan_in = SHAPES[k.replace("_b", "_w")][1]
        bound = 1.0 / math.sqrt(fan_in)
        self.w[k].copy_((torch.rand(shape, generator=gen) * 2 - 1) * bound)
      o += cnt
    self.model = ops.color_model(self.w)
    self.grads = ops.color_model(self.g)
    # GLOTable: zero-initialised embeddings (reference scene/color_model.py:16), one row per training image
    self.glo = torch.zeros((rows, GLO_FEATURES), device=self.device)
    self.glo_grad = self.exchange[total:total + rows * GLO_FEATURES].view(rows, GLO_FEATURES)
    self.touched = self.exchange[total + rows * GLO_FEATURES:]  # > 0: the row was used since the last step
    self.glo_m, self.glo_v = torch.zeros_like(self.glo), torch.zeros_like(self.glo)
    self.glo_steps = torch.zeros((rows,), device=self.device, dtype=torch.int32)
    self._zero_glo = torch.zeros((GLO_FEATURES,), device=self.device)
    self.lr, self.lr_glo, self.specular_reg = lr, lr_glo, specular_reg
    self.steps = 0

  def glo_row(self, image_idx):
    """the image's embedding (reference MLPScene.eval_colors, scene/mlp_scene.py:355-360: zeros for a view that is not a
    training image)."""
    return self._zero_glo if image_idx is None else self.glo[image_idx]

  def glo_grad_row(self, image_idx):
    if image_idx is None:
      return torch.zeros_like(self._zero_glo)
    self.touched[image_idx] = 1.0
    return self.glo_grad[image_idx]

  def forward(self, feature, position, cam_pos, image_idx, in_view, rgb, specular=None, specular_weight=1.0):
    return ops.color_mlp_fwd(self.model, feature, position, cam_pos, self.glo_row(image_idx), in_view=in_view,
                             specular_weight=specular_weight, rgb=rgb, specular=specular)

  def backward(self, feature, position, cam_pos, image_idx, row_w, d_rgb, d_feature, accumulate, reg_sums=None,
               specular_weight=1.0):
    ops.color_mlp_bwd(self.model, self.grads, feature, position, cam_pos, self.glo_row(image_idx), row_w, d_rgb, d_feature,
                      self.glo_grad_row(image_idx), specular_weight=specular_weight,
                      w_specular_reg=self.specular_reg if reg_sums is not None else 0.0, reg_sums=reg_sums,
                      accumulate=accumulate)

  def step(self):
    """Adam on the MLP weights (betas (0.9, 0.999), eps 1e-8) and on the GLO rows touched since the last step
    (betas (0.8, 0.95), SparseAdam semantics); clears gradients and flags.  Two launches, no host read-back."""
    self.steps += 1
    if self.flat.is_cuda:
      ops.adam_step(self.flat, self.grad, self.exp_avg, self.exp_avg_sq, self.lr, self.steps)
      ops.adam_rows(self.glo, self.glo_grad, self.glo_m, self.glo_v, self.glo_steps, self.touched, self.lr_glo)
      return
    # host-side replica (the gloo tests): the same update in torch ops
    def adam(p, g, m, v, lr, t, b1, b2):
      m.mul_(b1).add_(g, alpha=1 - b1)
      v.mul_(b2).addcmul_(g, g, value=1 - b2)
      p.addcdiv_(m, (v / (1 - b2 ** t)).sqrt_().add_(1e-8), value=-lr / (1 - b1 ** t))
      g.zero_()
    adam(self.flat, self.grad, self.exp_avg, self.exp_avg_sq, self.lr, self.steps, 0.9, 0.999)
    for r in (self.touched > 0).nonzero().flatten().tolist():
      self.glo_steps[r] += 1
      adam(self.glo[r], self.glo_grad[r], self.glo_m[r], self.glo_v[r], self.lr_glo, int(self.glo_steps[r]), 0.8, 0.95)
    self.touched.zero_()

  _STATE = ("flat", "exp_avg", "exp_avg_sq", "glo", "glo_m", "glo_v", "glo_steps")

  def state_dict(self):
    sd = {k: getattr(self, k).clone() for k in self._STATE}
    sd["steps"] = self.steps
    return sd

  def load_state_dict(self, sd):
    for k in self._STATE:
      getattr(self, k).copy_(sd[k])  # in place: the kernels hold pointers into `flat`
    self.steps = int(sd["steps"])

  def all_reduce(self, dist, group=None):
    """camera-batch DP: the colour model is replicated; its (tiny) dense gradients, the GLO gradients and the touched
    flags are summed over the ranks in one collective."""
    dist.all_reduce(self.exchange, group=group)
