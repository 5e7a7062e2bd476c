"""evaluate_sh_at (reference scene/transfer_sh.py:49): colour of the indexed points seen from cam_pos."""
import torch

from gsb200 import ops


class _EvalSH(torch.autograd.Function):
  @staticmethod
  def forward(ctx, sh, position, indexes, cam_pos):
    sh, position = sh.contiguous(), position.contiguous()
    k = sh.shape[2]
    deg = int(round(k ** 0.5)) - 1
    assert (deg + 1) ** 2 == k and deg <= 3, f"SH with {k} coefficients per channel"
    ctx.deg, ctx.cam_pos = deg, cam_pos.detach().cpu()
    ctx.save_for_backward(sh, position, indexes)
    return ops.sh_eval_fwd(indexes, sh, deg, position, ctx.cam_pos)

  @staticmethod
  def backward(ctx, d_rgb):
    sh, position, indexes = ctx.saved_tensors
    g_sh, g_pos = ops.sh_eval_bwd(indexes, sh, ctx.deg, position, ctx.cam_pos, d_rgb)
    return g_sh, g_pos, None, None


def evaluate_sh_at(sh_params, positions, indexes, camera_pos):
  """sh_params (N,3,K), positions (N,3), indexes (V,) int64, camera_pos (3,) -> (V,3) = 0.5 + Σ_k Y_k(dir)·sh."""
  return _EvalSH.apply(sh_params, positions, indexes.to(torch.int64), camera_pos)
