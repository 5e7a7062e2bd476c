"""fused_ssim stand-in (reference trainer/trainer.py:17,112,450-462): the published fused-ssim function — 11x11 Gaussian
window (sigma 1.5), C1 = 0.01², C2 = 0.03², mean over the map — in plain torch with autograd.  `padding="valid"` is what
the reference passes.  The training hot path of gsb200 does not use this: its loss kernels are csrc/ssim.cu."""
import torch
import torch.nn.functional as F


def _window(device, dtype):
  x = torch.arange(11, device=device, dtype=dtype) - 5
  g = torch.exp(-(x * x) / (2 * 1.5 * 1.5))
  return g / g.sum()


def fused_ssim(img1, img2, padding="same", train=True):
  """img1, img2: (B, C, H, W) in [0,1] -> scalar mean SSIM."""
  C = img1.shape[1]
  g = _window(img1.device, img1.dtype)
  pad = 5 if padding == "same" else 0

  def blur(x):
    x = F.conv2d(x, g.view(1, 1, 1, 11).expand(C, 1, 1, 11), padding=(0, pad), groups=C)
    return F.conv2d(x, g.view(1, 1, 11, 1).expand(C, 1, 11, 1), padding=(pad, 0), groups=C)

  mu1, mu2 = blur(img1), blur(img2)
  s11 = blur(img1 * img1) - mu1 * mu1
  s22 = blur(img2 * img2) - mu2 * mu2
  s12 = blur(img1 * img2) - mu1 * mu2
  C1, C2 = 0.01 ** 2, 0.03 ** 2
  m = ((2 * mu1 * mu2 + C1) * (2 * s12 + C2)) / ((mu1 * mu1 + mu2 * mu2 + C1) * (s11 + s22 + C2))
  return m.mean()
