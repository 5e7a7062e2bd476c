"""`taichi_splatting`-named operator shim over libgsb200: put the directory that holds this package
(`splat-trainer_b200/gsb200/compat`) and `splat-trainer_b200` on sys.path and the reference's scene / controller /
trainer modules import these names unchanged.  Everything computes in the sm_100a kernels; CPU tensors raise."""
from .data_types import Gaussians3D, RasterConfig, RenderedPoints, Rendering, TensorBag  # noqa: F401
from .renderer import project_to_image, render_gaussians, render_projected  # noqa: F401
from .spherical_harmonics import evaluate_sh_at  # noqa: F401
from .taichi_queue import TaichiQueue  # noqa: F401
from . import optim, perspective, renderer, torch_lib  # noqa: F401

__version__ = "0.31.0+gsb200"
