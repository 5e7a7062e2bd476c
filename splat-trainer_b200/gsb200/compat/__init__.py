"""Import shims that let the reference's unchanged Python run over libgsb200.

  compat/taichi_splatting   the operator interface (L1) the reference imports by that name, over the C-ABI
  compat/standins           minimal stand-ins for third-party packages this image lacks (tensordict, roma, ...)

`enable()` puts the operator shim FIRST on sys.path (it is the deliverable) and the stand-ins LAST (a real install of any
of those packages wins).  Nothing here is imported by the fused training path (gsb200.scene / gsb200.trainer)."""
import os
import sys

_HERE = os.path.dirname(os.path.abspath(__file__))


def enable(reference_root=None):
  """reference_root: directory holding the reference's `splat_trainer` package (appended to sys.path if given)."""
  if _HERE not in sys.path:
    sys.path.insert(0, _HERE)
  standins = os.path.join(_HERE, "standins")
  if standins not in sys.path:
    sys.path.append(standins)
  if reference_root and reference_root not in sys.path:
    sys.path.append(reference_root)
