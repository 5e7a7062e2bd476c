"""gsb200 — host side of the B200 Gaussian-splatting training hot path (C-ABI: include/gsb200.h)."""
from . import lib, ops  # noqa: F401
