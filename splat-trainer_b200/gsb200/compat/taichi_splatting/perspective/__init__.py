from .params import CameraParams  # noqa: F401
