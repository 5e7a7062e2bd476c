"""ctypes binding of libgsb200.so (C-ABI: include/gsb200.h).

PyTorch is plumbing only: tensors supply device pointers and the current CUDA stream.  There is
no CPU fallback and no other backend: if the shared library is missing, loading raises.
"""
import ctypes as C
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
# GSB200_LIB: another build of the same library (A/B measurements of kernel variants); default = the in-tree build
LIB_PATH = os.environ.get("GSB200_LIB") or os.path.join(os.path.dirname(_HERE), "libgsb200.so")

GS_GROUP_SCALAR, GS_GROUP_VECTOR, GS_GROUP_LOCAL_VECTOR = 0, 1, 2
GS_POST_NONE, GS_POST_NORMALIZE, GS_POST_CLAMP = 0, 1, 2


class Camera(C.Structure):  # gs_camera
  _fields_ = [("T", C.c_float * 16), ("fx", C.c_float), ("fy", C.c_float), ("cx", C.c_float), ("cy", C.c_float),
              ("width", C.c_int32), ("height", C.c_int32), ("near_plane", C.c_float), ("far_plane", C.c_float)]


class RasterConfig(C.Structure):  # gs_raster_config
  _fields_ = [("tile_size", C.c_int32), ("alpha_threshold", C.c_float), ("clamp_max_alpha", C.c_float),
              ("saturate_threshold", C.c_float), ("blur_cov", C.c_float), ("antialias", C.c_int32),
              ("compute_visibility", C.c_int32), ("compute_point_heuristic", C.c_int32),
              ("render_median_depth", C.c_int32), ("frustum_clamp", C.c_float)]


class OptimConfig(C.Structure):  # gs_optim_config
  _fields_ = [("mode", C.c_int32), ("beta1", C.c_float), ("beta2", C.c_float), ("eps", C.c_float),
              ("vis_beta", C.c_float), ("vis_smooth", C.c_float), ("grad_clip", C.c_float),
              ("bias_correction", C.c_int32)]


GS_EXCHANGE_MAX_GROUPS, GS_EXCHANGE_MAX_IMAGES = 8, 64


class ExchangeGroups(C.Structure):  # gs_exchange_groups
  _fields_ = [("count", C.c_int32), ("dim", C.c_int32 * GS_EXCHANGE_MAX_GROUPS),
              ("grad", C.c_void_p * GS_EXCHANGE_MAX_GROUPS)]


GS_OPTIM_MAX_GROUPS = 8


class OptimGroup(C.Structure):  # gs_optim_group
  _fields_ = [("kind", C.c_int32), ("dim", C.c_int32), ("lr", C.c_float), ("param", C.c_void_p), ("grad", C.c_void_p),
              ("exp_avg", C.c_void_p), ("exp_avg_sq", C.c_void_p), ("basis_ls", C.c_void_p), ("basis_q", C.c_void_p),
              ("post", C.c_int32), ("post_lo", C.c_float), ("post_hi", C.c_float)]


class ColorMLP(C.Structure):  # gs_color_mlp
  _fields_ = [("point_features", C.c_int32), ("glo_features", C.c_int32), ("hidden", C.c_int32), ("sh_degree", C.c_int32),
              ("weights", C.c_void_p * 10)]


class RegFuse(C.Structure):  # gs_reg_fuse
  _fields_ = [("w_scale", C.c_float), ("w_opacity", C.c_float), ("w_aspect", C.c_float), ("sums", C.c_void_p),
              ("depth", C.c_void_p)]


class RasterAux(C.Structure):  # gs_raster_aux
  _fields_ = [("records", C.c_void_p), ("tile_order", C.c_void_p), ("ticket", C.c_void_p), ("tile_work", C.c_void_p),
              ("counters", C.c_void_p)]


GS_REC_FLOATS = 12

_P, _I, _L, _F, _Z = C.c_void_p, C.c_int, C.c_int64, C.c_float, C.c_size_t
_CAM, _CFG, _OPT = C.POINTER(Camera), C.POINTER(RasterConfig), C.POINTER(OptimConfig)
_XG = C.POINTER(ExchangeGroups)
_REG = C.POINTER(RegFuse)
_AUX = C.POINTER(RasterAux)

# name -> (restype, argtypes); one entry per symbol declared in include/gsb200.h
SIGNATURES = {
  "gs_default_raster_config": (None, [_CFG]),
  "gs_version": (_I, []),
  "gs_last_error": (C.c_char_p, []),
  "gs_launch_count": (C.c_ulonglong, []),
  "gs_project_fwd": (_I, [_CAM, _CFG, _L, _P, _P, _P, _P, _P, _I, _P, _P, _P, _P, _I, _P, _P, _P]),
  "gs_project_bwd": (_I, [_CAM, _CFG, _L, _P, _P, _P, _P, _P, _I, _P, _P, _I, _P, _I, _P, _P, _I, _REG,
                          _P, _P, _P, _P, _P, _I, _P]),
  "gs_sh_eval_fwd": (_I, [_L, _P, _P, _I, _P, C.POINTER(C.c_float), _P, _P]),
  "gs_sh_eval_bwd": (_I, [_L, _P, _P, _I, _P, C.POINTER(C.c_float), _P, _P, _P, _P]),
  "gs_tile_info_bytes": (_Z, [_L, _I, _I]),
  "gs_tile_map_workspace_bytes": (_Z, [_L, _L, _I, _I]),
  "gs_tile_map": (_I, [_CFG, _L, _P, _P, _P, _I, _I, _L, _P, _Z, _P, _P, _P, _P]),
  "gs_tile_count": (_I, [_CFG, _L, _P, _P, _I, _I, _P, _P, _P]),
  "gs_radix_sort_workspace_bytes": (_Z, [_L]),
  "gs_radix_sort_pairs": (_I, [_P, _P, _P, _P, _L, _P, _I, _I, _P, _Z, _P]),
  "gs_raster_fwd": (_I, [_CFG, _I, _L, _P, _P, _P, _P, _P, _I, _I, _P, _P, _P, _P, _P, _AUX, _P]),
  "gs_raster_records": (_I, [_CFG, _L, _P, _P, _P, _P]),
  "gs_tile_order": (_I, [_I, _P, _I, _P, _P, _P]),
  "gs_raster_acc_stride": (_I, [_I]),
  "gs_raster_bwd": (_I, [_CFG, _I, _L, _P, _P, _P, _P, _I, _I, _P, _P, _P, _P, _I, _P, _AUX, _P]),
  "gs_loss_l1_mse": (_I, [_L, _P, _P, _P, _F, _F, _P, _P, _P]),
  "gs_optim_begin_workspace_bytes": (_Z, [_L]),
  "gs_optim_begin": (_I, [_OPT, _L, _P, _P, _P, _P, _P, _P, _Z, _P]),
  "gs_optim_vis_update": (_I, [_OPT, _L, _P, _P, _P, _P]),
  "gs_visible_indexes_workspace_bytes": (_Z, [_L]),
  "gs_visible_indexes": (_I, [_L, _P, _P, _P, _P, _Z, _P]),
  "gs_visible_indexes_static_max_rows": (_L, []),
  "gs_visible_indexes_static": (_I, [_L, _P, _P, _P, _P]),
  "gs_color_mlp_fwd": (_I, [C.POINTER(ColorMLP), _L, _P, _I, _P, _P, C.POINTER(C.c_float), _P, _F, _P, _P, _P]),
  "gs_color_mlp_bwd": (_I, [C.POINTER(ColorMLP), C.POINTER(ColorMLP), _L, _P, _I, _P, _P, C.POINTER(C.c_float), _P, _F, _P, _I,
                             _F, _P, _P, _P, _I, _P]),
  "gs_kmeans_workspace_bytes": (_Z, [_I, _I]),
  "gs_assign_clusters": (_I, [_L, _I, _I, _P, _P, _P, _P]),
  "gs_kmeans": (_I, [_L, _I, _I, _I, _P, _P, _P, _P, _Z, _P]),
  "gs_view_features": (_I, [_L, _P, _P, _L, _P, _F, _I, _P, _P]),
  "gs_point_visibility": (_I, [_L, _P, _I, _P, _P, _P]),
  "gs_knn_scale_workspace_bytes": (_Z, [_L]),
  "gs_knn_scale": (_I, [_L, _P, _I, _P, _P, _Z, _P]),
  "gs_adam_rows": (_I, [_I, _I, _P, _P, _P, _P, _P, _P, _F, _F, _F, _F, _P]),
  "gs_adam_step": (_I, [_L, _P, _P, _P, _P, _F, _F, _F, _F, _I, _I, _P]),
  "gs_optim_step_groups": (_I, [_OPT, _L, _P, _P, _I, C.POINTER(OptimGroup), _P, _P, _P, _I, _P]),
  "gs_optim_step": (_I, [_OPT, _I, _L, _P, _P, _I, _F, _P, _P, _P, _P, _P, _P, _P, _P, _P, _I, _F, _F, _I, _P]),
  "gs_accumulate_stats": (_I, [_L, _P, _P, _P, _I, _P, _I, _F, _F, _P, _P, _P, _P, _P, _P, _P, _I, _P, _P]),
  "gs_select_topk_workspace_bytes": (_Z, [_L]),
  "gs_select_topk_mask": (_I, [_P, _L, _L, _I, _P, _P, _Z, _P]),
  "gs_compact_workspace_bytes": (_Z, [_L]),
  "gs_compact_slots": (_I, [_P, _L, _P, _P, _P, _Z, _P]),
  "gs_compact_rows": (_I, [_P, _P, _L, _I, _P, _P]),
  "gs_split_positions": (_I, [_L, _P, _P, _P, _P, _P, _F, _P, _P, _L, _P]),
  "gs_gather_rows": (_I, [_P, _P, _L, _P, _I, _I, _L, _P]),
  "gs_exchange_record_bytes": (_Z, [_L, _L, _I]),
  "gs_exchange_pack_workspace_bytes": (_Z, [_L]),
  "gs_exchange_pack": (_I, [_L, _P, _P, _I, _P, _I, _P, _P, _I, _P, _XG, _P, _L, _P, _Z, _P]),
  "gs_ssim_workspace_bytes": (_Z, [_I, _I, _I, _I]),
  "gs_ssim_loss": (_I, [_I, _I, _I, _I, _P, _P, _P, _F, _P, _P, _P, _Z, _P]),
  "gs_photometric_loss": (_I, [_I, _I, _I, _I, _P, _P, _P, _F, _F, _F, _P, _P, _P, _P, _Z, _P]),
  "gs_mcmc_noise": (_I, [_L, _P, _P, _P, _P, _I, _F, _F, _P, _P, _P]),
  "gs_reg_loss": (_I, [_L, _F, _F, _F, _P, _P, _P, _P, _I, _P, _I, _P, _I, _P, _P, _P]),
  "gs_exchange_fold": (_I, [_L, _I, C.POINTER(C.c_void_p), _L, _XG, _F, _F, _P, _P, _P, _P, _L, _L, _L, _L, _P, _P]),
  "gs_exchange_publish": (_I, [_P, _P, _L, _L, _XG, _P, _I, C.POINTER(C.c_void_p), _P]),
}

_lib = None


def load(path=None):
  """Load libgsb200.so and bind every symbol; raises if the library or a symbol is missing."""
  global _lib
  if _lib is not None and path is None:
    return _lib
  path = path or LIB_PATH
  if not os.path.exists(path):
    raise RuntimeError(f"libgsb200.so not found at {path}: build it with `make -C splat-trainer_b200/csrc` "
                       "(there is no CPU fallback)")
  lib = C.CDLL(path)
  for name, (res, args) in SIGNATURES.items():
    fn = getattr(lib, name)  # AttributeError if the library does not export it
    fn.restype = res
    fn.argtypes = args
  _lib = lib
  return lib


class GsError(RuntimeError):
  pass


def check(rc):
  if rc != 0:
    msg = load().gs_last_error()
    raise GsError(f"libgsb200 status {rc}: {msg.decode() if msg else ''}")


def stream():
  return torch.cuda.current_stream().cuda_stream


def ptr(t):
  """device pointer of a tensor (None -> NULL); the tensor must be CUDA and contiguous."""
  if t is None:
    return None
  if not t.is_cuda:
    raise GsError("libgsb200 operates on CUDA tensors only (no CPU fallback)")
  if not t.is_contiguous():
    raise GsError("tensor must be contiguous")
  return t.data_ptr()


def default_raster_config(**kw):
  cfg = RasterConfig()
  load().gs_default_raster_config(C.byref(cfg))
  for k, v in kw.items():
    if not hasattr(cfg, k):
      raise TypeError(f"RasterConfig has no field {k}")
    setattr(cfg, k, v)
  return cfg


def make_camera(T_camera_world, fx, fy, cx, cy, width, height, near=0.1, far=100.0):
  cam = Camera()
  flat = [float(x) for x in torch.as_tensor(T_camera_world, dtype=torch.float32).reshape(-1).tolist()]
  assert len(flat) == 16
  for i in range(16):
    cam.T[i] = flat[i]
  cam.fx, cam.fy, cam.cx, cam.cy = float(fx), float(fy), float(cx), float(cy)
  cam.width, cam.height = int(width), int(height)
  cam.near_plane, cam.far_plane = float(near), float(far)
  return cam


def make_optim_config(mode=1, beta1=0.8, beta2=0.95, eps=1e-8, vis_beta=0.999, vis_smooth=0.01, grad_clip=2.0,
                      bias_correction=1):
  """defaults = reference config/scene/mlp.yaml:25-31 (LaProp, betas (.8,.95), vis_beta .999, vis_smooth .01, clip 2)."""
  return OptimConfig(mode=mode, beta1=beta1, beta2=beta2, eps=eps, vis_beta=vis_beta, vis_smooth=vis_smooth,
                     grad_clip=grad_clip, bias_correction=bias_correction)
