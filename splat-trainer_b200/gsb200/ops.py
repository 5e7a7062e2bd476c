"""Tensor-level wrappers over the C-ABI, one per entry point of include/gsb200.h.

Each function enqueues work on torch's current CUDA stream and returns torch tensors that own the
outputs.  Nothing here computes on the host; nothing falls back to torch ops.
"""
import ctypes as C

import torch

from . import lib as L


def _f32(t):
  assert t.dtype == torch.float32, f"expected float32, got {t.dtype}"
  return t


def tiles_of(width, height):
  return (width + 15) // 16, (height + 15) // 16


# ---- (1) projection + SH ---------------------------------------------------------------------------
def project_fwd(cam, cfg, position, log_scaling, rotation, alpha_logit, sh=None, sh_degree=0, acc=None,
                out=None, tile_info=None, records=None):
  """gs_project_fwd.  Returns (g2d (n,6), depth (n), rgb (n,3)); row i belongs to Gaussian i.
  tile_info: a TileInfo (filled for gs_tile_map(tile_info=...))."""
  n = position.shape[0]
  dev = position.device
  if out is None:
    g2d = torch.empty((n, 6), device=dev, dtype=torch.float32)
    depth = torch.empty((n,), device=dev, dtype=torch.float32)
    rgb = torch.empty((n, 3), device=dev, dtype=torch.float32)
  else:
    g2d, depth, rgb = out
  L.check(L.load().gs_project_fwd(C.byref(cam), C.byref(cfg), n, L.ptr(_f32(position)), L.ptr(_f32(log_scaling)),
                                  L.ptr(_f32(rotation)), L.ptr(_f32(alpha_logit)), L.ptr(sh), sh_degree,
                                  L.ptr(g2d), L.ptr(depth), L.ptr(rgb), L.ptr(acc),
                                  0 if acc is None else acc.shape[1], None if tile_info is None else L.ptr(tile_info.buffer),
                                  L.ptr(records), L.stream()))
  return g2d, depth, rgb


class TileInfo:
  """Opaque per-row tile-footprint buffer written by gs_project_fwd and consumed by gs_tile_map."""

  def __init__(self, n, width, height, device, buffer=None):
    """buffer: an existing uint8 allocation of at least gs_tile_info_bytes(n, W, H) bytes to lay the info out in."""
    self.n, self.width, self.height = n, width, height
    nbytes = L.load().gs_tile_info_bytes(n, width, height)
    if buffer is not None and buffer.numel() < nbytes:
      raise L.GsError("tile info buffer too small")
    self.buffer = buffer if buffer is not None else torch.empty((nbytes,), device=device, dtype=torch.uint8)

  @property
  def counts(self):
    """per-row tile counts (int32 view)"""
    off = (max(self.n, 1) * 16 + 255) // 256 * 256
    return self.buffer[off:off + 4 * self.n].view(torch.int32)


def project_bwd(cam, cfg, position, log_scaling, rotation, alpha_logit, sh, sh_degree, g2d, d_g2d, d_rgb,
                d_depth=None, row_vis=None, grads=None, accumulate=False, reg=None):
  """gs_project_bwd.  d_g2d / d_rgb / row_vis may be column views of one (n, stride) accumulator.
  grads: dict(position, log_scaling, rotation, alpha_logit, sh) to write/accumulate into (allocated if None).
  reg: (weights (scale, opacity, aspect), sums (4,), depth (n,)) — fuse the regularisation gradients of reg_loss."""
  n = position.shape[0]
  dev = position.device
  if grads is None:
    grads = dict(position=torch.zeros((n, 3), device=dev), log_scaling=torch.zeros((n, 3), device=dev),
                 rotation=torch.zeros((n, 4), device=dev), alpha_logit=torch.zeros((n,), device=dev),
                 sh=None if sh is None else torch.zeros_like(sh))

  def strided(t):
    if t is None:
      return None, 0
    assert t.dtype == torch.float32 and t.is_cuda
    if t.dim() == 1:
      return t.data_ptr(), t.stride(0)
    assert t.stride(1) == 1
    return t.data_ptr(), t.stride(0)

  p_dg, s_dg = strided(d_g2d)
  p_dr, s_dr = strided(d_rgb)
  p_rv, s_rv = strided(row_vis)
  fuse = None
  if reg is not None:
    wts, sums, depth = reg
    fuse = C.byref(L.RegFuse(float(wts[0]), float(wts[1]), float(wts[2]), L.ptr(sums), L.ptr(_f32(depth))))
  L.check(L.load().gs_project_bwd(C.byref(cam), C.byref(cfg), n, L.ptr(position), L.ptr(log_scaling), L.ptr(rotation),
                                  L.ptr(alpha_logit), L.ptr(sh), sh_degree, L.ptr(g2d), p_dg, s_dg, p_dr, s_dr,
                                  L.ptr(d_depth), p_rv, s_rv, fuse, L.ptr(grads["position"]), L.ptr(grads["log_scaling"]),
                                  L.ptr(grads["rotation"]), L.ptr(grads["alpha_logit"]), L.ptr(grads.get("sh")),
                                  int(accumulate), L.stream()))
  return grads


def _cam_pos3(cam_pos):
  v = [float(x) for x in torch.as_tensor(cam_pos).reshape(-1).tolist()]
  assert len(v) == 3
  return (C.c_float * 3)(*v)


def sh_eval_fwd(indexes, sh, sh_degree, position, cam_pos):
  v = position.shape[0] if indexes is None else indexes.shape[0]
  rgb = torch.empty((v, 3), device=position.device, dtype=torch.float32)
  if indexes is not None:
    assert indexes.dtype == torch.int64
  L.check(L.load().gs_sh_eval_fwd(v, L.ptr(indexes), L.ptr(_f32(sh)), sh_degree, L.ptr(_f32(position)),
                                  _cam_pos3(cam_pos), L.ptr(rgb), L.stream()))
  return rgb


def sh_eval_bwd(indexes, sh, sh_degree, position, cam_pos, d_rgb):
  v = position.shape[0] if indexes is None else indexes.shape[0]
  g_sh = torch.zeros_like(sh)
  g_pos = torch.zeros_like(position)
  L.check(L.load().gs_sh_eval_bwd(v, L.ptr(indexes), L.ptr(sh), sh_degree, L.ptr(position), _cam_pos3(cam_pos),
                                  L.ptr(_f32(d_rgb.contiguous())), L.ptr(g_sh), L.ptr(g_pos), L.stream()))
  return g_sh, g_pos


# ---- (2) tile mapping --------------------------------------------------------------------------------
def tile_count(cfg, g2d, depth, width, height):
  m = g2d.shape[0]
  counts = torch.empty((max(m, 1),), device=g2d.device, dtype=torch.int32)
  counters = torch.empty((4,), device=g2d.device, dtype=torch.int32)
  L.check(L.load().gs_tile_count(C.byref(cfg), m, L.ptr(g2d), L.ptr(depth), width, height, L.ptr(counts),
                                 L.ptr(counters), L.stream()))
  return counts[:m], counters


class TileMapBuffers:
  """Caller-owned buffers of gs_tile_map for (m rows, max_overlaps, image size); reusable across calls."""

  def __init__(self, m, max_overlaps, width, height, device):
    self.m, self.max_overlaps, self.width, self.height = m, int(max_overlaps), width, height
    tx, ty = tiles_of(width, height)
    nbytes = L.load().gs_tile_map_workspace_bytes(m, self.max_overlaps, width, height)
    self.workspace = torch.empty((nbytes,), device=device, dtype=torch.uint8)
    self.overlap_to_point = torch.empty((max(self.max_overlaps, 1),), device=device, dtype=torch.int32)
    self.tile_ranges = torch.empty((tx * ty, 2), device=device, dtype=torch.int32)
    self.counters = torch.zeros((4,), device=device, dtype=torch.int32)


def tile_map(cfg, g2d, depth, width, height, max_overlaps=None, buffers=None, tile_info=None):
  """gs_tile_map.  With max_overlaps=None the exact total is read back first (one host sync, like the
  reference's mapper); the training path passes reusable `buffers` and never syncs.
  Returns (overlap_to_point[max_overlaps], tile_ranges (T,2), counters[4], buffers)."""
  m = g2d.shape[0]
  if buffers is None:
    if max_overlaps is None:
      _, counters = tile_count(cfg, g2d, depth, width, height)
      max_overlaps = int(counters[0].item())
    buffers = TileMapBuffers(m, max_overlaps, width, height, g2d.device)
  b = buffers
  assert b.m >= m and b.width == width and b.height == height
  if tile_info is not None:
    assert tile_info.n == m and tile_info.width == width and tile_info.height == height
  L.check(L.load().gs_tile_map(C.byref(cfg), m, L.ptr(g2d), L.ptr(depth),
                               None if tile_info is None else L.ptr(tile_info.buffer), width, height, b.max_overlaps,
                               L.ptr(b.workspace), b.workspace.numel(), L.ptr(b.overlap_to_point),
                               L.ptr(b.tile_ranges), L.ptr(b.counters), L.stream()))
  return b.overlap_to_point, b.tile_ranges, b.counters, b


def radix_sort_pairs(keys, vals, begin_bit=0, end_bit=32, n_dev=None):
  n = keys.shape[0]
  assert keys.dtype == torch.int32 and vals.dtype == torch.int32  # bit patterns of u32
  kin, vin = keys.clone(), vals.clone()
  kout, vout = torch.empty_like(keys), torch.empty_like(vals)
  nbytes = L.load().gs_radix_sort_workspace_bytes(n)
  ws = torch.empty((nbytes,), device=keys.device, dtype=torch.uint8)
  L.check(L.load().gs_radix_sort_pairs(L.ptr(kin), L.ptr(vin), L.ptr(kout), L.ptr(vout), n, L.ptr(n_dev), begin_bit,
                                       end_bit, L.ptr(ws), nbytes, L.stream()))
  return kout, vout


# ---- (3)/(4) rasterisation -------------------------------------------------------------------------
def acc_stride(channels):
  return L.load().gs_raster_acc_stride(channels)


class RasterAux:
  """fast-path buffers of the rasteriser for one image size (gs_raster_aux): staged splat records, the heaviest-first
  tile queue of the forward and of the backward, the per-tile work the forward measures.
  persistent=False (default): one CTA per tile in tile order — measured faster on the BASELINE scenes than the
  heaviest-first queue (neighbouring tiles run together and share their splat records in L2; profiles/README.md r2a)."""

  def __init__(self, m, width, height, device, persistent=False):
    tx, ty = tiles_of(width, height)
    self.m, self.tiles = m, tx * ty
    self.records = torch.empty((m, L.GS_REC_FLOATS), device=device, dtype=torch.float32)
    self.tile_work = torch.zeros((self.tiles,), device=device, dtype=torch.int32)
    self.order = torch.empty((2, self.tiles), device=device, dtype=torch.int32) if persistent else None
    self.ticket = torch.zeros((2,), device=device, dtype=torch.int32) if persistent else None
    self.counters = None

  def struct(self, which, counters=None):
    """which: 0 = forward (queue sorted by list length), 1 = backward (queue sorted by the forward's tile_work)."""
    a = L.RasterAux()
    a.records = L.ptr(self.records)
    a.tile_work = L.ptr(self.tile_work)
    if self.order is not None:
      a.tile_order = L.ptr(self.order[which])
      a.ticket = self.ticket[which:].data_ptr()
    a.counters = L.ptr(counters)
    return a

  def order_forward(self, tile_ranges):
    if self.order is not None:
      L.check(L.load().gs_tile_order(self.tiles, L.ptr(tile_ranges), 2, L.ptr(self.order[0]), self.ticket.data_ptr(),
                                     L.stream()))

  def order_backward(self):
    if self.order is not None:
      L.check(L.load().gs_tile_order(self.tiles, L.ptr(self.tile_work), 1, L.ptr(self.order[1]),
                                     self.ticket[1:].data_ptr(), L.stream()))


def raster_records(cfg, g2d, features, records):
  """gs_raster_records: staged 48-byte records of every in-view row (3 feature channels)."""
  L.check(L.load().gs_raster_records(C.byref(cfg), g2d.shape[0], L.ptr(g2d), L.ptr(_f32(features)), L.ptr(records), L.stream()))
  return records


def raster_fwd(cfg, g2d, features, depth, overlap_to_point, tile_ranges, width, height, visibility=False,
               median=False, out=None, aux=None, counters=None):
  """aux: a RasterAux whose records are current for (g2d, features) -> record-staged asynchronous kernels."""
  m, ch = g2d.shape[0], features.shape[1]
  dev = g2d.device
  if out is None:
    image = torch.empty((height, width, ch), device=dev, dtype=torch.float32)
    final_T = torch.empty((height, width), device=dev, dtype=torch.float32)
    n_contrib = torch.empty((height, width), device=dev, dtype=torch.int32)
  else:
    image, final_T, n_contrib = out
  if visibility is True:
    vis = torch.zeros((m,), device=dev, dtype=torch.float32)
  elif visibility is False or visibility is None:
    vis = None
  else:
    vis = visibility  # caller's accumulator (+=)
  med = torch.empty((height, width), device=dev, dtype=torch.float32) if median else None
  ax = None
  if aux is not None:
    aux.order_forward(tile_ranges)
    ax = C.byref(aux.struct(0, counters))
  L.check(L.load().gs_raster_fwd(C.byref(cfg), ch, m, L.ptr(g2d), L.ptr(_f32(features)), L.ptr(depth),
                                 L.ptr(overlap_to_point), L.ptr(tile_ranges), width, height, L.ptr(image),
                                 L.ptr(final_T), L.ptr(n_contrib), L.ptr(vis), L.ptr(med), ax, L.stream()))
  return image, final_T, n_contrib, vis, med


def raster_bwd(cfg, g2d, features, overlap_to_point, tile_ranges, width, height, final_T, n_contrib, d_image,
               acc=None, vis_plane=None, aux=None):
  """gs_raster_bwd; acc (m, acc_stride) must be zero on entry (allocated zeroed if None).
  vis_plane (m,): receives the visibility sums instead of acc[:, 8+C] (must be zero on entry too)."""
  m, ch = g2d.shape[0], features.shape[1]
  if acc is None:
    acc = torch.zeros((m, acc_stride(ch)), device=g2d.device, dtype=torch.float32)
  ax = None
  if aux is not None:
    aux.order_backward()
    ax = C.byref(aux.struct(1))
  L.check(L.load().gs_raster_bwd(C.byref(cfg), ch, m, L.ptr(g2d), L.ptr(features), L.ptr(overlap_to_point),
                                 L.ptr(tile_ranges), width, height, L.ptr(final_T), L.ptr(n_contrib),
                                 L.ptr(_f32(d_image)), L.ptr(acc), acc.shape[1], L.ptr(vis_plane), ax, L.stream()))
  return acc


def loss_l1_mse(pred, target, w_l1, w_mse, d_pred=None, sums=None):
  """target: float32 in [0,1] or uint8.  Returns (d_pred, sums[2] = (Σ|c−t|, Σ(c−t)²))."""
  n = pred.numel()
  assert target.numel() == n
  if d_pred is None:
    d_pred = torch.empty_like(pred)
  if sums is None:
    sums = torch.zeros((2,), device=pred.device, dtype=torch.float32)
  tf = target if target.dtype == torch.float32 else None
  tu = target if target.dtype == torch.uint8 else None
  assert tf is not None or tu is not None
  L.check(L.load().gs_loss_l1_mse(n, L.ptr(_f32(pred)), L.ptr(tf), L.ptr(tu), w_l1, w_mse, L.ptr(d_pred), L.ptr(sums),
                                  L.stream()))
  return d_pred, sums


# ---- (5) optimiser / statistics / densification ----------------------------------------------------
def optim_vis_update(ocfg, visible, running_vis, step):
  L.check(L.load().gs_optim_vis_update(C.byref(ocfg), visible.shape[0], L.ptr(_f32(visible)), L.ptr(running_vis),
                                       L.ptr(step), L.stream()))


class VisibleIndexes:
  """Device-side `visible.nonzero()`: indexes (n,) int32 (first count[0] entries valid) + reusable workspace."""

  def __init__(self, n, device):
    self.n = n
    self.indexes = torch.empty((max(n, 1),), device=device, dtype=torch.int32)
    self.count = torch.zeros((1,), device=device, dtype=torch.int32)
    self.nbytes = L.load().gs_visible_indexes_workspace_bytes(n)
    self.workspace = torch.empty((self.nbytes,), device=device, dtype=torch.uint8)

  def begin(self, ocfg, visible, running_vis, step):
    """gs_optim_begin: visibility state update + index list in one kernel."""
    assert visible.shape[0] == self.n
    if not hasattr(self, "begin_ws"):
      self.begin_bytes = L.load().gs_optim_begin_workspace_bytes(self.n)
      self.begin_ws = torch.empty((self.begin_bytes,), device=visible.device, dtype=torch.uint8)
    L.check(L.load().gs_optim_begin(C.byref(ocfg), self.n, L.ptr(_f32(visible)), L.ptr(running_vis), L.ptr(step),
                                    L.ptr(self.indexes), L.ptr(self.count), L.ptr(self.begin_ws), self.begin_bytes,
                                    L.stream()))
    return self

  def update(self, visible):
    assert visible.shape[0] == self.n
    L.check(L.load().gs_visible_indexes(self.n, L.ptr(_f32(visible)), L.ptr(self.indexes), L.ptr(self.count),
                                        L.ptr(self.workspace), self.nbytes, L.stream()))
    return self


def optim_step(ocfg, kind, lr, param, grad, exp_avg, exp_avg_sq, visible, running_vis, step, basis_ls=None,
               basis_q=None, post=L.GS_POST_NONE, post_lo=0.0, post_hi=0.0, zero_grad=True, indexes=None):
  """indexes: a VisibleIndexes (updated from the same `visible`) -> only those rows are touched; None = dense."""
  n = param.shape[0]
  dim = param.numel() // max(n, 1) if n > 0 else 1
  L.check(L.load().gs_optim_step(C.byref(ocfg), kind, n, None if indexes is None else L.ptr(indexes.indexes),
                                 None if indexes is None else L.ptr(indexes.count), dim, lr, L.ptr(_f32(param)), L.ptr(_f32(grad)), L.ptr(exp_avg),
                                 L.ptr(exp_avg_sq), L.ptr(visible), L.ptr(running_vis), L.ptr(step), L.ptr(basis_ls),
                                 L.ptr(basis_q), post, post_lo, post_hi, int(zero_grad), L.stream()))


def optim_group(kind, lr, param, grad, exp_avg, exp_avg_sq, basis_ls=None, basis_q=None, post=L.GS_POST_NONE, post_lo=0.0,
                post_hi=0.0):
  """one gs_optim_group entry for optim_step_groups."""
  n = param.shape[0]
  dim = param.numel() // max(n, 1) if n > 0 else 1
  return L.OptimGroup(kind, dim, lr, L.ptr(_f32(param)), L.ptr(_f32(grad)), L.ptr(exp_avg), L.ptr(exp_avg_sq),
                      L.ptr(basis_ls), L.ptr(basis_q), post, post_lo, post_hi)


def optim_step_groups(ocfg, n, groups, visible, running_vis, step, zero_grad=True, indexes=None):
  """gs_optim_step_groups: the listed groups (optim_group entries) in one launch; bit-identical to one optim_step per
  group as long as no group of the call updates another one's basis (the library refuses that)."""
  arr = (L.OptimGroup * len(groups))(*groups)
  L.check(L.load().gs_optim_step_groups(C.byref(ocfg), n, None if indexes is None else L.ptr(indexes.indexes),
                                        None if indexes is None else L.ptr(indexes.count), len(groups), arr, L.ptr(visible),
                                        L.ptr(running_vis), L.ptr(step), int(zero_grad), L.stream()))


def accumulate_stats(idx, g2d, vis, heur, split_alpha, prune_alpha, state, scene_visible=None, clear_acc=None,
                     clear_vis=None):
  """vis: (m,) or a column view; heur: (m,2)-like view whose row stride is heur.stride(0) with
  [.,0]=prune_cost, [.,1]=split_score.  state: dict of the PointState tensors (updated in place)."""
  m = g2d.shape[0]
  assert state["points_in_view"].dtype == torch.int16
  L.check(L.load().gs_accumulate_stats(m, L.ptr(idx), L.ptr(g2d), vis.data_ptr(), vis.stride(0), heur.data_ptr(),
                                       heur.stride(0), split_alpha, prune_alpha, L.ptr(state["prune_cost"]),
                                       L.ptr(state["split_score"]), L.ptr(state["max_scale_px"]),
                                       L.ptr(state["points_in_view"]), L.ptr(state["visibility"]),
                                       L.ptr(scene_visible), L.ptr(clear_acc),
                                       0 if clear_acc is None else clear_acc.shape[1], L.ptr(clear_vis), L.stream()))


def select_topk_mask(values, k, descending):
  """take_n: bool mask with exactly min(k,n) True entries (k smallest, or largest if descending)."""
  n = values.shape[0]
  mask = torch.zeros((n,), device=values.device, dtype=torch.uint8)
  nbytes = L.load().gs_select_topk_workspace_bytes(n)
  ws = torch.empty((nbytes,), device=values.device, dtype=torch.uint8)
  L.check(L.load().gs_select_topk_mask(L.ptr(_f32(values)), n, int(k), int(descending), L.ptr(mask), L.ptr(ws), nbytes,
                                       L.stream()))
  return mask.view(torch.bool) if hasattr(mask, "view") else mask.bool()


def compact_slots(keep_mask):
  n = keep_mask.shape[0]
  km = keep_mask.view(torch.uint8) if keep_mask.dtype == torch.bool else keep_mask
  assert km.dtype == torch.uint8
  slots = torch.empty((max(n, 1),), device=km.device, dtype=torch.int32)
  count = torch.zeros((1,), device=km.device, dtype=torch.int32)
  nbytes = L.load().gs_compact_workspace_bytes(n)
  ws = torch.empty((nbytes,), device=km.device, dtype=torch.uint8)
  L.check(L.load().gs_compact_slots(L.ptr(km), n, L.ptr(slots), L.ptr(count), L.ptr(ws), nbytes, L.stream()))
  return slots[:n], count


def compact_rows(src, slots, dst):
  n = src.shape[0]
  row_bytes = src.element_size() * (src.numel() // max(n, 1)) if n > 0 else src.element_size()
  L.check(L.load().gs_compact_rows(L.ptr(src), L.ptr(dst), n, row_bytes, L.ptr(slots), L.stream()))
  return dst


def split_positions(split_idx, position, log_scaling, rotation, uniform, sep, out_position, out_log_scaling,
                    out_offset=0):
  s = split_idx.shape[0]
  assert split_idx.dtype == torch.int64
  L.check(L.load().gs_split_positions(s, L.ptr(split_idx), L.ptr(position), L.ptr(log_scaling), L.ptr(rotation),
                                      L.ptr(uniform), sep, L.ptr(out_position), L.ptr(out_log_scaling), out_offset,
                                      L.stream()))


def gather_rows(src, dst, idx, repeat=1, out_offset=0):
  s = idx.shape[0]
  n = src.shape[0]
  row_bytes = src.element_size() * (src.numel() // max(n, 1)) if n > 0 else src.element_size()
  L.check(L.load().gs_gather_rows(L.ptr(src), L.ptr(dst), s, L.ptr(idx), row_bytes, repeat, out_offset, L.stream()))
  return dst


# ---- camera-batch data parallelism: per-image exchange records (include/gsb200.h section 6) ---------------------
def exchange_groups(grad_planes):
  """grad_planes: list of contiguous (n, ...) float32 CUDA tensors (one per parameter group, record-plane order)."""
  g = L.ExchangeGroups()
  assert 1 <= len(grad_planes) <= L.GS_EXCHANGE_MAX_GROUPS
  g.count = len(grad_planes)
  total = 0
  for k, t in enumerate(grad_planes):
    dim = int(t.numel() // t.shape[0])
    g.dim[k] = dim
    g.grad[k] = L.ptr(t)
    total += dim
  return g, total


def exchange_record_bytes(n, cap, row_floats):
  return int(L.load().gs_exchange_record_bytes(n, cap, row_floats))


def exchange_pack(g2d, vis, heur, max_scale_px, groups, record, cap, workspace, clear_acc=None, clear_vis=None):
  """record: uint8 CUDA tensor of >= exchange_record_bytes; vis (n,) / heur (n,2)-like strided views as in
  accumulate_stats."""
  n = g2d.shape[0]
  L.check(L.load().gs_exchange_pack(n, L.ptr(g2d), vis.data_ptr(), vis.stride(0), heur.data_ptr(), heur.stride(0),
                                    L.ptr(max_scale_px), L.ptr(clear_acc), 0 if clear_acc is None else clear_acc.shape[1],
                                    L.ptr(clear_vis), C.byref(groups), record.data_ptr(), int(cap), L.ptr(workspace),
                                    workspace.numel(), L.stream()))


def exchange_fold(n, record_ptrs, cap, groups, split_alpha, prune_alpha, state, scene_visible=None, stats_rows=None,
                  grad_rows=None):
  """record_ptrs: device (or peer-mapped) addresses of the step's records in image order.  stats_rows: (begin, end) of
  the rows whose controller statistics this call folds (default: all); grad_rows: likewise for the gradient rows and
  `visible` (the row-sharded optimiser folds its own slice only)."""
  arr = (C.c_void_p * len(record_ptrs))(*[int(p) for p in record_ptrs])
  b, e = (0, n) if stats_rows is None else stats_rows
  gb, ge = (0, n) if grad_rows is None else grad_rows
  L.check(L.load().gs_exchange_fold(n, len(record_ptrs), arr, int(cap), C.byref(groups), split_alpha, prune_alpha,
                                    L.ptr(state["prune_cost"]), L.ptr(state["split_score"]),
                                    L.ptr(state["points_in_view"]), L.ptr(state["visibility"]), int(b), int(e), int(gb), int(ge),
                                    L.ptr(scene_visible), L.stream()))


def exchange_publish(indexes, row_begin, param_groups, local_base, peer_bases):
  """gs_exchange_publish: the parameter rows `indexes` (a VisibleIndexes over the slice starting at row_begin) go to
  every peer's copy of the packed parameter buffer.  param_groups: exchange_groups() over the PARAMETER planes."""
  arr = (C.c_void_p * max(len(peer_bases), 1))(*[int(p) for p in peer_bases])
  L.check(L.load().gs_exchange_publish(L.ptr(indexes.indexes), L.ptr(indexes.count), indexes.n, int(row_begin),
                                       C.byref(param_groups), int(local_base), len(peer_bases), arr, L.stream()))


# ---- per-step controller / scene passes (include/gsb200.h section 7) -----------------------------------------------
def mcmc_noise(alpha_logit, log_scaling, rotation, points_in_view, min_views, opacity_threshold, noise_level, randn,
               position):
  """in place on position (n,3); randn (n,3) N(0,1) draws supplied by the caller."""
  n = position.shape[0]
  assert points_in_view.dtype == torch.int16
  L.check(L.load().gs_mcmc_noise(n, L.ptr(_f32(alpha_logit)), L.ptr(_f32(log_scaling)), L.ptr(_f32(rotation)),
                                 L.ptr(points_in_view), int(min_views), float(opacity_threshold), float(noise_level),
                                 L.ptr(_f32(randn)), L.ptr(_f32(position)), L.stream()))


def reg_loss(weights, g2d, depth, log_scaling, vis, sums=None, compute_sums=True, d_opacity=None, d_depth=None,
             d_log_scaling=None):
  """weights = (scale, opacity, aspect).  vis / d_opacity may be strided column views.  Returns sums (4,) on the device:
  Σ scale terms, Σ opacity terms, Σ aspect terms, number of blended rows."""
  n = g2d.shape[0]
  if sums is None:
    sums = torch.zeros((4,), device=g2d.device, dtype=torch.float32)
  L.check(L.load().gs_reg_loss(n, float(weights[0]), float(weights[1]), float(weights[2]), L.ptr(_f32(g2d)),
                               L.ptr(_f32(depth)), L.ptr(_f32(log_scaling)), vis.data_ptr(), vis.stride(0), L.ptr(sums),
                               int(compute_sums), None if d_opacity is None else d_opacity.data_ptr(),
                               0 if d_opacity is None else d_opacity.stride(0), L.ptr(d_depth), L.ptr(d_log_scaling),
                               L.stream()))
  return sums


# ---- multi-scale SSIM loss (include/gsb200.h section 8) -------------------------------------------------------------
def ssim_workspace(height, width, channels, levels, device="cuda"):
  nbytes = int(L.load().gs_ssim_workspace_bytes(height, width, channels, levels))
  return torch.empty((nbytes,), device=device, dtype=torch.uint8)


def ssim_counts(height, width, channels, levels):
  """number of map entries in the valid crop of every level (denominators of ssim_l)."""
  out = []
  for _ in range(levels):
    out.append(max(height - 10, 0) * max(width - 10, 0) * channels)
    height, width = height // 2, width // 2
  return out


def ssim_loss(pred, target, levels=3, weight=1.0, d_pred=None, sums=None, workspace=None):
  """pred (H,W,C) raw render; target float32 or uint8 (H,W,C).  d_pred: gradient ADDED in place.  Returns sums (levels,)
  on the device: Σ of the SSIM map over the valid crop per level (loss = weight·mean_l(1 − sums[l]/ssim_counts[l]))."""
  H, W, C_ = pred.shape
  if workspace is None:
    workspace = ssim_workspace(H, W, C_, levels, pred.device)
  if sums is None:
    sums = torch.zeros((levels,), device=pred.device, dtype=torch.float32)
  tf = target if target.dtype == torch.float32 else None
  tu = target if target.dtype == torch.uint8 else None
  assert tf is not None or tu is not None, "target must be float32 or uint8"
  L.check(L.load().gs_ssim_loss(H, W, C_, int(levels), L.ptr(_f32(pred)), L.ptr(tf), L.ptr(tu), float(weight), L.ptr(d_pred),
                                L.ptr(sums), L.ptr(workspace), workspace.numel(), L.stream()))
  return sums


def photometric_loss(pred, target, w_l1, w_mse, w_ssim, levels, d_pred, loss_sums, ssim_sums, workspace):
  """gs_photometric_loss: loss_l1_mse + ssim_loss in one call (reference trainer/trainer.py:465-498); d_pred WRITTEN,
  loss_sums (2,) and ssim_sums (levels,) zeroed and filled on the device."""
  H, W, C_ = pred.shape
  tf = target if target.dtype == torch.float32 else None
  tu = target if target.dtype == torch.uint8 else None
  assert tf is not None or tu is not None, "target must be float32 or uint8"
  L.check(L.load().gs_photometric_loss(H, W, C_, int(levels), L.ptr(_f32(pred)), L.ptr(tf), L.ptr(tu), float(w_l1), float(w_mse),
                                       float(w_ssim), L.ptr(d_pred), L.ptr(loss_sums), L.ptr(ssim_sums), L.ptr(workspace),
                                       workspace.numel(), L.stream()))
  return loss_sums, ssim_sums


# ---- colour MLP path (include/gsb200.h section 9) -------------------------------------------------------------------
COLOR_WEIGHTS = ("base_w1", "base_b1", "base_w2", "base_b2", "dir_wsh", "dir_bsh", "dir_w1", "dir_b1", "dir_w2", "dir_b2")


def color_model(weights, point_features=16, glo_features=32, hidden=32, sh_degree=5):
  """gs_color_mlp from a dict COLOR_WEIGHTS -> fp32 CUDA tensors shaped like the reference's nn.Linear parameters
  (ColorModel.base_model / directional_model, scene/color_model.py:106-121).  Keeps the tensors alive."""
  ts = [_f32(weights[k]) for k in COLOR_WEIGHTS]
  m = L.ColorMLP(point_features, glo_features, hidden, sh_degree, (C.c_void_p * 10)(*[t.data_ptr() for t in ts]))
  m._keep = ts
  return m


def color_mlp_fwd(model, feature, position, cam_pos, glo, in_view=None, specular_weight=1.0, rgb=None, specular=None):
  """gs_color_mlp_fwd.  in_view: (n,) or a column view (e.g. g2d[:, 5]); rows with in_view <= 0 are left untouched.
  Returns (rgb (n,3), specular or None)."""
  n = feature.shape[0]
  if rgb is None:
    rgb = torch.zeros((n, 3), device=feature.device, dtype=torch.float32)
  cam = _cam_pos3(cam_pos)
  L.check(L.load().gs_color_mlp_fwd(C.byref(model), n, None if in_view is None else in_view.data_ptr(),
                                    0 if in_view is None else in_view.stride(0), L.ptr(_f32(feature)), L.ptr(_f32(position)),
                                    cam, L.ptr(_f32(glo)), float(specular_weight), L.ptr(rgb), L.ptr(specular), L.stream()))
  return rgb, specular


def color_mlp_bwd(model, grads, feature, position, cam_pos, glo, row_w, d_rgb, d_feature, d_glo, specular_weight=1.0,
                  w_specular_reg=0.0, reg_sums=None, accumulate=False):
  """gs_color_mlp_bwd.  grads: a color_model() over the gradient tensors (accumulated); row_w / d_rgb may be column
  views; d_feature (n, point_features) is written for the rows with row_w > 0, d_glo accumulated."""
  n = feature.shape[0]
  assert row_w.dtype == torch.float32 and d_rgb.dtype == torch.float32 and row_w.is_cuda and d_rgb.is_cuda
  assert d_rgb.dim() == 2 and d_rgb.stride(1) == 1 and d_feature.is_contiguous()
  L.check(L.load().gs_color_mlp_bwd(C.byref(model), C.byref(grads), n, row_w.data_ptr(), row_w.stride(0),
                                    L.ptr(_f32(feature)), L.ptr(_f32(position)), _cam_pos3(cam_pos), L.ptr(_f32(glo)),
                                    float(specular_weight), d_rgb.data_ptr(), d_rgb.stride(0), float(w_specular_reg),
                                    L.ptr(reg_sums), L.ptr(d_feature), L.ptr(d_glo), int(accumulate), L.stream()))


# ---- point clustering (include/gsb200.h section 10) -----------------------------------------------------------------
def assign_clusters(x, centroids):
  """reference visibility/cluster.py:139-148: nearest centroid of every point -> (n,) int32."""
  n, d = x.shape
  labels = torch.empty((n,), device=x.device, dtype=torch.int32)
  L.check(L.load().gs_assign_clusters(n, d, centroids.shape[0], L.ptr(_f32(x)), L.ptr(_f32(centroids)), L.ptr(labels),
                                      L.stream()))
  return labels


def kmeans(x, k=None, iters=100, centroids=None, generator=None):
  """reference visibility/cluster.py:150-180: kmeans(x, k, iters) -> (labels (n,) int32, centroids (k,d)).
  centroids: initial values (copied); None draws k distinct points with torch.randperm like the reference."""
  n, d = x.shape
  if centroids is None:
    centroids = x[torch.randperm(n, device=x.device, generator=generator)[:k]]
  centroids = _f32(centroids).clone().contiguous()
  k = centroids.shape[0]
  labels = torch.empty((n,), device=x.device, dtype=torch.int32)
  nbytes = int(L.load().gs_kmeans_workspace_bytes(k, d))
  ws = torch.empty((nbytes,), device=x.device, dtype=torch.uint8)
  L.check(L.load().gs_kmeans(n, d, k, iters, L.ptr(_f32(x)), L.ptr(centroids), L.ptr(labels), L.ptr(ws), nbytes, L.stream()))
  return labels, centroids


def view_features(labels, num_clusters, vis, indexes=None, threshold=0.01):
  """PointClusters.view_features (reference visibility/cluster.py:40-55): (num_clusters,) Σ of the visibilities > threshold
  per cluster.  vis: the (n,) visibility plane of a render, or — with indexes (v,) int64 — the compacted (idx, visibility)
  lists of query_visibility."""
  out = torch.empty((num_clusters,), device=labels.device, dtype=torch.float32)
  n = labels.shape[0]
  v = 0 if indexes is None else indexes.shape[0]
  assert labels.dtype == torch.int32 and (indexes is None or indexes.dtype == torch.int64)
  L.check(L.load().gs_view_features(n, L.ptr(_f32(vis)), L.ptr(indexes), v, L.ptr(labels), float(threshold), int(num_clusters),
                                    L.ptr(out), L.stream()))
  return out


def camera_table(cameras):
  """list of lib cameras (gs_camera) -> (C, 20) float32 host tensor: image_t_world = K·T_camera_world | W, H, near, far."""
  rows = []
  for c in cameras:
    T = torch.tensor([c.T[i] for i in range(16)], dtype=torch.float32).reshape(4, 4)
    K = torch.zeros((4, 4), dtype=torch.float32)
    K[0, 0], K[1, 1], K[0, 2], K[1, 2], K[2, 2], K[3, 3] = c.fx, c.fy, c.cx, c.cy, 1.0, 1.0
    rows.append(torch.cat([(K @ T).reshape(-1), torch.tensor([c.width, c.height, c.near_plane, c.far_plane], dtype=torch.float32)]))
  return torch.stack(rows)


def point_visibility(points, cameras):
  """reference visibility/query_points.py:89-92: (n,) int32 number of cameras that see each point.  cameras: (C,20) table
  (camera_table) or a list of lib cameras."""
  if not isinstance(cameras, torch.Tensor):
    cameras = camera_table(cameras)
  cams = cameras.to(points.device, torch.float32).contiguous()
  counts = torch.empty((points.shape[0],), device=points.device, dtype=torch.int32)
  L.check(L.load().gs_point_visibility(points.shape[0], L.ptr(_f32(points)), cams.shape[0], L.ptr(cams), L.ptr(counts), L.stream()))
  return counts


def estimate_scale(points, num_neighbors=3):
  """reference gaussians/loading.py:45-61: mean distance to the num_neighbors nearest neighbours -> (n,) float32."""
  n = points.shape[0]
  assert points.shape[1] == 3
  scales = torch.empty((n,), device=points.device, dtype=torch.float32)
  nbytes = int(L.load().gs_knn_scale_workspace_bytes(n))
  ws = torch.empty((nbytes,), device=points.device, dtype=torch.uint8)
  L.check(L.load().gs_knn_scale(n, L.ptr(_f32(points)), num_neighbors, L.ptr(scales), L.ptr(ws), nbytes, L.stream()))
  return scales


def adam_step(param, grad, exp_avg, exp_avg_sq, lr, step, betas=(0.9, 0.999), eps=1e-8, zero_grad=True):
  """gs_adam_step on contiguous fp32 CUDA tensors of equal size (views into larger buffers are fine)."""
  assert param.is_contiguous() and grad.is_contiguous() and exp_avg.is_contiguous() and exp_avg_sq.is_contiguous()
  L.check(L.load().gs_adam_step(param.numel(), L.ptr(_f32(param)), L.ptr(_f32(grad)), L.ptr(exp_avg), L.ptr(exp_avg_sq),
                                float(lr), float(betas[0]), float(betas[1]), float(eps), int(step), int(zero_grad), L.stream()))


def adam_rows(param, grad, exp_avg, exp_avg_sq, steps, touched, lr, betas=(0.8, 0.95), eps=1e-8):
  """gs_adam_rows on a contiguous (rows, dim) table; steps int32 (rows,), touched float32 (rows,) on the device."""
  rows, dim = param.shape
  assert steps.dtype == torch.int32 and touched.dtype == torch.float32 and param.is_contiguous() and grad.is_contiguous()
  L.check(L.load().gs_adam_rows(rows, dim, L.ptr(_f32(param)), L.ptr(_f32(grad)), L.ptr(exp_avg), L.ptr(exp_avg_sq),
                                L.ptr(steps), L.ptr(touched), float(lr), float(betas[0]), float(betas[1]), float(eps),
                                L.stream()))
