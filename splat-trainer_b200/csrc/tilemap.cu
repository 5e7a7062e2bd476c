// tilemap.cu — tile-overlap mapping (mapper inside render_projected, reference
// scene/mlp_scene.py:418,377).  count -> depth sort of rows -> scan -> emit (tile,row) in depth
// order -> stable sort by tile -> ranges.  Everything stays on the device: the overlap total is a
// device counter, buffers are sized by the caller's max_overlaps (no host read-back).
#include "common.cuh"
#include "footprint.cuh"

int gs_radix_sort_pingpong(uint32_t *ka, uint32_t *va, uint32_t *kb, uint32_t *vb, int64_t n, const int32_t *n_dev,
                           int begin_bit, int end_bit, void *workspace, int *result_in_b, cudaStream_t st);

__device__ __forceinline__ Footprint load_footprint(const float *__restrict__ g2d, const float *__restrict__ depth,
                                                    int64_t i, float thr, int W, int H) {
  const float2 *g = reinterpret_cast<const float2 *>(g2d + GS_G2D_STRIDE * i);
  float2 a = g[0], b = g[1], c = g[2];
  return row_footprint(a.x, a.y, b.x, b.y, c.x, c.y, depth[i], thr, W, H);
}

// per row: tiles touched
__global__ void __launch_bounds__(256) k_tile_count(const float *__restrict__ g2d, const float *__restrict__ depth,
                                                    int64_t m, float thr, int W, int H, int32_t *__restrict__ counts) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= m) return;
  Footprint fp = load_footprint(g2d, depth, i, thr, W, H);
  counts[i] = footprint_count(fp, W, H);
}

// rows that touch at least one tile, compacted in row order: (depth bits, row) at rank[i]
__global__ void __launch_bounds__(256) k_compact_live(const int32_t *__restrict__ counts, const float *__restrict__ depth,
                                                      int64_t m, const int32_t *__restrict__ rank,
                                                      uint32_t *__restrict__ depth_keys, uint32_t *__restrict__ rows) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= m || counts[i] <= 0) return;
  int r = rank[i];
  depth_keys[r] = __float_as_uint(depth[i]);
  rows[r] = (uint32_t)i;
}

__global__ void __launch_bounds__(256) k_sum_i32(const int32_t *__restrict__ v, int64_t n, int32_t *__restrict__ out) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int x = i < n ? v[i] : 0;
  x = __reduce_add_sync(GS_FULL, x);
  __shared__ int s[8];
  if ((threadIdx.x & 31) == 0) s[threadIdx.x >> 5] = x;
  __syncthreads();
  if (threadIdx.x == 0) {
    int t = 0;
    for (int k = 0; k < 8; ++k) t += s[k];
    if (t) atomicAdd(out, t);
  }
}

// counters: [0] total overlaps, [1] overflow flag, [2] min(total, capacity), [3] rows touching >= 1 tile
__global__ void k_finish_counters(int32_t *counters, int64_t max_overlaps) {
  int o = counters[0];
  counters[1] = (o > max_overlaps) ? 1 : 0;
  counters[2] = (int32_t)min((int64_t)o, max_overlaps);
}

// thread j = j-th live row in depth order; writes its (tile, row) pairs at offsets[j]...
__global__ void __launch_bounds__(256) k_tile_emit(const float *__restrict__ g2d, const float *__restrict__ depth,
                                                   const int32_t *__restrict__ live_dev, int64_t m, float thr, int W, int H,
                                                   const uint32_t *__restrict__ sorted_rows,
                                                   const int32_t *__restrict__ offsets, int64_t max_overlaps,
                                                   uint32_t *__restrict__ tile_keys, uint32_t *__restrict__ tile_vals) {
  int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= min((int64_t)*live_dev, m)) return;
  int64_t i = sorted_rows[j];
  Footprint fp = load_footprint(g2d, depth, i, thr, W, H);
  if (!fp.valid) return;
  int TX = (W + GS_TILE - 1) / GS_TILE;
  int64_t o = offsets[j];
  for (int ty = fp.ty0; ty <= fp.ty1; ++ty)
    for (int tx = fp.tx0; tx <= fp.tx1; ++tx)
      if (tile_hit(fp, tx, ty, W, H)) {
        if (o < max_overlaps) {
          tile_keys[o] = (uint32_t)(ty * TX + tx);
          tile_vals[o] = (uint32_t)i;
        }
        ++o;
      }
}

__global__ void __launch_bounds__(256) k_tile_ranges(const uint32_t *__restrict__ keys, const int32_t *__restrict__ n_dev,
                                                     int64_t n_cap, int32_t *__restrict__ ranges) {
  int64_t n = min((int64_t)*n_dev, n_cap);
  int64_t o = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (o >= n) return;
  uint32_t t = keys[o];
  if (o == 0 || keys[o - 1] != t) ranges[2 * t] = (int32_t)o;
  if (o == n - 1 || keys[o + 1] != t) ranges[2 * t + 1] = (int32_t)(o + 1);
}

// workspace layout (all 256-B aligned)
struct TileWs {
  int32_t *counts;      // m
  uint32_t *dk[2];      // m each: depth keys ping/pong
  uint32_t *dv[2];      // m each: rows ping/pong
  int32_t *offsets;     // m
  uint32_t *tk[2];      // max_overlaps each
  uint32_t *tv;         // max_overlaps: the other value buffer is the caller's overlap_to_point
  int32_t *partials;    // scan partials
  void *sort_ws;
  size_t total;
};

static TileWs carve(void *base, int64_t m, int64_t maxo) {
  TileWs w;
  char *p = (char *)base;
  auto take = [&](size_t bytes) { char *r = p; p += gs_align256(bytes); return r; };
  size_t mm = (size_t)(m > 0 ? m : 1), oo = (size_t)(maxo > 0 ? maxo : 1);
  w.counts = (int32_t *)take(mm * 4);
  w.dk[0] = (uint32_t *)take(mm * 4); w.dk[1] = (uint32_t *)take(mm * 4);
  w.dv[0] = (uint32_t *)take(mm * 4); w.dv[1] = (uint32_t *)take(mm * 4);
  w.offsets = (int32_t *)take(mm * 4);
  w.tk[0] = (uint32_t *)take(oo * 4); w.tk[1] = (uint32_t *)take(oo * 4);
  w.tv = (uint32_t *)take(oo * 4);
  w.partials = (int32_t *)take((size_t)gs_scan_partials(m > 0 ? m : 1) * 4);
  int64_t big = m > maxo ? m : maxo;
  w.sort_ws = take(gs_radix_sort_workspace_bytes(big));
  w.total = (size_t)(p - (char *)base);
  return w;
}

extern "C" size_t gs_tile_map_workspace_bytes(int64_t m, int64_t max_overlaps, int width, int height) {
  (void)width; (void)height;
  return carve(nullptr, m, max_overlaps).total + 256;
}

extern "C" int gs_tile_count(const gs_raster_config *h_cfg, int64_t m, const float *g2d, const float *depth, int width,
                             int height, int32_t *counts, int32_t *counters, void *stream) {
  GS_REQUIRE(h_cfg && h_cfg->tile_size == GS_TILE, "tile_size must be 16");
  cudaStream_t st = (cudaStream_t)stream;
  GS_CUDA(cudaMemsetAsync(counters, 0, 4 * sizeof(int32_t), st));
  if (m <= 0) return GS_OK;
  k_tile_count<<<gs_div_up(m, 256), 256, 0, st>>>(g2d, depth, m, h_cfg->alpha_threshold, width, height, counts);
  GS_LAUNCH_CHECK();
  k_sum_i32<<<gs_div_up(m, 256), 256, 0, st>>>(counts, m, counters);
  GS_LAUNCH_CHECK();
  return GS_OK;
}

extern "C" int gs_tile_map(const gs_raster_config *h_cfg, int64_t m, const float *g2d, const float *depth,
                           const int32_t *counts_in, int width,
                           int height, int64_t max_overlaps, void *workspace, size_t workspace_bytes,
                           int32_t *overlap_to_point, int32_t *tile_ranges, int32_t *counters, void *stream) {
  GS_REQUIRE(h_cfg && h_cfg->tile_size == GS_TILE, "tile_size must be 16");
  GS_REQUIRE(m >= 0 && m < ((int64_t)1 << 31) && max_overlaps >= 0 && max_overlaps < ((int64_t)1 << 31), "sizes must fit int32");
  GS_REQUIRE(width > 0 && height > 0, "bad image size");
  cudaStream_t st = (cudaStream_t)stream;
  int TX = (width + GS_TILE - 1) / GS_TILE, TY = (height + GS_TILE - 1) / GS_TILE;
  int64_t T = (int64_t)TX * TY;
  uintptr_t aligned = ((uintptr_t)workspace + 255) & ~(uintptr_t)255;
  TileWs w = carve((void *)aligned, m, max_overlaps);
  if (workspace_bytes < w.total + (aligned - (uintptr_t)workspace)) {
    gs_set_error("gs_tile_map: workspace too small (%zu < %zu)", workspace_bytes, w.total + 256);
    return GS_ERR_WORKSPACE;
  }
  GS_CUDA(cudaMemsetAsync(counters, 0, 4 * sizeof(int32_t), st));
  GS_CUDA(cudaMemsetAsync(tile_ranges, 0, (size_t)T * 2 * sizeof(int32_t), st));
  if (m == 0) return GS_OK;
  float thr = h_cfg->alpha_threshold;
  int blocks = gs_div_up(m, 256);
  const int32_t *counts = counts_in;  // the fused path gets them from gs_project_fwd (same footprint code)
  if (!counts) {
    k_tile_count<<<blocks, 256, 0, st>>>(g2d, depth, m, thr, width, height, w.counts);
    GS_LAUNCH_CHECK();
    counts = w.counts;
  }
  // (a) compact the live rows (>= 1 tile) in row order, then sort them by depth: stable 32-bit LSD sort, ties keep
  //     row order.  Culled / off-screen rows never enter the sort.
  int32_t *live = counters + 3;
  int rc = gs_exclusive_scan_i32(counts, nullptr, w.offsets, m, nullptr, w.partials, live, st, /*as_flag=*/1);
  if (rc != GS_OK) return rc;
  k_compact_live<<<blocks, 256, 0, st>>>(counts, depth, m, w.offsets, w.dk[0], w.dv[0]);
  GS_LAUNCH_CHECK();
  int in_b = 0;
  rc = gs_radix_sort_pingpong(w.dk[0], w.dv[0], w.dk[1], w.dv[1], m, live, 0, 32, w.sort_ws, &in_b, st);
  if (rc != GS_OK) return rc;
  uint32_t *srows = w.dv[in_b];
  // (b) offsets of each depth-ordered row's overlaps; total -> counters[0]
  rc = gs_exclusive_scan_i32(counts, (const int32_t *)srows, w.offsets, m, live, w.partials, counters, st);
  if (rc != GS_OK) return rc;
  k_finish_counters<<<1, 1, 0, st>>>(counters, max_overlaps);
  if (max_overlaps == 0) return GS_OK;
  // (c)+(d) emit (tile, row) pairs in depth order, then stable sort by tile id.  The value buffers are
  // chosen so that the last sort pass writes overlap_to_point directly (no copy).
  int bits = 1;
  while (((int64_t)1 << bits) < T) ++bits;
  int passes = (bits + 7) / 8;
  uint32_t *ka = w.tk[0], *kb = w.tk[1];
  uint32_t *va = (passes % 2) ? w.tv : (uint32_t *)overlap_to_point;
  uint32_t *vb = (passes % 2) ? (uint32_t *)overlap_to_point : w.tv;
  k_tile_emit<<<blocks, 256, 0, st>>>(g2d, depth, live, m, thr, width, height, srows, w.offsets, max_overlaps, ka, va);
  GS_LAUNCH_CHECK();
  rc = gs_radix_sort_pingpong(ka, va, kb, vb, max_overlaps, counters + 2, 0, bits, w.sort_ws, &in_b, st);
  if (rc != GS_OK) return rc;
  uint32_t *tkeys = in_b ? kb : ka;
  k_tile_ranges<<<gs_div_up(max_overlaps, 256), 256, 0, st>>>(tkeys, counters + 2, max_overlaps, tile_ranges);
  GS_LAUNCH_CHECK();
  return GS_OK;
}
