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

// ---- live-row compaction (single pass, decoupled look-back) -------------------------------------------
// rows that touch at least one tile, compacted in row order: (depth bits, row) at their rank; live count -> *live_out
constexpr int LC_THREADS = 256;
constexpr int LC_ITEMS = 8;
constexpr int LC_TILE = LC_THREADS * LC_ITEMS;

__device__ __forceinline__ int block_scan_256(int v, int *s_warp, int &block_total) {  // exclusive, 256 threads
  int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  int inc = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    int t = __shfl_up_sync(GS_FULL, inc, o);
    if (lane >= o) inc += t;
  }
  if (lane == 31) s_warp[w] = inc;
  __syncthreads();
  int wo = 0, tot = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    int c = s_warp[i];
    if (i < w) wo += c;
    tot += c;
  }
  block_total = tot;
  return wo + inc - v;
}

__global__ void __launch_bounds__(LC_THREADS) k_live_compact(const int32_t *__restrict__ counts, const float *__restrict__ depth,
                                                             int64_t m, uint32_t *__restrict__ status, uint32_t *__restrict__ ticket,
                                                             uint32_t *__restrict__ depth_keys, uint32_t *__restrict__ rows,
                                                             int32_t *__restrict__ live_out) {
  __shared__ int s_warp[8];
  __shared__ int s_tile;
  __shared__ uint32_t s_prefix;
  if (threadIdx.x == 0) s_tile = (int)atomicAdd(ticket, 1u);
  __syncthreads();
  const int tile = s_tile;
  int64_t base = (int64_t)tile * LC_TILE + (int64_t)threadIdx.x * LC_ITEMS;
  int c[LC_ITEMS];
  if (base + LC_ITEMS <= m) {
    const int4 *p = reinterpret_cast<const int4 *>(counts + base);
    int4 a = p[0], b = p[1];
    c[0] = a.x; c[1] = a.y; c[2] = a.z; c[3] = a.w; c[4] = b.x; c[5] = b.y; c[6] = b.z; c[7] = b.w;
  } else {
#pragma unroll
    for (int i = 0; i < LC_ITEMS; ++i) c[i] = (base + i < m) ? counts[base + i] : 0;
  }
  int sum = 0;
#pragma unroll
  for (int i = 0; i < LC_ITEMS; ++i) sum += c[i] > 0;
  int tot;
  int ex = block_scan_256(sum, s_warp, tot);
  if (threadIdx.x == 0) gs_st_status(status + tile, (uint32_t)tot | (tile == 0 ? GS_LB_INCL : GS_LB_LOCAL));
  if (threadIdx.x < 32) {
    uint32_t prefix = gs_lookback_warp(status, tile);
    if (threadIdx.x == 0) {
      if (tile > 0) gs_st_status(status + tile, (prefix + (uint32_t)tot) | GS_LB_INCL);
      s_prefix = prefix;
      if (tile == (int)gridDim.x - 1) *live_out = (int32_t)(prefix + (uint32_t)tot);
    }
  }
  __syncthreads();
  uint32_t pos = s_prefix + (uint32_t)ex;
#pragma unroll
  for (int i = 0; i < LC_ITEMS; ++i) {
    if (c[i] > 0) {
      depth_keys[pos] = __float_as_uint(depth[base + i]);
      rows[pos] = (uint32_t)(base + i);
      ++pos;
    }
  }
}

// ---- emission (single pass: chained scan of the per-row tile counts in depth order + the pairs) --------
// thread handles EM_ITEMS consecutive live rows of the depth order; writes (tile, row) pairs at their offsets.
// The last live tile publishes the counters: [0] total overlaps, [1] overflow flag, [2] min(total, capacity).
constexpr int EM_THREADS = 256;
constexpr int EM_ITEMS = 1;
constexpr int EM_TILE = EM_THREADS * EM_ITEMS;

__global__ void __launch_bounds__(EM_THREADS) k_tile_emit(const float *__restrict__ g2d, const float *__restrict__ depth,
                                                          const int32_t *__restrict__ counts, const int32_t *__restrict__ live_dev,
                                                          int64_t m, float thr, int W, int H,
                                                          const uint32_t *__restrict__ sorted_rows, int64_t max_overlaps,
                                                          uint32_t *__restrict__ status, uint32_t *__restrict__ ticket,
                                                          uint32_t *__restrict__ tile_keys, uint32_t *__restrict__ tile_vals,
                                                          int32_t *__restrict__ counters) {
  __shared__ int s_warp[8];
  __shared__ int s_tile;
  __shared__ uint32_t s_prefix;
  if (threadIdx.x == 0) s_tile = (int)atomicAdd(ticket, 1u);
  __syncthreads();
  const int tile = s_tile;
  int64_t live = min((int64_t)*live_dev, m);
  int64_t tbase = (int64_t)tile * EM_TILE;
  if (tbase >= live) return;
  int64_t j0 = tbase + (int64_t)threadIdx.x * EM_ITEMS;
  int64_t row[EM_ITEMS];
  int c[EM_ITEMS];
  int sum = 0;
#pragma unroll
  for (int i = 0; i < EM_ITEMS; ++i) {
    row[i] = (j0 + i < live) ? (int64_t)sorted_rows[j0 + i] : -1;
    c[i] = row[i] >= 0 ? counts[row[i]] : 0;
    sum += c[i];
  }
  int tot;
  int ex = block_scan_256(sum, s_warp, tot);
  if (threadIdx.x == 0) gs_st_status(status + tile, (uint32_t)tot | (tile == 0 ? GS_LB_INCL : GS_LB_LOCAL));
  if (threadIdx.x < 32) {
    uint32_t prefix = gs_lookback_warp(status, tile);
    if (threadIdx.x == 0) {
      if (tile > 0) gs_st_status(status + tile, (prefix + (uint32_t)tot) | GS_LB_INCL);
      s_prefix = prefix;
      if (tbase + EM_TILE >= live) {  // last live tile
        int64_t o = (int64_t)prefix + tot;
        counters[0] = (int32_t)o;
        counters[1] = (o > max_overlaps) ? 1 : 0;
        counters[2] = (int32_t)min(o, max_overlaps);
      }
    }
  }
  __syncthreads();
  int TX = (W + GS_TILE - 1) / GS_TILE;
  int64_t o = (int64_t)s_prefix + ex;
#pragma unroll 1
  for (int i = 0; i < EM_ITEMS; ++i) {
    if (c[i] <= 0) continue;
    Footprint fp = load_footprint(g2d, depth, row[i], thr, W, H);
    for (int ty = fp.ty0; ty <= fp.ty1; ++ty)
      for (int tx = fp.tx0; tx <= fp.tx1; ++tx)
        if (tile_hit(fp, tx, ty, W, H)) {
          if (o < max_overlaps) {
            tile_keys[o] = (uint32_t)(ty * TX + tx);
            tile_vals[o] = (uint32_t)row[i];
          }
          ++o;
        }
  }
}

// 4 keys per thread (one 16-B load) + the two neighbours
__global__ void __launch_bounds__(256) k_tile_ranges(const uint32_t *__restrict__ keys, const int32_t *__restrict__ n_dev,
                                                     int64_t n_cap, int32_t *__restrict__ ranges) {
  int64_t n = min((int64_t)*n_dev, n_cap);
  int64_t o = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 4;
  if (o >= n) return;
  uint32_t k[6];
  if (o + 4 <= n) {
    uint4 q = *reinterpret_cast<const uint4 *>(keys + o);
    k[1] = q.x; k[2] = q.y; k[3] = q.z; k[4] = q.w;
  } else {
#pragma unroll
    for (int i = 0; i < 4; ++i) k[1 + i] = (o + i < n) ? keys[o + i] : 0xffffffffu;
  }
  k[0] = o > 0 ? keys[o - 1] : 0xffffffffu;
  k[5] = (o + 4 < n) ? keys[o + 4] : 0xffffffffu;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    if (o + i >= n) break;
    uint32_t t = k[1 + i];
    if (o + i == 0 || k[i] != t) ranges[2 * t] = (int32_t)(o + i);
    if (o + i == n - 1 || k[2 + i] != t) ranges[2 * t + 1] = (int32_t)(o + i + 1);
  }
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

// workspace layout (all 256-B aligned)
struct TileWs {
  int32_t *counts;      // m
  uint32_t *dk[2];      // m each: depth keys ping/pong
  uint32_t *dv[2];      // m each: rows ping/pong
  uint32_t *tk[2];      // max_overlaps each
  uint32_t *tv;         // max_overlaps: the other value buffer is the caller's overlap_to_point
  uint32_t *scan_status;  // [0] ticket(compact) [1] ticket(emit) ... then LC tiles + EM tiles status words
  size_t scan_status_bytes;
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
  w.tk[0] = (uint32_t *)take(oo * 4); w.tk[1] = (uint32_t *)take(oo * 4);
  w.tv = (uint32_t *)take(oo * 4);
  size_t lc = (mm + LC_TILE - 1) / LC_TILE, em = (mm + EM_TILE - 1) / EM_TILE;
  w.scan_status_bytes = (64 + lc + em) * 4;
  w.scan_status = (uint32_t *)take(w.scan_status_bytes);
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
  GS_REQUIRE(m >= 0 && m < ((int64_t)1 << 30) && max_overlaps >= 0 && max_overlaps < ((int64_t)1 << 30), "m and max_overlaps must be < 2^30");
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
  // (a) compact the live rows (>= 1 tile) in row order (one chained-scan kernel), then sort them by depth: stable
  //     32-bit LSD sort, ties keep row order.  Culled / off-screen rows never enter the sort.
  int32_t *live = counters + 3;
  GS_CUDA(cudaMemsetAsync(w.scan_status, 0, w.scan_status_bytes, st));
  int lc_tiles = gs_div_up(m, LC_TILE), em_tiles = gs_div_up(m, EM_TILE);
  uint32_t *lc_status = w.scan_status + 64, *em_status = lc_status + lc_tiles;
  k_live_compact<<<lc_tiles, LC_THREADS, 0, st>>>(counts, depth, m, lc_status, w.scan_status, w.dk[0], w.dv[0], live);
  GS_LAUNCH_CHECK();
  int in_b = 0;
  int rc = gs_radix_sort_pingpong(w.dk[0], w.dv[0], w.dk[1], w.dv[1], m, live, 0, 32, w.sort_ws, &in_b, st);
  if (rc != GS_OK) return rc;
  uint32_t *srows = w.dv[in_b];
  // (b)+(c) offsets of each depth-ordered row's overlaps (chained scan) and the (tile, row) pairs, one kernel;
  // the value buffers are chosen so that the last sort pass writes overlap_to_point directly (no copy).
  int bits = 1;
  while (((int64_t)1 << bits) < T) ++bits;
  int passes = (bits + 7) / 8;
  uint32_t *ka = w.tk[0], *kb = w.tk[1];
  uint32_t *va = (passes % 2) ? w.tv : (uint32_t *)overlap_to_point;
  uint32_t *vb = (passes % 2) ? (uint32_t *)overlap_to_point : w.tv;
  k_tile_emit<<<em_tiles, EM_THREADS, 0, st>>>(g2d, depth, counts, live, m, thr, width, height, srows, max_overlaps,
                                               em_status, w.scan_status + 1, ka, va, counters);
  GS_LAUNCH_CHECK();
  if (max_overlaps == 0) return GS_OK;
  // (d) stable sort by tile id
  rc = gs_radix_sort_pingpong(ka, va, kb, vb, max_overlaps, counters + 2, 0, bits, w.sort_ws, &in_b, st);
  if (rc != GS_OK) return rc;
  uint32_t *tkeys = in_b ? kb : ka;
  k_tile_ranges<<<gs_div_up(max_overlaps, 1024), 256, 0, st>>>(tkeys, counters + 2, max_overlaps, tile_ranges);
  GS_LAUNCH_CHECK();
  return GS_OK;
}
