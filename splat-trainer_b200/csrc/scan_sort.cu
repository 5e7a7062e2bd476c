// scan_sort.cu — exclusive scan and stable LSD radix sort of (u32 key, u32 value) pairs.
//
// Both are deterministic three-kernel schemes (tile reduce -> scan of tile totals -> tile apply):
// HBM-bound integer work, coalesced 128-B warp accesses, no atomics on the data path.  The sort
// serves SURVEY §8 a8 (the (tile, depth) key sort inside render_projected, reference
// scene/mlp_scene.py:418): rows are sorted by the 32 depth bits (4 passes over M rows), overlaps
// are emitted in that order and stably sorted by tile id (1–2 passes over O pairs), which equals
// one stable sort by the 64-bit key tile<<32|depth but moves ~4.5x fewer bytes.
#include <stdarg.h>

#include "common.cuh"

// ---- error plumbing ---------------------------------------------------------------------------
static thread_local char g_err[512] = "";
void gs_set_error(const char *fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
extern "C" const char *gs_last_error(void) { return g_err; }
extern "C" int gs_version(void) { return GSB200_VERSION; }
extern "C" void gs_default_raster_config(gs_raster_config *c) {
  c->tile_size = GS_TILE;
  c->alpha_threshold = 1.0f / 255.0f;
  c->clamp_max_alpha = 0.99f;
  c->saturate_threshold = 0.9999f;
  c->blur_cov = 0.3f;
  c->antialias = 0;
  c->compute_visibility = 0;
  c->compute_point_heuristic = 0;
  c->render_median_depth = 0;
  c->frustum_clamp = 1.3f;
}

// ---- exclusive scan ---------------------------------------------------------------------------
constexpr int SC_THREADS = 256;
constexpr int SC_ITEMS = 16;
constexpr int SC_TILE = SC_THREADS * SC_ITEMS;

int64_t gs_scan_partials(int64_t n) { return (n + SC_TILE - 1) / SC_TILE + 8; }

__device__ __forceinline__ int block_exclusive_scan(int v, int *s_warp, int &block_total) {
  // inclusive warp scan
  int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  int inc = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    int t = __shfl_up_sync(GS_FULL, inc, o);
    if (lane >= o) inc += t;
  }
  if (lane == 31) s_warp[w] = inc;
  __syncthreads();
  if (w == 0) {
    int nw = blockDim.x >> 5;
    int x = lane < nw ? s_warp[lane] : 0;
    int xi = x;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      int t = __shfl_up_sync(GS_FULL, xi, o);
      if (lane >= o) xi += t;
    }
    if (lane < nw) s_warp[lane] = xi - x;  // exclusive warp offsets
    if (lane == nw - 1) s_warp[32] = xi;
  }
  __syncthreads();
  block_total = s_warp[32];
  return s_warp[w] + inc - v;
}

__global__ void __launch_bounds__(SC_THREADS) k_scan_reduce(const int32_t *__restrict__ in, const int32_t *__restrict__ idx,
                                                            int64_t n_cap, const int32_t *__restrict__ n_dev,
                                                            int32_t *__restrict__ partials, int as_flag) {
  __shared__ int s_warp[33];
  int64_t n = n_dev ? min((int64_t)*n_dev, n_cap) : n_cap;
  int64_t base = (int64_t)blockIdx.x * SC_TILE;
  int sum = 0;
#pragma unroll
  for (int i = 0; i < SC_ITEMS; ++i) {
    int64_t j = base + i * SC_THREADS + threadIdx.x;
    if (j < n) {
      int v = idx ? in[idx[j]] : in[j];
      sum += as_flag ? (v > 0) : v;
    }
  }
  int tot;
  block_exclusive_scan(sum, s_warp, tot);
  if (threadIdx.x == 0) partials[blockIdx.x] = tot;
}

// single block: exclusive scan of the tile totals in place; grand total -> total
__global__ void __launch_bounds__(1024) k_scan_partials(int32_t *__restrict__ partials, int nblk, int32_t *__restrict__ total) {
  __shared__ int s_warp[33];
  __shared__ int s_carry;
  if (threadIdx.x == 0) s_carry = 0;
  __syncthreads();
  for (int base = 0; base < nblk; base += 1024) {
    int j = base + threadIdx.x;
    int v = j < nblk ? partials[j] : 0;
    int tot;
    int ex = block_exclusive_scan(v, s_warp, tot);
    int carry = s_carry;
    if (j < nblk) partials[j] = carry + ex;
    __syncthreads();
    if (threadIdx.x == 0) s_carry = carry + tot;
    __syncthreads();
  }
  if (threadIdx.x == 0 && total) *total = s_carry;
}

__global__ void __launch_bounds__(SC_THREADS) k_scan_apply(const int32_t *__restrict__ in, const int32_t *__restrict__ idx,
                                                           int32_t *__restrict__ out, int64_t n_cap,
                                                           const int32_t *__restrict__ n_dev,
                                                           const int32_t *__restrict__ partials, int as_flag) {
  __shared__ int s_warp[33];
  int64_t n = n_dev ? min((int64_t)*n_dev, n_cap) : n_cap;
  int64_t base = (int64_t)blockIdx.x * SC_TILE + (int64_t)threadIdx.x * SC_ITEMS;
  int v[SC_ITEMS];
  int sum = 0;
#pragma unroll
  for (int i = 0; i < SC_ITEMS; ++i) {
    int64_t j = base + i;
    v[i] = (j < n) ? (idx ? in[idx[j]] : in[j]) : 0;
    if (as_flag) v[i] = v[i] > 0;
    sum += v[i];
  }
  int tot;
  int ex = block_exclusive_scan(sum, s_warp, tot) + partials[blockIdx.x];
#pragma unroll
  for (int i = 0; i < SC_ITEMS; ++i) {
    int64_t j = base + i;
    if (j < n) out[j] = ex;
    ex += v[i];
  }
}

int gs_exclusive_scan_i32(const int32_t *in, const int32_t *idx, int32_t *out, int64_t n, const int32_t *n_dev,
                          int32_t *partials, int32_t *total, cudaStream_t st, int as_flag) {
  if (n <= 0) {
    if (total) cudaMemsetAsync(total, 0, sizeof(int32_t), st);
    return GS_OK;
  }
  int nblk = gs_div_up(n, SC_TILE);
  k_scan_reduce<<<nblk, SC_THREADS, 0, st>>>(in, idx, n, n_dev, partials, as_flag);
  k_scan_partials<<<1, 1024, 0, st>>>(partials, nblk, total);
  k_scan_apply<<<nblk, SC_THREADS, 0, st>>>(in, idx, out, n, n_dev, partials, as_flag);
  GS_LAUNCH_CHECK();
  return GS_OK;
}

// ---- radix sort -------------------------------------------------------------------------------
constexpr int RS_THREADS = 256;
constexpr int RS_ITEMS = 16;
constexpr int RS_TILE = RS_THREADS * RS_ITEMS;  // 4096 pairs per CTA
constexpr int RS_WARPS = RS_THREADS / 32;

// digit histogram of each 4096-pair tile, stored digit-major: hist[d * nblk + blk]
__global__ void __launch_bounds__(RS_THREADS) k_rs_hist(const uint32_t *__restrict__ keys, int64_t n_cap,
                                                        const int32_t *__restrict__ n_dev, int shift, uint32_t mask,
                                                        int32_t *__restrict__ hist, int nblk) {
  __shared__ int h[256];
  h[threadIdx.x] = 0;
  __syncthreads();
  int64_t n = n_dev ? min((int64_t)*n_dev, n_cap) : n_cap;
  int64_t base = (int64_t)blockIdx.x * RS_TILE;
  int lane = threadIdx.x & 31;
#pragma unroll 4
  for (int i = 0; i < RS_ITEMS; ++i) {
    int64_t j = base + i * RS_THREADS + threadIdx.x;
    bool valid = j < n;
    int d = valid ? (int)((keys[j] >> shift) & mask) : 256 + lane;
    unsigned peers = __match_any_sync(GS_FULL, d);
    if (valid && lane == __ffs(peers) - 1) atomicAdd(&h[d], __popc(peers));
  }
  __syncthreads();
  hist[(int64_t)threadIdx.x * nblk + blockIdx.x] = h[threadIdx.x];
}

// stable scatter: rank of a pair = global base of (digit, tile) + pairs of the same digit in earlier
// warps of the tile + earlier pairs of the same digit in this warp (match_any ballots, no atomics)
__global__ void __launch_bounds__(RS_THREADS) k_rs_scatter(const uint32_t *__restrict__ keys_in, const uint32_t *__restrict__ vals_in,
                                                           uint32_t *__restrict__ keys_out, uint32_t *__restrict__ vals_out,
                                                           int64_t n_cap, const int32_t *__restrict__ n_dev, int shift,
                                                           uint32_t mask, const int32_t *__restrict__ hist_scanned, int nblk) {
  __shared__ int wcnt[RS_WARPS][256];
  int64_t n = n_dev ? min((int64_t)*n_dev, n_cap) : n_cap;
  int64_t base = (int64_t)blockIdx.x * RS_TILE;
  if (base >= n) return;
  int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
#pragma unroll
  for (int i = 0; i < RS_WARPS; ++i) wcnt[i][threadIdx.x] = 0;
  __syncthreads();
  uint32_t k[RS_ITEMS], v[RS_ITEMS];
  int rank[RS_ITEMS];
  int64_t wbase = base + (int64_t)w * (RS_ITEMS * 32);
  unsigned lt = (1u << lane) - 1u;
#pragma unroll
  for (int r = 0; r < RS_ITEMS; ++r) {
    int64_t j = wbase + r * 32 + lane;
    bool valid = j < n;
    k[r] = valid ? keys_in[j] : 0xffffffffu;
    v[r] = valid ? vals_in[j] : 0u;
    int d = valid ? (int)((k[r] >> shift) & mask) : 256 + lane;
    unsigned peers = __match_any_sync(GS_FULL, d);
    int leader = __ffs(peers) - 1;
    int pre = 0;
    if (valid && lane == leader) {
      pre = wcnt[w][d];
      wcnt[w][d] = pre + __popc(peers);
    }
    pre = __shfl_sync(GS_FULL, pre, leader);
    rank[r] = pre + __popc(peers & lt);
    __syncwarp();
  }
  __syncthreads();
  {
    int d = threadIdx.x;
    int run = hist_scanned[(int64_t)d * nblk + blockIdx.x];
#pragma unroll
    for (int i = 0; i < RS_WARPS; ++i) {
      int c = wcnt[i][d];
      wcnt[i][d] = run;
      run += c;
    }
  }
  __syncthreads();
#pragma unroll
  for (int r = 0; r < RS_ITEMS; ++r) {
    int64_t j = wbase + r * 32 + lane;
    if (j < n) {
      int d = (int)((k[r] >> shift) & mask);
      int pos = wcnt[w][d] + rank[r];
      keys_out[pos] = k[r];
      vals_out[pos] = v[r];
    }
  }
}

extern "C" size_t gs_radix_sort_workspace_bytes(int64_t n) {
  int64_t nblk = (n + RS_TILE - 1) / RS_TILE;
  if (nblk < 1) nblk = 1;
  size_t hist = gs_align256((size_t)256 * nblk * sizeof(int32_t));
  size_t part = gs_align256((size_t)gs_scan_partials(256 * nblk) * sizeof(int32_t));
  return hist + part;
}

// internal: ping-pong sort; returns (via *result_in_b) which buffer pair holds the result
int gs_radix_sort_pingpong(uint32_t *ka, uint32_t *va, uint32_t *kb, uint32_t *vb, int64_t n, const int32_t *n_dev,
                           int begin_bit, int end_bit, void *workspace, int *result_in_b, cudaStream_t st) {
  *result_in_b = 0;
  if (n <= 0 || end_bit <= begin_bit) return GS_OK;
  int nblk = gs_div_up(n, RS_TILE);
  int32_t *hist = (int32_t *)workspace;
  int32_t *partials = (int32_t *)((char *)workspace + gs_align256((size_t)256 * nblk * sizeof(int32_t)));
  uint32_t *ks = ka, *vs = va, *kd = kb, *vd = vb;
  for (int bit = begin_bit; bit < end_bit; bit += 8) {
    int bits = min(8, end_bit - bit);
    uint32_t mask = (1u << bits) - 1u;
    k_rs_hist<<<nblk, RS_THREADS, 0, st>>>(ks, n, n_dev, bit, mask, hist, nblk);
    int rc = gs_exclusive_scan_i32(hist, nullptr, hist, (int64_t)256 * nblk, nullptr, partials, nullptr, st);
    if (rc != GS_OK) return rc;
    k_rs_scatter<<<nblk, RS_THREADS, 0, st>>>(ks, vs, kd, vd, n, n_dev, bit, mask, hist, nblk);
    uint32_t *t;
    t = ks; ks = kd; kd = t;
    t = vs; vs = vd; vd = t;
    *result_in_b ^= 1;
  }
  GS_LAUNCH_CHECK();
  return GS_OK;
}

extern "C" int gs_radix_sort_pairs(uint32_t *keys_in, uint32_t *vals_in, uint32_t *keys_out, uint32_t *vals_out, int64_t n,
                                   const int32_t *n_dev, int begin_bit, int end_bit, void *workspace,
                                   size_t workspace_bytes, void *stream) {
  GS_REQUIRE(n >= 0 && begin_bit >= 0 && end_bit <= 32 && begin_bit <= end_bit, "bad range");
  GS_REQUIRE(n < ((int64_t)1 << 31), "n must fit int32");
  if (workspace_bytes < gs_radix_sort_workspace_bytes(n)) {
    gs_set_error("gs_radix_sort_pairs: workspace too small");
    return GS_ERR_WORKSPACE;
  }
  cudaStream_t st = (cudaStream_t)stream;
  int in_b = 0;
  int rc = gs_radix_sort_pingpong(keys_in, vals_in, keys_out, vals_out, n, n_dev, begin_bit, end_bit, workspace, &in_b, st);
  if (rc != GS_OK) return rc;
  if (!in_b && n > 0) {  // even number of passes (or none): result sits in the input pair
    GS_CUDA(cudaMemcpyAsync(keys_out, keys_in, (size_t)n * 4, cudaMemcpyDeviceToDevice, st));
    GS_CUDA(cudaMemcpyAsync(vals_out, vals_in, (size_t)n * 4, cudaMemcpyDeviceToDevice, st));
  }
  return GS_OK;
}
