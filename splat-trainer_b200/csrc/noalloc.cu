// noalloc.cu — gs_visible_indexes_static: `visible.nonzero()` (reference scene/mlp_scene.py:217) for callers that cannot
// allocate (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  It has no counterpart
// in the reference.
//
// gs_visible_indexes needs a caller-provided workspace of ~8·n bytes (flags, ranks, scan partials).  This variant is a
// single-pass chained-scan compaction whose only scratch — one work ticket and one look-back status word per 2048-row
// tile — lives in a file-scope device array of fixed size, so the caller passes inputs and outputs and nothing else:
//   * limit: n <= GS_VISIBLE_INDEXES_STATIC_MAX_ROWS (2^24 rows; the array is 33 KB and every GPU that loads the library
//     holds it until the process exits);
//   * the entry point calls no cudaMalloc / cudaFree / allocating library routine: one cudaMemsetAsync of the part of the
//     array the call uses (the array keeps the previous call's contents) and one kernel launch, both on `stream`;
//   * every call on a device shares the one array: calls that overlap on the GPU (different streams, or concurrent host
//     threads launching onto the same device) race.  Serialise them on one stream.
// In host code the array's name is a host-side stand-in, not a device address: the address the kernel is given comes from
// cudaGetSymbolAddress, looked up once per device with that device current.
#include <mutex>

#include "common.cuh"

namespace {

constexpr int NA_THREADS = 256;
constexpr int NA_ITEMS = 8;
constexpr int NA_TILE = NA_THREADS * NA_ITEMS;  // 2048 rows per CTA
constexpr int64_t NA_MAX_ROWS = GS_VISIBLE_INDEXES_STATIC_MAX_ROWS;
constexpr int NA_MAX_TILES = (int)((NA_MAX_ROWS + NA_TILE - 1) / NA_TILE);
constexpr int NA_HEAD = 64;                     // words before the status array: [0] = work ticket
constexpr int NA_MAX_DEVICES = 64;

// scratch of gs_visible_indexes_static: ticket + per-tile look-back status.  (NA_HEAD + NA_MAX_TILES) * 4 = 33 024 bytes.
static __device__ __align__(256) uint32_t g_visible_scratch[NA_HEAD + NA_MAX_TILES];

// Tiles are handed out by ticket (a tile only waits for tiles that have started); a thread owns 8 consecutive rows.
__global__ void __launch_bounds__(NA_THREADS) k_visible_indexes_static(int64_t n, const float *__restrict__ visible,
                                                                       uint32_t *__restrict__ scratch,
                                                                       int32_t *__restrict__ indexes,
                                                                       int32_t *__restrict__ count_dev) {
  __shared__ int s_warp[NA_THREADS / 32];
  __shared__ int s_tile;
  __shared__ uint32_t s_prefix;
  uint32_t *ticket = scratch, *status = scratch + NA_HEAD;
  if (threadIdx.x == 0) s_tile = (int)atomicAdd(ticket, 1u);
  __syncthreads();
  const int tile = s_tile;
  const int64_t base = (int64_t)tile * NA_TILE + (int64_t)threadIdx.x * NA_ITEMS;
  float v[NA_ITEMS];
  if (base + NA_ITEMS <= n && (reinterpret_cast<uintptr_t>(visible) & 15u) == 0) {
    const float4 *p = reinterpret_cast<const float4 *>(visible + base);
    float4 a = p[0], b = p[1];
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
  } else {
#pragma unroll
    for (int i = 0; i < NA_ITEMS; ++i) v[i] = (base + i < n) ? visible[base + i] : 0.f;
  }
  int sum = 0;
#pragma unroll
  for (int i = 0; i < NA_ITEMS; ++i) sum += v[i] > 0.f;
  // exclusive scan over the CTA's 256 threads
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  int inc = sum;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    int t = __shfl_up_sync(GS_FULL, inc, o);
    if (lane >= o) inc += t;
  }
  if (lane == 31) s_warp[w] = inc;
  __syncthreads();
  int wo = 0, tot = 0;
#pragma unroll
  for (int i = 0; i < NA_THREADS / 32; ++i) {
    int c = s_warp[i];
    if (i < w) wo += c;
    tot += c;
  }
  const int ex = wo + inc - sum;
  // chained scan over the tiles (decoupled look-back)
  if (threadIdx.x == 0) gs_st_status(status + tile, (uint32_t)tot | (tile == 0 ? GS_LB_INCL : GS_LB_LOCAL));
  if (threadIdx.x < 32) {
    uint32_t prefix = gs_lookback_warp(status, tile);
    if (threadIdx.x == 0) {
      if (tile > 0) gs_st_status(status + tile, (prefix + (uint32_t)tot) | GS_LB_INCL);
      s_prefix = prefix;
      if (tile == (int)gridDim.x - 1) *count_dev = (int32_t)(prefix + (uint32_t)tot);
    }
  }
  __syncthreads();
  uint32_t pos = s_prefix + (uint32_t)ex;
#pragma unroll
  for (int i = 0; i < NA_ITEMS; ++i)
    if (v[i] > 0.f) indexes[pos++] = (int32_t)(base + i);
}

// device address of g_visible_scratch on the current device, looked up once per device
int scratch_address(uint32_t **out) {
  static std::mutex mu;
  static uint32_t *addr[NA_MAX_DEVICES] = {nullptr};
  int dev = -1;
  GS_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= NA_MAX_DEVICES) {
    gs_set_error("gs_visible_indexes_static: device ordinal %d not supported (max %d)", dev, NA_MAX_DEVICES - 1);
    return GS_ERR_UNSUPPORTED;
  }
  std::lock_guard<std::mutex> lock(mu);
  if (!addr[dev]) {
    void *p = nullptr;
    GS_CUDA(cudaGetSymbolAddress(&p, g_visible_scratch));
    if (!p) {
      gs_set_error("gs_visible_indexes_static: cudaGetSymbolAddress returned a null address");
      return GS_ERR_CUDA;
    }
    addr[dev] = (uint32_t *)p;
  }
  *out = addr[dev];
  return GS_OK;
}

}  // namespace

extern "C" int64_t gs_visible_indexes_static_max_rows(void) { return NA_MAX_ROWS; }

extern "C" int gs_visible_indexes_static(int64_t n, const float *visible, int32_t *indexes, int32_t *count_dev, void *stream) {
  GS_REQUIRE(count_dev, "null count");
  if (n < 0 || n > NA_MAX_ROWS) {
    gs_set_error("gs_visible_indexes_static: n = %lld is outside [0, %lld] (the static scratch is sized for "
                 "GS_VISIBLE_INDEXES_STATIC_MAX_ROWS rows; use gs_visible_indexes with a workspace for more)",
                 (long long)n, (long long)NA_MAX_ROWS);
    return GS_ERR_INVALID;
  }
  cudaStream_t st = (cudaStream_t)stream;
  if (n == 0) { GS_CUDA(cudaMemsetAsync(count_dev, 0, 4, st)); return GS_OK; }
  GS_REQUIRE(visible && indexes, "null argument");
  uint32_t *scratch = nullptr;
  int rc = scratch_address(&scratch);
  if (rc != GS_OK) return rc;
  const int tiles = gs_div_up(n, NA_TILE);
  // the array holds the previous call's ticket and status words: clear the part this call uses
  GS_CUDA(cudaMemsetAsync(scratch, 0, (size_t)(NA_HEAD + tiles) * sizeof(uint32_t), st));
  k_visible_indexes_static<<<tiles, NA_THREADS, 0, st>>>(n, visible, scratch, indexes, count_dev);
  GS_LAUNCH_CHECK();
  return GS_OK;
}
